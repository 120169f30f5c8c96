#!/bin/bash
# A/B of the kernel-variant switches on one GPU: runs bench.py (cfg2, no CPU leg) once per setting and keeps the JSON
# lines under gpurun_out/<tag>_<name>.json. Usage: tools/ab_options.sh <tag> <steps> name:VAR=val,VAR=val ...
tag=$1; steps=$2; shift 2
mkdir -p gpurun_out
for spec in "$@"; do
    name=${spec%%:*}; envs=${spec#*:}
    env $(echo "$envs" | tr ',' ' ') python bench.py --steps "$steps" --warmup 5 --no-cpu-baseline \
        > "gpurun_out/${tag}_${name}.json" 2> "gpurun_out/${tag}_${name}.err" || echo "FAILED ${name} (see gpurun_out/${tag}_${name}.err)"
done
python - "$tag" <<'PY'
import glob, json, sys
tag = sys.argv[1]
keys = ("gsb_sh_colors_fwd", "gsb_sh_colors_bwd", "gsb_projection_bwd", "gsb_isect_finalize_cap", "gsb_radix_sort_pairs_u32",
        "gsb_radix_sort_pairs_u32_prepared")
for f in sorted(glob.glob(f"gpurun_out/{tag}_*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
    except Exception as e:
        print(f, "no JSON line:", e); continue
    st = d.get("stages", {})
    print(f.split("/")[-1], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"],
          {k.replace("gsb_", ""): st[k]["ms_per_step"] for k in keys if k in st})
PY
