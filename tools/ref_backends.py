"""Same-box comparison of the b200 backend with the reference's OWN GPU backends.

    python tools/ref_backends.py [--configs cfg1,cfg2,cfg3] [--out profiles/r02_ref_backends.json]

Needs the unmodified reference under baseline/_ref (tools/install_reference.sh; git-ignored, travels with gpurun).
Every (backend, config) runs in its own process so the `GSPLAT_BACKEND` import-time switch of
`gsplat/rendering.py:11-29` is honoured:
  * `b200`   : this repo's package (gsplat-triton_b200/);
  * `triton` : baseline/_ref, GSPLAT_BACKEND=triton (unpacked, pinhole only; `triton_impl/_wrapper.py:523`);
  * `cuda`   : baseline/_ref, GSPLAT_BACKEND=cuda, the reference CUDA extension built for sm_100 by
               tools/build_reference_cuda.py.
Per run: (i) perf with the reference's protocol (profiling/main.py:28-38: warm-ups, then CUDA-event-timed
forward and forward+backward iterations); (ii) outputs dumped for the parity table computed by the parent:
radii exact-match rate and max |delta| (triton_tests/test_fused_proj.py:116), image / alpha / gradient errors, and
— feeding the REFERENCE's projection outputs into b200's isect_tiles (protocol of triton_tests/test_isect.py:66-89) —
`torch.equal` on isect_ids / flatten_ids / isect_offsets.
A backend that cannot be built or run on the box is recorded with its exact error instead.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")


def make_scene(name):
    import numpy as np
    import torch

    if name == "cfg1":  # reference gsplat/_helper.py:9-55 on the committed fixture, camera 0
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from conftest import garden_scene

        means, quats, scales, opac, colors, viewmats, Ks, W, H = garden_scene()
        return dict(means=means, quats=quats, scales=scales, opacities=opac, colors=colors, viewmats=viewmats[:1],
                    Ks=Ks[:1], W=W, H=H, kw=dict(packed=False))
    if name == "cfg2":
        sys.path.insert(0, ROOT)
        from bench import make_cfg2

        s = make_cfg2("cpu")
        return dict(means=s["means"], quats=s["quats"], scales=s["scales"], opacities=s["opacities"], colors=s["shs"],
                    viewmats=s["viewmats"], Ks=s["Ks"], W=s["width"], H=s["height"], kw=dict(packed=False, sh_degree=3))
    if name in ("cfg3", "cfg3u"):  # 3M Gaussians, 8 ring cameras at 1297x840, packed + antialiased
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from scenes import ring_scene

        N = 3_000_000 if name == "cfg3" else 1_000_000
        means, quats, scales, opac, shs, viewmats, Ks = ring_scene(N, 8, 1297, 840)
        kw = dict(packed=(name == "cfg3"), sh_degree=3, rasterize_mode="antialiased")
        return dict(means=means, quats=quats, scales=scales, opacities=opac, colors=shs, viewmats=viewmats, Ks=Ks,
                    W=1297, H=840, kw=kw)
    raise ValueError(name)


def worker(args):
    backend = args.worker
    if backend == "b200":
        sys.path.insert(0, os.path.join(ROOT, "gsplat-triton_b200"))
    else:
        sys.path.insert(0, REF)
    os.environ["GSPLAT_BACKEND"] = backend
    import torch

    from gsplat.rendering import rasterization

    dev = torch.device("cuda", 0)
    s = make_scene(args.config)
    names = ("means", "quats", "scales", "opacities", "colors")
    params = [s[k].to(dev).requires_grad_(True) for k in names]
    viewmats, Ks, W, H, kw = s["viewmats"].to(dev), s["Ks"].to(dev), s["W"], s["H"], s["kw"]
    g = torch.Generator().manual_seed(1)
    C = viewmats.shape[0]
    Gw = torch.rand(C, H, W, 3, generator=g).to(dev)
    Aw = torch.rand(C, H, W, 1, generator=g).to(dev)

    def step(backward=True):
        for p in params:
            p.grad = None
        rc, ra, meta = rasterization(*params, viewmats, Ks, W, H, **kw)
        if backward:
            ((rc * Gw).sum() + (ra * Aw).sum()).backward()
        return rc, ra, meta

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    t0 = time.time()
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    warm_s = time.time() - t0

    def timed(backward):
        ms = []
        for _ in range(args.steps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            step(backward)
            b.record()
            torch.cuda.synchronize()
            ms.append(a.elapsed_time(b))
        ms.sort()
        return ms[len(ms) // 2], ms[0]

    full_med, full_min = timed(True)
    with torch.no_grad():
        fwd_med, fwd_min = timed(False)
    torch.cuda.reset_peak_memory_stats()
    rc, ra, meta = step()
    torch.cuda.synchronize()
    out = {"backend": backend, "config": args.config, "ms_fwd_bwd_median": full_med, "ms_fwd_bwd_min": full_min,
           "ms_fwd_median": fwd_med, "ms_fwd_min": fwd_min, "warmup_s": warm_s, "steps": args.steps,
           "n_isects": int(meta["flatten_ids"].numel()), "peak_mem_gb": torch.cuda.max_memory_allocated() / 2**30,
           "torch": torch.__version__}
    dump = {"render_colors": rc.detach().cpu(), "render_alphas": ra.detach().cpu()}
    for k in ("radii", "means2d", "depths", "conics", "opacities", "isect_ids", "flatten_ids", "isect_offsets",
              "camera_ids", "gaussian_ids"):
        v = meta.get(k)
        if v is not None and args.config in ("cfg1", "cfg2"):
            dump[k] = v.detach().cpu()
    for n, p in zip(names, params):
        dump["v_" + n] = p.grad.detach().cpu() if args.config in ("cfg1", "cfg2") else p.grad.detach()[::64].cpu()
    torch.save(dump, args.dump)
    print("WORKER_JSON " + json.dumps(out), flush=True)


def isect_from(args):
    """b200 isect_tiles / isect_offset_encode on the projection outputs a reference backend dumped."""
    sys.path.insert(0, os.path.join(ROOT, "gsplat-triton_b200"))
    os.environ["GSPLAT_BACKEND"] = "b200"
    if args.tile_bits:
        os.environ["GSPLAT_B200_TILE_BITS"] = args.tile_bits
    import torch

    from gsplat.b200_impl._wrapper import isect_offset_encode, isect_tiles

    d = torch.load(args.isect_from)
    dev = torch.device("cuda", 0)
    means2d, radii, depths = d["means2d"].to(dev), d["radii"].to(dev), d["depths"].to(dev)
    C, th, tw = d["isect_offsets"].shape
    _, ids, flat = isect_tiles(means2d, radii, depths, 16, tw, th)
    offs = isect_offset_encode(ids, C, tw, th)
    res = {"isect_ids_equal": bool(torch.equal(ids.cpu(), d["isect_ids"])),
           "flatten_ids_equal": bool(torch.equal(flat.cpu(), d["flatten_ids"].to(flat.dtype))),
           "isect_offsets_equal": bool(torch.equal(offs.cpu(), d["isect_offsets"].to(offs.dtype))),
           "n_isects": int(ids.numel()), "n_isects_ref": int(d["isect_ids"].numel())}
    if not res["flatten_ids_equal"] and res["isect_ids_equal"]:
        # equal keys (same tile, same depth bits) may be ordered differently by an unstable tie-break: count them
        diff = flat.cpu() != d["flatten_ids"].to(flat.dtype)
        ties = d["isect_ids"][diff]
        res["flatten_ids_mismatches"] = int(diff.sum())
        res["mismatches_inside_equal_key_runs"] = bool(
            torch.equal(torch.sort(flat.cpu()[diff]).values, torch.sort(d["flatten_ids"][diff].to(flat.dtype)).values))
        res["distinct_tied_keys"] = int(torch.unique(ties).numel())
    print("WORKER_JSON " + json.dumps(res), flush=True)


def run_child(extra, timeout):
    env = dict(os.environ)
    env.setdefault("TORCH_EXTENSIONS_DIR", os.path.join(REF, "torch_ext"))
    env.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    cmd = [sys.executable, os.path.abspath(__file__)] + extra
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
    except subprocess.TimeoutExpired:
        return {"error": f"timeout after {timeout} s"}
    for line in r.stdout.splitlines():
        if line.startswith("WORKER_JSON "):
            return json.loads(line[len("WORKER_JSON "):])
    return {"error": (r.stderr or r.stdout)[-1500:], "rc": r.returncode}


def compare(ours, theirs, config):
    import torch

    a, b = torch.load(ours), torch.load(theirs)
    out = {}
    for k in ("render_colors", "render_alphas"):
        err = (a[k] - b[k]).abs()
        out[k] = {"max_abs_err": float(err.max()),
                  "frac_beyond_rtol1e-4_atol1e-4": float((err > 1e-4 + 1e-4 * b[k].abs()).float().mean())}
    if "radii" in a and "radii" in b and a["radii"].shape == b["radii"].shape:
        ra_, rb_ = a["radii"].long(), b["radii"].long()
        out["radii"] = {"exact_match_rate": float((ra_ == rb_).float().mean()), "max_abs_delta": int((ra_ - rb_).abs().max()),
                        "visible_ours": int((ra_ > 0).sum()), "visible_ref": int((rb_ > 0).sum())}
        both = (ra_ > 0) & (rb_ > 0)
        for k in ("means2d", "depths", "conics"):
            if k in a and k in b:
                d = (a[k] - b[k]).abs()[both]
                out[k] = {"max_abs_err": float(d.max()), "max_rel_err": float((d / (b[k][both].abs() + 1e-6)).max())}
    for k in ("v_means", "v_quats", "v_scales", "v_opacities", "v_colors"):
        ga, gb = a[k].float(), b[k].float()
        err = (ga - gb).abs()
        scale = float(gb.abs().max()) + 1e-20
        out[k] = {"max_err_over_max_grad": float(err.max()) / scale,
                  "frac_beyond_rtol1e-3": float((err > 1e-3 * gb.abs() + 1e-5 * scale).float().mean())}
    if "isect_ids" in a and "isect_ids" in b:
        out["n_isects"] = {"ours": int(a["isect_ids"].numel()), "ref": int(b["isect_ids"].numel())}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="cfg1,cfg2,cfg3")
    ap.add_argument("--backends", default="b200,triton,cuda")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r02_ref_backends.json"))
    ap.add_argument("--worker")
    ap.add_argument("--config")
    ap.add_argument("--dump")
    ap.add_argument("--isect-from")
    ap.add_argument("--tile-bits", default="")
    ap.add_argument("--timeout", type=int, default=900)
    args = ap.parse_args()
    if args.isect_from:
        return isect_from(args)
    if args.worker:
        return worker(args)
    if not os.path.isdir(os.path.join(REF, "gsplat")):
        raise SystemExit("baseline/_ref is missing: run tools/install_reference.sh in the build container")
    tmp = tempfile.mkdtemp(prefix="gsb_ref_")
    report = {"protocol": f"{args.warmup} warm-ups, {args.steps} timed fwd+bwd iterations and {args.steps} forward-only "
                          "iterations, CUDA events, median; 256 MB L2 flush between iterations", "runs": {}, "parity": {}}
    for config in args.configs.split(","):
        dumps = {}
        for backend in args.backends.split(","):
            if backend == "triton" and config == "cfg3":
                report["runs"][f"{config}/{backend}"] = {"skipped": "Triton backend has no packed mode "
                                                                    "(triton_impl/_wrapper.py:523) and miscomputes camera ids beyond 2^24 rows"}
                continue
            dump = os.path.join(tmp, f"{config}_{backend}.pt")
            res = run_child(["--worker", backend, "--config", config, "--dump", dump, "--steps", str(args.steps),
                             "--warmup", str(args.warmup)], args.timeout)
            report["runs"][f"{config}/{backend}"] = res
            print(config, backend, json.dumps(res)[:400], flush=True)
            if "error" not in res:
                dumps[backend] = dump
        for backend in ("triton", "cuda"):
            if backend in dumps and "b200" in dumps:
                try:
                    par = compare(dumps["b200"], dumps[backend], config)
                    if config in ("cfg1", "cfg2"):
                        # Triton sizes the tile field as (n_tiles-1).bit_length(), CUDA as n_tiles.bit_length() (SURVEY 8a hazard 3)
                        par["b200_isect_on_reference_projection"] = run_child(
                            ["--isect-from", dumps[backend], "--tile-bits", "cuda" if backend == "cuda" else "triton"],
                            args.timeout)
                    report["parity"][f"{config}/b200_vs_{backend}"] = par
                    print(config, "parity vs", backend, json.dumps(par)[:600], flush=True)
                except Exception as e:  # noqa: BLE001
                    report["parity"][f"{config}/b200_vs_{backend}"] = {"error": repr(e)}
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(report, open(args.out, "w"), indent=1)
    print("written", args.out)


if __name__ == "__main__":
    main()
