"""Turns the two ncu artefacts of a profiling call into the files kept under profiles/.

    python tools/ncu_extract.py summary  gpurun_out/r01u_full.ncu-rep  profiles/r01u_ncu_full_summary.csv
    python tools/ncu_extract.py step     gpurun_out/r01u_launches.csv            # share-of-step table (markdown)
    python tools/ncu_extract.py lines    gpurun_out/r01u_full.ncu-rep rasterize_bwd [--top 40]

`summary`: one column per captured launch, one row per metric (`ncu -i REP --page raw --csv`).
`step`:    aggregates the launch list (`--metrics gpu__time_duration.sum`) of the last complete step, delimited by
           two consecutive `projection_fwd` launches, into time and share per kernel.
`lines`:   executed warp instructions and stall samples per source line of one kernel (`--page source --csv`;
           needs `-lineinfo`, which build.py passes, and `--import-source on` at capture time).
"""
import csv
import io
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed_op_global_red.sum",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
]


def _ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def summary(rep, dst):
    rows = _ncu_csv(rep, "raw")
    head, units, launches = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(head)}
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + [r[col["Kernel Name"]][:48] for r in launches])
        for m in ["Grid Size", "Block Size"] + METRICS:
            if m in col:
                w.writerow([m, units[col[m]]] + [r[col[m]] for r in launches])
    print(f"{dst}: {len(launches)} launches")


def step(path):
    rows = list(csv.reader(open(path)))
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    head = rows[start]
    name, value = head.index("Kernel Name"), head.index("Metric Value")
    launches = [(r[name], float(r[value].replace(",", ""))) for r in rows[start + 1:] if len(r) > value and r[value]]
    marks = [i for i, (k, _) in enumerate(launches) if "projection_fwd" in k]
    one = launches[marks[-2]:marks[-1]]
    total = sum(v for _, v in one)
    print(f"{len(one)} launches, {total / 1000:.1f} us\n\n| kernel | launches | us | share |\n|---|---|---|---|")
    agg = {}
    for k, v in one:
        key = k.split("(")[0].replace("void ", "")[:70]
        t = agg.setdefault(key, [0.0, 0])
        t[0] += v
        t[1] += 1
    for k, (v, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"| `{k}` | {n} | {v / 1000:.1f} | {100 * v / total:.1f} % |")


def lines(rep, kernel, top):
    rows = _ncu_csv(rep, "source", ["--kernel-name", f"regex:{kernel}", "--print-source", "cuda,sass"])
    current, head, agg, text = None, None, {}, {}
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            current, head = r[1].split("/")[-1], None
        elif r and r[0] == "Line No":
            head = r
        elif head and current and len(r) > 8 and r[0].isdigit():
            try:
                inst = float(r[head.index("Instructions Executed")] or 0)
                smp = float(r[head.index("# Samples")] or 0)
            except ValueError:
                continue
            key = (current, int(r[0]))
            t = agg.setdefault(key, [0.0, 0.0])
            t[0] += inst
            t[1] += smp
            text[key] = r[1].strip()
    ti, ts = sum(a[0] for a in agg.values()) or 1.0, sum(a[1] for a in agg.values()) or 1.0
    print(f"{ti / 1e6:.1f} M warp instructions, {ts:.0f} samples")
    for key, (inst, smp) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{key[0]:24s}{key[1]:5d} {100 * inst / ti:5.1f}% inst {100 * smp / ts:5.1f}% smp | {text[key][:100]}")


if __name__ == "__main__":
    cmd = sys.argv[1] if len(sys.argv) > 1 else ""
    if cmd == "summary" and len(sys.argv) == 4:
        summary(sys.argv[2], sys.argv[3])
    elif cmd == "step" and len(sys.argv) == 3:
        step(sys.argv[2])
    elif cmd == "lines" and len(sys.argv) >= 4:
        lines(sys.argv[2], sys.argv[3], int(sys.argv[5]) if len(sys.argv) > 5 and sys.argv[4] == "--top" else 40)
    else:
        sys.exit(__doc__)
