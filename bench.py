"""bench.py — fwd+bwd render iterations/s of `gsplat.rendering.rasterization()` under GSPLAT_BACKEND=b200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N=1 workload = BASELINE.json configs[1] ("cfg2"): synthetic 1M Gaussians, SH degree 3, one
1920x1080 pinhole camera, packed=False, loss = sum(colors*G) + sum(alphas*g).  N>1 (torchrun, one rank
per GPU, NCCL): the same 1M-Gaussian scene strided over the ranks (`means[rank::N]`), one camera per
rank, `distributed=True` (per-camera all-to-all of the projected rows) — per-GPU work is constant
(weak scaling); `value` = camera renders (fwd+bwd) per second over all ranks.

One JSON line on stdout (rank 0): see the keys in `main`. `--impl reference` times the CPU oracle
port of the reference's torch path (oracle/gs_oracle.py) on the host cores instead.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "gsplat-triton_b200"))
os.environ.setdefault("GSPLAT_BACKEND", "b200")

WIDTH, HEIGHT, N_GAUSS, SH_DEGREE = 1920, 1080, 1_000_000, 3
SM_COUNT, FP32_LANES = 148, 128
# FP32 issue slots per composited pixel-Gaussian pair (14 fwd incl. the D=3 blend, 40 bwd); DESIGN.md 4
SLOTS_FWD, SLOTS_BWD = 14, 40


def make_cfg2(device="cpu", n=N_GAUSS, seed=0):
    """Gaussians uniformly filling the frustum of an identity camera, fov 60 deg."""
    g = torch.Generator().manual_seed(seed)
    fx = 0.5 * WIDTH / math.tan(math.radians(30.0))
    u = torch.rand(n, generator=g) * 2.1 - 1.05
    v = torch.rand(n, generator=g) * 2.1 - 1.05
    z = torch.rand(n, generator=g) * 8.0 + 2.0
    means = torch.stack([u * z * WIDTH / (2 * fx), v * z * HEIGHT / (2 * fx), z], dim=-1)
    quats = torch.randn(n, 4, generator=g)
    lo, hi = math.log(0.002), math.log(0.02)
    scales = torch.exp(torch.rand(n, 3, generator=g) * (hi - lo) + lo)
    opacities = torch.rand(n, generator=g)
    shs = torch.randn(n, (SH_DEGREE + 1) ** 2, 3, generator=g) * 0.3
    shs[:, 0, :] += 0.5
    viewmats = torch.eye(4)[None]
    Ks = torch.tensor([[[fx, 0.0, WIDTH / 2], [0.0, fx, HEIGHT / 2], [0.0, 0.0, 1.0]]])
    s = dict(means=means, quats=quats, scales=scales, opacities=opacities, shs=shs, viewmats=viewmats, Ks=Ks)
    s = {k: t.to(device) for k, t in s.items()}
    s.update(width=WIDTH, height=HEIGHT)
    return s


def loss_weights(device, seed=1):
    g = torch.Generator().manual_seed(seed)
    G = torch.rand(1, HEIGHT, WIDTH, 3, generator=g)
    a = torch.rand(1, HEIGHT, WIDTH, 1, generator=g)
    return G.to(device), a.to(device)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, index, enabled=True):
        self.rows, self.proc, self.index, self.enabled = [], None, index, enabled

    def __enter__(self):
        if not self.enabled:
            return self
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc is not None:
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm = sorted(float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, c in zip(self.NAMES, r[2:6]) if c == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": float(d["hbm_gbs"]), "sm_max_mhz": float(d.get("sm_max_mhz", 1965.0)), "source": "measured"}
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "source": "fallback"}


# ------------------------------------------------------------------------------------------------
# the b200 arm
# ------------------------------------------------------------------------------------------------
def _raster_work_counts(meta, device):
    """(E, Hits): E = reference-algorithm evaluations 256 * sum_tiles(last contributing index - tile start + 1)
    (SURVEY 8d); Hits = pixel-Gaussian pairs actually composited (alpha >= 1/255, before termination)."""
    from gsplat.b200_impl._lib import lib, ptr, stream_of

    means2d, conics, opac = (meta[k].detach().contiguous() for k in ("means2d", "conics", "opacities"))
    offs, flat = meta["isect_offsets"], meta["flatten_ids"]
    C, th, tw = offs.shape
    P, W, H = opac.numel(), meta["width"], meta["height"]
    rec = torch.empty((2, P, 4), device=device)
    ones = torch.ones((P,), device=device)
    cnt, ra = torch.empty((C, H, W), device=device), torch.empty((C, H, W), device=device)
    last = torch.empty((C, H, W), dtype=torch.int32, device=device)
    st = stream_of(rec)
    lib.gsb_rasterize_pack(P, ptr(means2d), ptr(conics), ptr(opac), ptr(rec[0]), ptr(rec[1]), st)
    lib.gsb_rasterize_count(C, flat.numel(), ptr(rec[0]), ptr(rec[1]), ptr(ones), W, H, 16, tw, th, ptr(offs),
                            ptr(flat), ptr(cnt), ptr(ra), ptr(last), st)
    hits = int(cnt.double().sum().item())
    last = torch.where(cnt > 0, last, torch.full_like(last, -1))
    pad_h, pad_w = th * 16 - H, tw * 16 - W
    last = torch.nn.functional.pad(last, (0, pad_w, 0, pad_h), value=-1)
    tile_last = last.reshape(C, th, 16, tw, 16).amax(dim=(2, 4))
    work = (tile_last - offs + 1).clamp(min=0)
    work = torch.where(tile_last >= 0, work, torch.zeros_like(work))
    return 256 * int(work.sum().item()), hits


def _stage_table(summary, steps, meta, n_local, C, D, K, work, peaks, sm_mhz):
    """Per-entry-point ms/step, algorithmic bytes (SURVEY 8d formulas) or FP32 slots, achieved vs peak."""
    P = C * n_local
    I = meta["flatten_ids"].numel()
    px = C * meta["width"] * meta["height"]
    n_tiles = meta["tile_width"] * meta["tile_height"]
    bits = 32 + (n_tiles - 1).bit_length() + max(C - 1, 0).bit_length()
    hbm = {
        "gsb_projection_fwd": 40 * n_local + 28 * P,
        "gsb_projection_bwd": 80 * n_local + 40 * P,
        "gsb_sh_fwd": (12 * K + 12) * n_local + 12 * P + P,
        "gsb_sh_colors_fwd": (12 * K + 12) * n_local + 12 * P + 4 * P,
        "gsb_sh_colors_bwd": (12 * K + 12) * n_local + 12 * P + 4 * P + (12 * K + 12) * n_local,
        "gsb_sh_bwd": 24 * P + P + 12 * K * n_local,
        "gsb_isect_count": 12 * P + 4 * P,
        "gsb_isect_row_keys": 4 * P + 12 * P,
        "gsb_isect_row_tiles": 12 * P + 24 * P,
        "gsb_isect_count_perm": 4 * P + 4 * P,
        "gsb_isect_emit_perm": 20 * P + 8 * I,
        "gsb_isect_emit_perm_cap": 20 * P + 8 * I,  # perm 4 + one 16-byte record per row, 8 per intersection out
        "gsb_radix_sort_pairs_u32": 4 * 16 * P + 4 * P + math.ceil((bits - 32) / 8) * 16 * I + 4 * I,
        "gsb_isect_finalize_cap": 8 * I + 4 * I + 8 * I + 4 * C * n_tiles,
        "gsb_a2a_put_fwd": 2 * (8 + D) * 4 * P,
        "gsb_a2a_put_bwd": 2 * (7 + D) * 4 * P,
        "gsb_a2a_wait_copy": (2 * (8 + D) + 2 * (7 + D)) * 4 * P,
        "gsb_sort_pairs_u32": math.ceil((bits - 32) / 8) * 16 * I + 4 * I,
        "gsb_isect_finalize": 8 * I + 4 * I + 8 * I + 4 * C * n_tiles,
        "gsb_isect_emit": 20 * P + 12 * I,
        "gsb_sort_pairs": 4 * 24 * P + 8 * P,  # row sort of the presorted path: 32 depth bits over P rows
        "gsb_isect_offsets": 8 * I + 4 * C * n_tiles,
        "gsb_rasterize_pack": 24 * P + 32 * P,
    }
    fp32_peak = SM_COUNT * FP32_LANES * 2 * peaks["sm_max_mhz"] * 1e6 / 1e12  # TFLOP/s at max clock, FMA = 2
    evals, hits = work if work is not None else (None, None)
    flops = {} if hits is None else {"gsb_rasterize_fwd": 2 * SLOTS_FWD * hits, "gsb_rasterize_fwd_cap": 2 * SLOTS_FWD * hits,
                                     "gsb_rasterize_bwd": 2 * SLOTS_BWD * hits}
    out = {}
    for name, (calls, ms) in summary.items():
        per_step = ms / steps
        row = {"ms_per_step": round(per_step, 4), "launches_per_step": calls / steps}
        if name in hbm and per_step > 0:
            a = hbm[name] / (per_step * 1e-3) / 1e9
            row.update(bound="hbm", achieved=round(a, 1), peak=peaks["hbm_gbs"], unit="GB/s",
                       frac=round(a / peaks["hbm_gbs"], 4), algorithmic_bytes=hbm[name])
        elif name in flops and per_step > 0:
            a = flops[name] / (per_step * 1e-3) / 1e12
            # headline = SURVEY 8(d): slots x the reference algorithm's evaluations E (every pixel of a tile against
            # every Gaussian up to the tile's last contributor). The kernels cull most of E exactly, so this can
            # exceed 1; `frac_composited_pairs` counts only the pairs that really blend (strict lower bound).
            a_ref = a * evals / max(hits, 1)
            row.update(bound="fp32", achieved=round(a_ref, 2), peak=round(fp32_peak, 2), unit="TFLOP/s",
                       frac=round(a_ref / fp32_peak, 4), reference_evaluations=evals, composited_pairs=hits,
                       achieved_composited_pairs=round(a, 2), frac_composited_pairs=round(a / fp32_peak, 4))
        out[name] = row
    return out


def run_b200(args):
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torch.distributed.run"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from gsplat.b200_impl._lib import lib, timing
    from gsplat.rendering import rasterization

    host = make_cfg2("cpu")
    names = ("means", "quats", "scales", "opacities", "shs")
    host_params = [host[k][rank::world].contiguous().pin_memory() for k in names]
    viewmats, Ks = host["viewmats"].to(dev), host["Ks"].to(dev)
    G, gA = loss_weights(dev)
    n_local = host_params[0].shape[0]

    from gsplat.distributed import host_trace

    def iteration(params):
        host_trace.mark()
        rc, ra, meta = rasterization(*params, viewmats, Ks, WIDTH, HEIGHT, sh_degree=SH_DEGREE, packed=False,
                                     distributed=(world > 1))
        host_trace.mark()
        loss = (rc * G).sum() + (ra * gA).sum()
        loss.backward()
        host_trace.mark("us_loss_backward")
        return rc, ra, loss, meta

    params = [t.to(dev).requires_grad_(True) for t in host_params]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_loop(fn, steps):
        evs = []
        for _ in range(steps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs)

    def resident_step():
        for p in params:
            p.grad = None
        iteration(params)

    for _ in range(max(args.warmup, 3)):
        resident_step()
    barrier()
    lib.gsb_reset_launch_count()
    timing.reset(False)  # the timed steps carry no per-launch instrumentation; the stage table is measured after
    # one sampler per NODE (local rank 0, its own GPU): N nvidia-smi pollers contend for the driver's locks
    clocks = ClockSampler(local_rank, enabled=(local_rank == 0))
    clocks.__enter__()  # sampled through both timed regions (resident steps, then end-to-end steps)
    host_trace.acc.clear()
    t0 = time.perf_counter()
    total_ms = timed_loop(resident_step, args.steps)
    host_resident = host_trace.summary()
    barrier()
    wall_s = time.perf_counter() - t0
    launches = int(lib.gsb_launch_count())
    # per-entry-point CUDA-event timing (the `stages` table / roofline): a second, instrumented pass
    stage_steps = min(args.steps, 20)
    timing.reset(True)
    timed_loop(resident_step, stage_steps)
    barrier()
    stage_summary = timing.summary()
    timing.reset(False)

    # end to end: host buffers in, image out, copies inside the timed region
    out_c = torch.empty((1, HEIGHT, WIDTH, 3), dtype=torch.float32).pin_memory()
    out_a = torch.empty((1, HEIGHT, WIDTH, 1), dtype=torch.float32).pin_memory()
    h2d = sum(t.numel() * t.element_size() for t in host_params)
    d2h = out_c.numel() * 4 + out_a.numel() * 4 + 4

    # double-buffered: step i+1's host->device copy (copy stream) overlaps step i's kernels; every step
    # still copies all of its inputs from pinned host memory and reads its image + loss back to the host
    main_stream = torch.cuda.current_stream()
    copy_stream = torch.cuda.Stream()
    d2h_stream = torch.cuda.Stream()
    bufs = [[torch.empty(t.shape, dtype=t.dtype, device=dev) for t in host_params] for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]

    def prefetch(i):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[i % 2])
            for b, h in zip(bufs[i % 2], host_params):
                b.copy_(h, non_blocking=True)
            ready[i % 2].record(copy_stream)

    loss_host = [torch.empty(1, dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ready = [torch.cuda.Event() for _ in range(2)]

    def e2e_loop(steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        copy_stream.wait_event(a)
        for e in consumed:
            e.record()
        prefetch(0)
        seen = 0.0
        for i in range(steps):
            if i + 1 < steps:
                prefetch(i + 1)
            main_stream.wait_event(ready[i % 2])
            ps = [t.detach().requires_grad_(True) for t in bufs[i % 2]]
            rc, ra, loss, _ = iteration(ps)
            consumed[i % 2].record()
            # read-back on its own stream: the next step's kernels do not queue behind 33 MB of D2H
            d2h_stream.wait_event(consumed[i % 2])
            with torch.cuda.stream(d2h_stream):
                rc_d, ra_d, loss_d = rc.detach(), ra.detach(), loss.detach().reshape(1)
                for t in (rc_d, ra_d, loss_d):
                    t.record_stream(d2h_stream)
                out_c.copy_(rc_d, non_blocking=True)
                out_a.copy_(ra_d, non_blocking=True)
                loss_host[i % 2].copy_(loss_d, non_blocking=True)
                loss_ready[i % 2].record(d2h_stream)
            if i >= 1:  # the host reads every step's loss, one step behind the launches (no pipeline drain)
                loss_ready[(i - 1) % 2].synchronize()
                seen += float(loss_host[(i - 1) % 2])
        loss_ready[(steps - 1) % 2].synchronize()
        seen += float(loss_host[(steps - 1) % 2])
        main_stream.wait_stream(d2h_stream)  # the closing event also covers the last read-back
        b.record()
        torch.cuda.synchronize()
        assert math.isfinite(seen)
        return a.elapsed_time(b)

    e2e_loop(3)
    barrier()
    e2e_ms = e2e_loop(args.steps)
    barrier()
    clocks.__exit__(None, None, None)

    ms = torch.tensor([total_ms / args.steps, e2e_ms / args.steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step, ms_e2e = ms.tolist()

    # one more (untimed) iteration on EVERY rank (it contains collectives) to read the stage geometry from `meta`
    for p in params:
        p.grad = None
    _, _, _, meta = iteration(params)
    barrier()
    if rank == 0:
        # exchange rows are re-laid out after `meta` is filled when distributed: count on a single GPU only
        work = _raster_work_counts(meta, dev) if world == 1 else None
        peaks = measured_peaks()
        clk = clocks.summary()
        C_local, D, K = 1, 3, (SH_DEGREE + 1) ** 2
        n_rows = meta["means2d"].shape[1]
        stages = _stage_table(stage_summary, stage_steps, meta, n_rows, C_local, D, K, work, peaks, clk["sm_mhz"])
        top = max((k for k in stages if "bound" in stages[k]), key=lambda k: stages[k]["ms_per_step"])
        roof = {k: stages[top][k] for k in ("bound", "achieved", "peak", "unit", "frac")}
        traffic_file = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        traffic = json.load(open(traffic_file)).get(top) if os.path.exists(traffic_file) else None
        roof.update(kernel=top, traffic=traffic, peak_source=(
            "148 SMs*128 lanes*2*sm_max_mhz (no measured FP32 peak in MEASURED_PEAKS.json)"
            if roof["bound"] == "fp32" else peaks["source"] + " MEASURED_PEAKS.json hbm_gbs"),
            share_of_step=round(stages[top]["ms_per_step"] / ms_step, 4))
        line = {
            "metric": "fwd+bwd render iters/s (1M Gaussians, 1080p)", "value": round(world * 1e3 / ms_step, 3),
            "unit": "iters/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "cfg2: 1M Gaussians SH deg 3, 1920x1080, 1 camera per GPU, packed=False"
                       + (f", Gaussians strided over {world} ranks, distributed=True" if world > 1 else ""),
                       "n_gaussians_per_gpu": n_local, "n_isects": meta["flatten_ids"].numel(),
                       "l2": "inputs 236 MB/step > 126 MB L2 and a 256 MB flush write between steps (untimed)",
                       "timing": "sum of per-step CUDA-event times on the launch stream, max over ranks"},
            "e2e": {"value": round(world * 1e3 / ms_e2e, 3), "unit": "iters/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": round(ms_e2e, 4),
                    "pipeline": "double-buffered H2D on a copy stream (step i+1's copy overlaps step i's kernels); "
                                "image, alpha and loss of every step are copied to pinned host memory on a read-back stream and the host "
                                "reads each loss one step behind the launches; one CUDA-event bracket around all K "
                                "steps; inputs (236 MB/step by DMA) exceed L2"},
            "gpu_launches": launches, "wall_s_timed_region": round(wall_s, 3), "clocks": clk,
            "roofline": roof, "stages": stages,
        }
        if host_trace.enabled:
            line["host_trace_us"] = host_resident
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(rows=2, gauss_stride=8)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# the CPU arm: oracle port of the reference's pure-torch path (_torch_impl) on the host cores
# ------------------------------------------------------------------------------------------------
_WARM = []


def _cpu_reference_step(scene, O, rows, gauss_stride):
    """One bounded sample of a cfg2 fwd+bwd iteration on the CPU; returns the estimated full-iteration seconds.

    Sample: (A) covariance + projection + SH colour forward and backward on every `gauss_stride`-th
    Gaussian (linear in N, scaled by gauss_stride); (B) tile intersection + sort + offsets on the full
    scene (unscaled); (C) compositor forward+backward on `rows` of the 68 tile rows (scaled by 68/rows).
    """
    W, H = scene["width"], scene["height"]
    names = ("means", "quats", "scales", "opacities", "shs")
    t = {}
    def stage_a():
        sub = [scene[k][::gauss_stride].clone().requires_grad_(True) for k in names]
        t0 = time.perf_counter()
        cov = O.quat_scale_to_covar(sub[1], sub[2])
        radii, means2d, depths, conics, _ = O.fully_fused_projection(sub[0], cov, scene["viewmats"], scene["Ks"], W, H)
        dirs = sub[0][None] - torch.inverse(scene["viewmats"])[:, None, :3, 3]
        col = torch.clamp_min(O.spherical_harmonics(SH_DEGREE, dirs, sub[4][None], masks=radii > 0) + 0.5, 0.0)
        opac = sub[3][None]
        t1 = time.perf_counter()
        outs = [means2d, conics, col, opac]
        torch.autograd.backward(outs, [torch.ones_like(x) for x in outs])
        return t1 - t0, time.perf_counter() - t1

    if not _WARM:
        stage_a()  # first call pays thread-pool / einsum-plan start-up: not part of the baseline
        _WARM.append(True)
    t["A_fwd"], t["A_bwd"] = stage_a()

    with torch.no_grad():  # full-size projected splats for (B) and (C)
        cov = O.quat_scale_to_covar(scene["quats"], scene["scales"])
        radii, means2d, depths, conics, _ = O.fully_fused_projection(scene["means"], cov, scene["viewmats"],
                                                                     scene["Ks"], W, H)
        col = torch.rand(1, radii.shape[1], 3)
        opac = scene["opacities"][None]
    tw, th = -(-W // 16), -(-H // 16)
    t0 = time.perf_counter()
    _, ids, flat = O.isect_tiles(means2d, radii, depths, 16, tw, th)
    offs = O.isect_offset_encode(ids, 1, tw, th)
    t["B_isect"] = time.perf_counter() - t0
    masks = torch.zeros(1, th, tw, dtype=torch.bool)
    pick = [int(round((i + 0.5) * th / rows)) for i in range(rows)]
    masks[0, pick, :] = True
    p = [x.clone().requires_grad_(True) for x in (means2d, conics, col, opac)]
    t0 = time.perf_counter()
    rc, ra = O.rasterize_to_pixels(p[0], p[1], p[2], p[3], W, H, 16, offs, flat, masks=masks)
    (rc.sum() + ra.sum()).backward()
    t["C_raster_sample"] = time.perf_counter() - t0
    # masked-out tiles still cost the python loop a little; measure that overhead and remove it
    t0 = time.perf_counter()
    O.rasterize_to_pixels(means2d, conics, col, opac, W, H, 16, offs, flat, masks=torch.zeros_like(masks))
    t["C_loop_overhead"] = time.perf_counter() - t0
    raster = max(t["C_raster_sample"] - t["C_loop_overhead"] * (th - rows) / th, 1e-6) * th / rows
    est = gauss_stride * (t["A_fwd"] + t["A_bwd"]) + t["B_isect"] + raster
    return est, t


def _load_oracle():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import gs_oracle

    return gs_oracle


def cpu_baseline(rows=2, gauss_stride=8):
    O = _load_oracle()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    scene = make_cfg2("cpu")
    est, parts = _cpu_reference_step(scene, O, rows, gauss_stride)
    return {"value": round(1.0 / est, 6), "unit": "iters/s", "cores": cores, "kind": "port",
            "sample": (f"oracle/gs_oracle.py (port of the reference's _torch_impl path): projection+SH fwd+bwd on every "
                       f"{gauss_stride}th Gaussian (x{gauss_stride}), isect+sort+offsets on the full 1M scene, compositor "
                       f"fwd+bwd on {rows} of 68 tile rows (x68/{rows}); estimated full iteration {est:.1f} s"),
            "parts_s": {k: round(v, 3) for k, v in parts.items()}}


def run_reference(args):
    """`--impl reference`: the reference's CPU path for cfg2 (oracle port) on all host cores, rank 0 only."""
    if int(os.environ.get("RANK", 0)) != 0:
        return
    O = _load_oracle()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    scene = make_cfg2("cpu")
    total = args.steps + args.warmup
    rows, stride = (2, 8) if total <= 12 else (1, 16)
    for _ in range(args.warmup):
        _cpu_reference_step(scene, O, rows, stride)
    ests = [_cpu_reference_step(scene, O, rows, stride)[0] for _ in range(args.steps)]
    est = sum(ests) / len(ests)
    value = round(1.0 / est, 6)
    sample = (f"per step: projection+SH fwd+bwd on every {stride}th Gaussian (x{stride}), isect+sort+offsets on the "
              f"full scene, compositor fwd+bwd on {rows} of 68 tile rows (x68/{rows}); oracle/gs_oracle.py")
    print(json.dumps({
        "impl": "reference", "metric": "fwd+bwd render iters/s (1M Gaussians, 1080p)", "value": value,
        "unit": "iters/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(est * 1e3, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "cfg2: 1M Gaussians SH deg 3, 1920x1080, 1 camera, packed=False (CPU, extrapolated "
                               "from a bounded sample per step)"},
        "cpu_baseline": {"value": value, "unit": "iters/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU leg (profiling runs)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 backend has no CPU path (use --impl reference)")
    run_b200(args)


if __name__ == "__main__":
    main()
