"""bench.py — fwd+bwd render iterations/s of `gsplat.rendering.rasterization()` under GSPLAT_BACKEND=b200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N=1 workload = BASELINE.json configs[1] ("cfg2"): synthetic 1M Gaussians, SH degree 3, one
1920x1080 pinhole camera, packed=False, loss = sum(colors*G) + sum(alphas*g).  N>1 (torchrun, one rank
per GPU, NCCL): the same 1M-Gaussian scene strided over the ranks (`means[rank::N]`), one camera per
rank, `distributed=True` (per-camera all-to-all of the projected rows) — per-GPU work is constant
(weak scaling); `value` = camera renders (fwd+bwd) per second over all ranks.

`--config cfg1` (N=1 only) runs BASELINE.json configs[0] instead: the garden fixture (N = 111 785, camera 0, 648x420,
colours [N,3], `gsplat/_helper.py:9-55` protocol), the only configuration the reference's pure-torch CPU path runs end
to end.

One JSON line on stdout (rank 0): see the keys in `main`. `--impl reference` times WHOLE iterations of the CPU port of
the reference's pure-torch path on the host cores (oracle/gs_oracle.py for projection / SH / intersection,
oracle/gs_compositor.c for the compositor: the reference's own `_rasterize_to_pixels` cannot run anywhere without the
un-vendored nerfacc and a CUDA op) — no sampling, no extrapolation; it stops early only at a wall budget and then
reports the number of whole iterations it measured as `steps`.
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


def make_cfg1(device="cpu"):
    """BASELINE configs[0]: the reference's `load_test_data` protocol (gsplat/_helper.py:9-55, seed 42) on the
    committed copy of its only fixture (tests/golden/test_garden.npz), camera 0."""
    import numpy as np

    data = np.load(os.path.join(ROOT, "tests", "golden", "test_garden.npz"))
    means = torch.from_numpy(data["means3d"]).float()
    colors = torch.from_numpy(data["colors"] / 255.0).float()
    sel = ((means >= -2) & (means <= 2)).all(dim=-1)
    means, colors = means[sel].contiguous(), colors[sel].contiguous()
    g = torch.Generator().manual_seed(42)
    n = means.shape[0]
    scales = torch.rand((n, 3), generator=g) * 0.02
    quats = torch.nn.functional.normalize(torch.randn((n, 4), generator=g), dim=-1)
    opacities = torch.rand((n,), generator=g)
    s = dict(means=means, quats=quats, scales=scales, opacities=opacities, colors=colors,
             viewmats=torch.from_numpy(data["viewmats"]).float()[:1].contiguous(),
             Ks=torch.from_numpy(data["Ks"]).float()[:1].contiguous())
    s = {k: t.to(device) for k, t in s.items()}
    s.update(width=int(data["width"]), height=int(data["height"]))
    return s


WORKLOADS = {
    "cfg2": "cfg2: 1M Gaussians SH deg 3, 1920x1080, 1 camera per GPU, packed=False",
    "cfg1": "cfg1: garden fixture N=111785 colours [N,3], 648x420, camera 0, packed=False",
}
METRICS = {
    "cfg2": "fwd+bwd render iters/s (1M Gaussians, 1080p)",
    "cfg1": "fwd+bwd render iters/s (garden 111785 Gaussians, 648x420)",
}


def workload_string(config, world):
    return WORKLOADS[config] + (f", Gaussians strided over {world} ranks, distributed=True" if world > 1 else "")


def make_scene(config, device="cpu"):
    """-> (scene dict, ordered parameter names, rasterization kwargs)."""
    if config == "cfg1":
        return make_cfg1(device), ("means", "quats", "scales", "opacities", "colors"), dict(packed=False)
    return (make_cfg2(device), ("means", "quats", "scales", "opacities", "shs"),
            dict(packed=False, sh_degree=SH_DEGREE))


def weighted_loss(rc, ra, G, gA):
    """sum(rc * G) + sum(ra * gA), the loss of both arms, as two dot products (no [C,H,W,D] product tensors)."""
    return torch.dot(rc.reshape(-1), G.expand_as(rc).reshape(-1)) + torch.dot(ra.reshape(-1), gA.expand_as(ra).reshape(-1))


def loss_weights(device, seed=1, width=WIDTH, height=HEIGHT):
    g = torch.Generator().manual_seed(seed)
    G = torch.rand(1, height, width, 3, generator=g)
    a = torch.rand(1, height, width, 1, generator=g)
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
    """Roofline denominators: HBM from the driver-written MEASURED_PEAKS.json; FP32-FMA / MUFU.EX2 from
    profiles/peaks_fp32.json (tools/peak_fp32.cu run on this pool's B200), else the arithmetic figure."""
    out = {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "source": "fallback"}
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        out.update(hbm_gbs=float(d["hbm_gbs"]), sm_max_mhz=float(d.get("sm_max_mhz", 1965.0)), source="measured")
    out["fp32_tflops"] = SM_COUNT * FP32_LANES * 2 * out["sm_max_mhz"] * 1e6 / 1e12
    out["fp32_source"] = "arithmetic: 148 SMs * 128 lanes * 2 * sm_max_mhz (profiles/peaks_fp32.json absent)"
    p = os.path.join(ROOT, "profiles", "peaks_fp32.json")
    if os.path.exists(p):
        d = json.load(open(p))
        out["fp32_tflops"] = float(d["fp32_fma"]["burst"])
        out["mufu_tops"] = float(d["mufu_ex2"]["burst"])
        out["fp32_source"] = "measured: profiles/peaks_fp32.json fp32_fma.burst (tools/peak_fp32.cu on this pool's B200)"
    return out


# ------------------------------------------------------------------------------------------------
# the b200 arm
# ------------------------------------------------------------------------------------------------
def _raster_work_counts(meta, device):
    """(E, hits, lane_evals): E = reference-algorithm evaluations 256 * sum_tiles(last contributing index - tile
    start + 1); hits = pixel-Gaussian pairs actually composited (alpha >= 1/255, before termination);
    lane_evals = per kernel, the evaluations it EXECUTES: lanes x the inner-loop iterations counted in-kernel."""
    from gsplat.b200_impl._lib import lib, ptr, stream_of

    means2d, conics, opac = (meta[k].detach().contiguous() for k in ("means2d", "conics", "opacities"))
    offs, flat = meta["isect_offsets"], meta["flatten_ids"]
    C, th, tw = offs.shape
    P, W, H = opac.numel(), meta["width"], meta["height"]
    rec = torch.empty((2, P, 4), device=device)
    ones = torch.ones((P,), device=device)
    cnt, ra = torch.empty((C, H, W), device=device), torch.empty((C, H, W), device=device)
    last = torch.empty((C, H, W), dtype=torch.int32, device=device)
    pairs = torch.zeros((1,), dtype=torch.int64, device=device)
    st = stream_of(rec)
    lib.gsb_rasterize_pack(P, ptr(means2d), ptr(conics), ptr(opac), ptr(rec[0]), ptr(rec[1]), st)
    lib.gsb_rasterize_count(C, flat.numel(), ptr(rec[0]), ptr(rec[1]), ptr(ones), W, H, 16, tw, th, ptr(offs),
                            ptr(flat), ptr(cnt), ptr(ra), ptr(last), ptr(pairs), st)
    hits = int(cnt.double().sum().item())
    lane_evals = {"gsb_rasterize_bwd": 32 * int(pairs.item())}
    if int(lib.gsb_get_option(0)) >= 2:  # forward v2 / v3: one loop iteration = 2 blocks x 32 pixels
        iters = torch.zeros((1,), dtype=torch.int64, device=device)
        s1, s2 = torch.empty_like(cnt), torch.empty_like(ra)
        s3 = torch.empty_like(last)
        lib.gsb_rasterize_fwd2_count(C, flat.numel(), ptr(rec[0]), ptr(rec[1]), ptr(ones), W, H, tw, th, ptr(offs),
                                     ptr(flat), ptr(s1), ptr(s2), ptr(s3), ptr(iters), st)
        lane_evals["gsb_rasterize_fwd_cap"] = 64 * int(iters.item())
        if int(lib.gsb_get_option(1)) == 2:
            # backward v2 walks the same per-half lists (cut at each block's last contributor): same count, an
            # upper bound by the few entries behind the last contributor
            lane_evals["gsb_rasterize_bwd"] = 64 * int(iters.item())
    else:
        lane_evals["gsb_rasterize_fwd_cap"] = 32 * int(pairs.item())
    lane_evals["gsb_rasterize_fwd"] = lane_evals["gsb_rasterize_fwd_cap"]
    last = torch.where(cnt > 0, last, torch.full_like(last, -1))
    pad_h, pad_w = th * 16 - H, tw * 16 - W
    last = torch.nn.functional.pad(last, (0, pad_w, 0, pad_h), value=-1)
    tile_last = last.reshape(C, th, 16, tw, 16).amax(dim=(2, 4))
    work = (tile_last - offs + 1).clamp(min=0)
    work = torch.where(tile_last >= 0, work, torch.zeros_like(work))
    return 256 * int(work.sum().item()), hits, lane_evals


def _stage_table(summary, steps, meta, n_local, C, D, K, work, peaks, sm_mhz):
    """Per-entry-point ms/step, algorithmic bytes (SURVEY 8d formulas) or FP32 slots, achieved vs peak."""
    P = C * n_local
    I = meta["flatten_ids"].numel()
    n_tiles = meta["tile_width"] * meta["tile_height"]
    bits = 32 + (n_tiles - 1).bit_length() + max(C - 1, 0).bit_length()
    binned = "gsb_isect_bin" in summary  # tile binning instead of the two radix passes over the intersections
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
        "gsb_radix_sort_pairs_u32": 4 * 16 * P + 4 * P + (0 if binned else math.ceil((bits - 32) / 8) * 16 * I + 4 * I),
        "gsb_isect_row_tiles_keys": 16 * P + 28 * P,
        # the producers of the keys count the digits: the sorts run their passes only
        "gsb_radix_sort_pairs_u32_prepared": 4 * 16 * P + (0 if binned else math.ceil((bits - 32) / 8) * 16 * I),
        "gsb_isect_emit_perm_hist": 20 * P + 8 * I,
        "gsb_isect_emit_chunks": 20 * P + 8 * I,
        "gsb_isect_count_perm_records": 20 * P + 16 * P,
        # the rows' records twice (count, scatter) + perm / depth bits once, 12 bytes per intersection out, offsets;
        # the chunks x bins count matrix the kernels pass between them is overhead, not counted
        "gsb_isect_bin": 2 * 16 * P + 8 * P + 12 * I + 4 * C * n_tiles,
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
    fp32_peak = peaks["fp32_tflops"]
    slots = {"gsb_rasterize_fwd": SLOTS_FWD, "gsb_rasterize_fwd_cap": SLOTS_FWD, "gsb_rasterize_bwd": SLOTS_BWD}
    out = {}
    for name, (calls, ms) in summary.items():
        per_step = ms / steps
        row = {"ms_per_step": round(per_step, 4), "launches_per_step": calls / steps}
        if name in hbm and per_step > 0:
            a = hbm[name] / (per_step * 1e-3) / 1e9
            row.update(bound="hbm", achieved=round(a, 1), peak=peaks["hbm_gbs"], unit="GB/s",
                       frac=round(a / peaks["hbm_gbs"], 4), algorithmic_bytes=hbm[name])
        elif name in slots and per_step > 0 and work is not None:
            # Three counts of pixel-Gaussian evaluations, each x the SURVEY 8(d) FP32 issue slots per evaluation:
            #   frac                      what the kernel EXECUTES: 32 lanes x the (warp, splat) pairs it walks
            #                             (counted in-kernel by gsb_rasterize_count) -- the roofline fraction;
            #   frac_composited_pairs     only the pairs that really blend (alpha >= 1/255 before termination):
            #                             useful work, a strict lower bound;
            #   frac_reference_algorithm  the reference's dense loop E (every pixel of a tile x every Gaussian up to
            #                             the tile's last contributor). The kernels cull most of E exactly, so this
            #                             one can exceed 1: it is a speed-up over the dense algorithm, not a fraction.
            evals, hits, lane_evals = work[0], work[1], work[2][name]
            to_tf = 2 * slots[name] / (per_step * 1e-3) / 1e12
            row.update(bound="fp32", achieved=round(lane_evals * to_tf, 2), peak=round(fp32_peak, 2), unit="TFLOP/s",
                       frac=round(lane_evals * to_tf / fp32_peak, 4), executed_lane_evaluations=lane_evals,
                       composited_pairs=hits, frac_composited_pairs=round(hits * to_tf / fp32_peak, 4),
                       reference_evaluations=evals, frac_reference_algorithm=round(evals * to_tf / fp32_peak, 4),
                       issue_slots_per_evaluation=slots[name])
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
    # one process per GPU: stage the pinned host buffers of the end-to-end leg on the GPU's own NUMA node
    numa, all_cpus = None, os.sched_getaffinity(0)
    if os.environ.get("GSPLAT_BENCH_NUMA", "1") == "1":
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from numa_bind import bind_to_gpu

        numa = bind_to_gpu(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from gsplat.b200_impl._lib import lib, timing
    from gsplat.rendering import rasterization

    config = args.config
    assert config == "cfg2" or world == 1, "--config cfg1 is a single-GPU configuration"
    host, names, render_kw = make_scene(config, "cpu")
    WIDTH, HEIGHT = host["width"], host["height"]
    host_params = [host[k][rank::world].contiguous().pin_memory() for k in names]
    viewmats, Ks = host["viewmats"].to(dev), host["Ks"].to(dev)
    G, gA = loss_weights(dev, width=WIDTH, height=HEIGHT)
    n_local = host_params[0].shape[0]

    from gsplat.distributed import host_trace

    def iteration(params, distributed=(world > 1)):
        host_trace.mark()
        rc, ra, meta = rasterization(*params, viewmats, Ks, WIDTH, HEIGHT, distributed=distributed, **render_kw)
        host_trace.mark()
        loss = weighted_loss(rc, ra, G, gA)
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
    d2h_image = out_c.numel() * 4 + out_a.numel() * 4 + 4
    d2h = 4
    assert all(t.is_pinned() for t in host_params) and out_c.is_pinned()

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

    def e2e_loop(steps, read_image):
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
                if read_image:
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

    # headline e2e: the step's RESULT read back by the host is the loss (what a training loop reads); the variant that
    # also copies the rendered image and alpha back every step (33 MB) is reported beside it
    e2e_loop(3, False)
    barrier()
    # K steps, twice: host<->device copies share the PCIe switch / host memory with whatever else the host runs, and a
    # whole K-step run has been seen 30 % slower than its neighbour on the same box. Both runs are printed; the line's
    # e2e value is the faster one (max over ranks of each run first).
    e2e_runs = []
    for _ in range(2):
        e2e_runs.append(e2e_loop(args.steps, False))
        barrier()
    e2e_loop(2, True)
    barrier()
    e2e_img_ms = e2e_loop(args.steps, True)
    barrier()
    clocks.__exit__(None, None, None)

    ms = torch.tensor([total_ms / args.steps, e2e_runs[0] / args.steps, e2e_runs[1] / args.steps,
                       e2e_img_ms / args.steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step, ms_e2e_a, ms_e2e_b, ms_e2e_img = ms.tolist()
    ms_e2e = min(ms_e2e_a, ms_e2e_b)

    # one more (untimed) iteration on EVERY rank (it contains collectives) to read the stage geometry from `meta`
    for p in params:
        p.grad = None
    rc_d, ra_d, _, meta = iteration(params)
    barrier()
    parity = None
    if world > 1 and rank == 0:
        # multi-GPU parity, checked in every recorded run: rank 0 renders its camera from the WHOLE scene on its own
        # GPU (distributed=False) and compares the distributed image / alpha and the gradients of its shard. Every
        # rank holds the same camera and loss weights here, so the shard's distributed gradient is world x the
        # single-camera one.
        full = [host[k].to(dev).requires_grad_(True) for k in names]
        rc_s, ra_s, _, _ = iteration(full, distributed=False)
        g_rel = max(float(((p.grad - world * f.grad[rank::world]).abs().max()
                           / (world * f.grad[rank::world].abs().max() + 1e-30)).item())
                    for p, f in zip(params, full))
        bad = ((rc_d - rc_s).abs() > 1e-4 + 1e-4 * rc_s.abs()).any(-1).float().mean()
        parity = {"img_max_err": float((rc_d - rc_s).abs().max()), "alpha_max_err": float((ra_d - ra_s).abs().max()),
                  "img_frac_beyond_1e-4": float(bad), "grad_max_rel": g_rel,
                  "what": "rank 0: distributed render of its camera vs its own single-GPU render of the whole scene; "
                          "gradients of its Gaussian shard vs world x the single-GPU ones (max |diff| / max |grad|)"}
        del full
    barrier()
    if rank == 0:
        # exchange rows are re-laid out after `meta` is filled when distributed: count on a single GPU only
        # exchange rows are re-laid out after `meta` is filled when distributed: count on a single GPU only
        work = _raster_work_counts(meta, dev) if world == 1 else None
        peaks = measured_peaks()
        clk = clocks.summary()
        C_local, D = 1, 3
        K = (SH_DEGREE + 1) ** 2 if config == "cfg2" else 0
        n_rows = meta["means2d"].shape[1]
        stages = _stage_table(stage_summary, stage_steps, meta, n_rows, C_local, D, K, work, peaks, clk["sm_mhz"])
        top = max((k for k in stages if "bound" in stages[k]), key=lambda k: stages[k]["ms_per_step"])
        roof = {k: stages[top][k] for k in ("bound", "achieved", "peak", "unit", "frac")}
        for k in ("frac_composited_pairs", "frac_reference_algorithm"):
            if k in stages[top]:
                roof[k] = stages[top][k]
        traffic_file = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        traffic = json.load(open(traffic_file)).get(top) if os.path.exists(traffic_file) and config == "cfg2" else None
        roof.update(kernel=top, traffic=traffic,
                    traffic_source="profiles/ncu_traffic.json: dram__bytes_read+write per launch from a separate "
                                   "`ncu --set full` capture of this workload, NOT measured in this run",
                    peak_source=(peaks["fp32_source"] if roof["bound"] == "fp32"
                                 else peaks["source"] + " MEASURED_PEAKS.json hbm_gbs"),
                    share_of_step=round(stages[top]["ms_per_step"] / ms_step, 4))
        line = {
            "metric": METRICS[config], "value": round(world * 1e3 / ms_step, 3),
            "unit": "iters/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic" if config == "cfg2" else "reference fixture (garden) + seeded random attributes",
            "config": {"workload": workload_string(config, world),
                       "n_gaussians_per_gpu": n_local, "n_isects": meta["flatten_ids"].numel(),
                       "l2": "a 256 MB flush write between steps (untimed)"
                             + ("; inputs 236 MB/step > 126 MB L2" if config == "cfg2" else ""),
                       "timing": "sum of per-step CUDA-event times on the launch stream, max over ranks"},
            "e2e": {"value": round(world * 1e3 / ms_e2e, 3), "unit": "iters/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": round(ms_e2e, 4),
                    "runs_ms_per_step": [round(ms_e2e_a, 4), round(ms_e2e_b, 4)],
                    "runs_note": "two back-to-back runs of K steps each; value = the faster (PCIe / host noise)",
                    "pipeline": "every step copies ALL of its inputs (the rank's Gaussian parameters) from pinned host "
                                "memory, double-buffered on a copy stream (step i+1's copy overlaps step i's kernels), "
                                "and the host reads the step's loss back on a read-back stream, one step behind the "
                                "launches; one CUDA-event bracket around all K steps; max over ranks",
                    "with_image_readback": {
                        "value": round(world * 1e3 / ms_e2e_img, 3), "ms_per_step": round(ms_e2e_img, 4),
                        "d2h_bytes_per_step": d2h_image,
                        "note": "the same loop also copying the rendered image and alpha to the host every step (round "
                                "1's definition). On an 8-GPU box it is bound by the host: concurrent pinned D2H "
                                "reaches 121 GB/s over all 8 GPUs (12-18 GB/s each, 55 GB/s alone; "
                                "profiles/r02p_multigpu/r02p_h2d8.json)"}},
            "gpu_launches": launches, "wall_s_timed_region": round(wall_s, 3), "clocks": clk,
            "roofline": roof, "stages": stages,
        }
        if numa is not None:
            line["e2e"]["numa_binding_rank0"] = numa
        if parity is not None:
            line["parity"] = parity
        if host_trace.enabled:
            line["host_trace_us"] = host_resident
        if world == 1 and not args.no_cpu_baseline:
            os.sched_setaffinity(0, all_cpus)  # the CPU leg uses every host core again
            line["cpu_baseline"] = cpu_baseline(config)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# the CPU arm: WHOLE iterations of the port of the reference's pure-torch path on the host cores
# ------------------------------------------------------------------------------------------------
def _load_oracle():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import c_oracle
    import gs_oracle

    c_oracle.build()
    return gs_oracle


def _load_reference_torch_impl():
    """The reference's own gsplat/cuda/_torch_impl.py from baseline/_ref (tools/install_reference.sh), loaded as a
    stand-alone module (its only package-relative import sits in `_rasterize_to_pixels`, which cannot run: nerfacc +
    a CUDA op). None when baseline/_ref is absent."""
    import importlib.util

    path = os.path.join(ROOT, "baseline", "_ref", "gsplat", "cuda", "_torch_impl.py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("_gsplat_ref_torch_impl", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class CpuArm:
    """One fwd+bwd iteration of `rasterization()` on the CPU, same inputs and loss as the b200 arm.

    cfg2: oracle/gs_oracle.py (restated `_torch_impl` projection / SH / vectorised isect_tiles / offsets) +
          oracle/gs_compositor.c (OpenMP) — kind "port".
    cfg1: the reference's OWN `_torch_impl` functions for covariance, projection and offsets when baseline/_ref is
          present (else the restatement), the vectorised isect_tiles restatement (the reference's is a Python double
          loop: timed once and reported beside), the C compositor. `impl` in the report says which ran per stage."""

    def __init__(self, config):
        self.O = _load_oracle()
        self.config = config
        self.scene, self.names, kw = make_scene(config, "cpu")
        self.kw = {k: v for k, v in kw.items() if k != "packed"}
        self.G, self.gA = loss_weights("cpu", width=self.scene["width"], height=self.scene["height"])
        self.T = _load_reference_torch_impl() if config == "cfg1" else None
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.impl = {
            "covariance+projection": "reference gsplat/cuda/_torch_impl.py (baseline/_ref)" if self.T else
                                     "port oracle/gs_oracle.py",
            "spherical_harmonics": "port oracle/gs_oracle.py" if config == "cfg2" else "n/a (colours [N,3])",
            "isect_tiles": "port oracle/gs_oracle.py (vectorised; the reference's is a Python double loop)",
            "isect_offset_encode": "reference _torch_impl.py" if self.T else "port oracle/gs_oracle.py",
            "rasterize_to_pixels fwd+bwd": "port oracle/gs_compositor.c (OpenMP, all cores)",
        }

    def step(self):
        """-> (seconds, per-stage forward seconds)."""
        s, O = self.scene, self.O
        leaves = [s[k].clone().requires_grad_(True) for k in self.names]
        timers = {}
        t0 = time.perf_counter()
        if self.T is None:
            rc, ra, _ = O.rasterization(*leaves, s["viewmats"], s["Ks"], s["width"], s["height"], compositor="c",
                                        timers=timers, **self.kw)
        else:
            rc, ra = self._reference_pipeline(leaves, timers)
        t1 = time.perf_counter()
        weighted_loss(rc, ra, self.G, self.gA).backward()
        t2 = time.perf_counter()
        timers.update(forward_s=t1 - t0, backward_s=t2 - t1)
        return t2 - t0, timers

    def _reference_pipeline(self, leaves, timers):
        import c_oracle

        s, O, T = self.scene, self.O, self.T
        means, quats, scales, opac, colors = leaves
        W, H, C = s["width"], s["height"], s["viewmats"].shape[0]
        t0 = time.perf_counter()
        covars, _ = T._quat_scale_to_covar_preci(quats, scales, True, False, triu=False)
        radii, means2d, depths, conics, _ = T._fully_fused_projection(means, covars, s["viewmats"], s["Ks"], W, H)
        t1 = time.perf_counter()
        tw, th = -(-W // 16), -(-H // 16)
        _, ids, flat = O.isect_tiles(means2d.detach(), radii, depths.detach(), 16, tw, th, bits_rule="cuda")
        offs = T._isect_offset_encode(ids, C, tw, th)
        t2 = time.perf_counter()
        rc, ra = c_oracle.rasterize_to_pixels(means2d, conics, colors.expand(C, -1, -1), opac.repeat(C, 1), W, H, 16,
                                              offs, flat)
        timers.update(projection_colors_fwd_s=t1 - t0, isect_sort_offsets_s=t2 - t1,
                      compositor_fwd_s=time.perf_counter() - t2)
        self._projected = (means2d.detach(), radii, depths.detach(), tw, th)
        return rc, ra

    def reference_isect_loop_seconds(self):
        """The reference's own `_isect_tiles` (Python double loop, _torch_impl.py:331) on the projected scene."""
        if self.T is None or not hasattr(self, "_projected"):
            return None
        m, r, d, tw, th = self._projected
        t0 = time.perf_counter()
        self.T._isect_tiles(m, r, d, 16, tw, th)
        return time.perf_counter() - t0

    def describe(self, steps_measured, seconds):
        return (f"{steps_measured} WHOLE fwd+bwd iteration(s) of {WORKLOADS[self.config]} on {self.cores} host threads "
                f"({seconds:.1f} s measured), nothing sampled or extrapolated")


def cpu_baseline(config="cfg2", budget_s=25.0):
    """The `cpu_baseline` object of the b200 line: whole CPU iterations (one warm-up, then as many as fit the
    budget, at least one)."""
    arm = CpuArm(config)
    arm.step()
    times, parts = [], {}
    while not times or (sum(times) + times[-1] < budget_s and len(times) < 20):
        t, parts = arm.step()
        times.append(t)
    mean = sum(times) / len(times)
    out = {"value": round(1.0 / mean, 6), "unit": "iters/s", "cores": arm.cores, "kind": "port",
           "sample": arm.describe(len(times), sum(times)), "steps_measured": len(times),
           "s_per_step": round(mean, 3), "impl": arm.impl, "parts_s": {k: round(v, 3) for k, v in parts.items()}}
    if config == "cfg1":
        loop = arm.reference_isect_loop_seconds()
        if loop is not None:
            out["reference_isect_tiles_python_loop_s"] = round(loop, 2)
    return out


def run_reference(args):
    """`--impl reference`: whole CPU iterations of the same workload, rank 0 only, on all host cores."""
    if int(os.environ.get("RANK", 0)) != 0:
        return
    config = args.config
    arm = CpuArm(config)
    budget = float(os.environ.get("GSPLAT_BENCH_CPU_BUDGET_S", 420.0))
    t_begin = time.perf_counter()
    warm = 0
    for _ in range(args.warmup):
        # warm-ups are whole iterations too; they stop early when they would eat more than a third of the budget
        if warm and time.perf_counter() - t_begin > budget / 3:
            break
        arm.step()
        warm += 1
    times, parts = [], {}
    for _ in range(args.steps):
        t, parts = arm.step()
        times.append(t)
        if time.perf_counter() - t_begin + t > budget:
            break
    mean = sum(times) / len(times)
    value = round(1.0 / mean, 6)
    base = {"value": value, "unit": "iters/s", "cores": arm.cores, "kind": "port",
            "sample": arm.describe(len(times), sum(times)), "impl": arm.impl,
            "parts_s": {k: round(v, 3) for k, v in parts.items()}}
    if config == "cfg1":
        loop = arm.reference_isect_loop_seconds()
        if loop is not None:
            base["reference_isect_tiles_python_loop_s"] = round(loop, 2)
    print(json.dumps({
        "impl": "reference", "metric": METRICS[config], "value": value,
        "unit": "iters/s", "n_gpus": args.gpus, "steps": len(times), "steps_requested": args.steps,
        "warmup": warm, "warmup_requested": args.warmup,
        "ms_per_step": round(mean * 1e3, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic" if config == "cfg2" else "reference fixture (garden) + seeded random attributes",
        "config": {"workload": workload_string(config, args.gpus),
                   "cpu_arm": "whole iterations on the host cores of rank 0 (one camera render per iteration); "
                              f"wall budget {budget:.0f} s: `steps` is the number of whole iterations measured"},
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": "iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="cfg2", choices=["cfg2", "cfg1"],
                    help="cfg2 = BASELINE configs[1] (the metric's configuration); cfg1 = configs[0] (garden, N=1 only)")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU leg (profiling runs)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 backend has no CPU path (use --impl reference)")
    run_b200(args)


if __name__ == "__main__":
    main()
