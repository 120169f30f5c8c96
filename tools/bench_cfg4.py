"""BASELINE.json configs[3] ("cfg4"): the cfg3 scene with 6 M Gaussians strided over the ranks of one box,
one local camera per rank (C_global = world size, ring of cameras at 1297x840), `distributed=True`, SH degree 3,
antialiased, fwd+bwd. Prints (rank 0) ms/step as the max over ranks, camera renders/s over all ranks, the exchange
volume per rank and direction and the per-entry-point table (CUDA events) of rank 0.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 \
        tools/bench_cfg4.py                      # CFG4_PACKED=1 for packed rows (NCCL transport), CFG4_N=6000000

Not the bench line (bench.py measures cfg2 and its sharded version); the data point for the configuration BASELINE.json
names for the distributed path. The exchange volume is counted from the rows really sent (unpacked: every local row to
every camera owner; packed: the visible rows), the achieved NVLink rate = bytes that leave this GPU / duration of the put
kernel (CUDA events), against 900 GB/s nominal / 770 GB/s measured peer copy per direction (B200_PROFILING.md).
"""
import json
import math
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gsplat-triton_b200"))
os.environ.setdefault("GSPLAT_BACKEND", "b200")


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from gsplat.b200_impl._lib import timing
    from gsplat.rendering import rasterization

    N_total, W, H = int(os.environ.get("CFG4_N", 6_000_000)), 1297, 840
    packed = os.environ.get("CFG4_PACKED", "0") == "1"
    steps = int(os.environ.get("CFG4_STEPS", 10))
    g = torch.Generator().manual_seed(0)  # every rank draws the same scene and keeps its stride
    means = (torch.rand(N_total, 3, generator=g) * 6 - 3)[rank::world].contiguous().to(dev)
    quats = torch.randn(N_total, 4, generator=g)[rank::world].contiguous().to(dev)
    lo, hi = math.log(0.002), math.log(0.02)
    scales = torch.exp(torch.rand(N_total, 3, generator=g) * (hi - lo) + lo)[rank::world].contiguous().to(dev)
    opac = torch.rand(N_total, generator=g)[rank::world].contiguous().to(dev)
    gs = torch.Generator().manual_seed(100 + rank)  # the SH table is the bulk of the data: drawn per rank
    n_local = means.shape[0]
    shs = torch.randn(n_local, 16, 3, generator=gs) * 0.3
    shs[:, 0] += 0.5
    shs = shs.to(dev)
    f = W / (2 * math.tan(math.radians(30)))
    az = 2 * math.pi * rank / max(world, 1)  # this rank's camera on the ring of radius 6, looking at the origin
    eye = torch.tensor([6 * math.cos(az), 0.0, 6 * math.sin(az)])
    fwd = -eye / eye.norm()
    right = torch.linalg.cross(torch.tensor([0.0, 1.0, 0.0]), fwd)
    right = right / right.norm()
    up = torch.linalg.cross(fwd, right)
    R = torch.stack([right, up, fwd])
    vm = torch.eye(4)
    vm[:3, :3], vm[:3, 3] = R, -R @ eye
    viewmats = vm[None].to(dev)
    Ks = torch.tensor([[[f, 0, W / 2], [0, f, H / 2], [0, 0, 1.0]]], device=dev)
    gw = torch.Generator().manual_seed(1)
    G = torch.rand(1, H, W, 3, generator=gw).to(dev)
    gA = torch.rand(1, H, W, 1, generator=gw).to(dev)
    params = [t.requires_grad_(True) for t in (means, quats, scales, opac, shs)]

    def step():
        for p in params:
            p.grad = None
        rc, ra, meta = rasterization(*params, viewmats, Ks, W, H, sh_degree=3, packed=packed,
                                     rasterize_mode="antialiased", distributed=world > 1)
        ((rc * G).sum() + (ra * gA).sum()).backward()
        return meta

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(3):
        meta = step()
    barrier()
    evs = []
    for _ in range(steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step()
        b.record()
        evs.append((a, b))
    barrier()
    ms = torch.tensor([sum(a.elapsed_time(b) for a, b in evs) / steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    timing.reset(True)
    for _ in range(5):
        step()
    table = {k: round(v[1] / 5, 4) for k, v in timing.summary().items()}
    timing.reset(False)
    barrier()
    if rank == 0:
        D = 3
        if packed:  # rows this rank keeps visible for any camera = rows it sends (incl. to itself)
            rows_out = int(meta["_rows_sent"]) if "_rows_sent" in meta else None
        else:
            rows_out = n_local * world  # every local Gaussian goes to every camera owner
        out = {
            "workload": f"cfg4: {N_total} Gaussians over {world} ranks, 1 camera per rank at {W}x{H}, packed={packed}, "
                        "antialiased, SH deg 3, distributed",
            "world": world, "packed": packed,
            "ms_per_step": round(float(ms), 3), "camera_renders_per_s": round(world * 1e3 / float(ms), 1),
            "n_local": n_local, "n_isects_rank0": int(meta["flatten_ids"].numel()),
            "rows_received_rank0": int(meta["radii"].numel()),
            "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 2), "stages_ms_rank0": table}
        if world > 1:
            rows_in = int(meta["radii"].numel())
            rows = rows_out if rows_out is not None else rows_in  # packed: symmetric scene, sent ~= received
            fwd_b = rows * ((8 + D) * 4 + (8 if packed else 0))
            bwd_b = rows_in * (7 + D) * 4
            nv = (world - 1) / world
            ex = {"fwd_MB_per_rank": round(fwd_b / 1e6, 1), "bwd_MB_per_rank": round(bwd_b / 1e6, 1),
                  "over_nvlink_fraction": round(nv, 3)}
            for key, name, b in (("fwd", "gsb_a2a_put_packed_fwd" if packed else "gsb_a2a_put_fwd", fwd_b),
                                 ("bwd", "gsb_a2a_put_packed_bwd" if packed else "gsb_a2a_put_bwd", bwd_b)):
                if name in table and table[name] > 0:
                    ex[f"{key}_put_ms"] = table[name]
                    ex[f"{key}_nvlink_GBps_achieved"] = round(b * nv / (table[name] * 1e-3) / 1e9, 1)
            ex["nvlink_GBps_nominal_per_direction"] = 900.0
            ex["nvlink_GBps_measured_peer_copy"] = 770.0
            out["exchange"] = ex
        print("CFG4 " + json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
