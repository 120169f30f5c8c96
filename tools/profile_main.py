"""The reference's profiling protocol (profiling/main.py:28-140) under GSPLAT_BACKEND=b200.

    python tools/profile_main.py [--reso 1080p] [--scene-grid 5] [--batch-size 1] [--channels 3]
                                 [--packed/--no-packed] [--sparse-grad] [--repeats 100]
    python -m torch.distributed.run --nproc-per-node N tools/profile_main.py ...   (Gaussians strided over ranks)

Same scene construction (garden fixture tiled `scene_grid` x `scene_grid`, random scales/quats/opacities,
first camera repeated `batch_size` times, intrinsics rescaled to the resolution), same arguments to
`rasterization()` (near 0.01, far 100, radius_clip 3), same measurements: 5 warm-ups, `repeats` forward calls,
then `repeats` backward calls of `render_colors.sum()` with retain_graph, wall clock around a synchronize, and
the peak allocation. Prints the reference's one-line summary.
"""
import argparse
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gsplat-triton_b200"))
os.environ.setdefault("GSPLAT_BACKEND", "b200")

RESOLUTIONS = {"360p": (640, 360), "720p": (1280, 720), "1080p": (1920, 1080), "4k": (3840, 2160)}


def tiled_garden(device, scene_grid=1, crop=2.0, seed=42):
    """gsplat/_helper.py:9-55 on the committed fixture: crop to [-2, 2]^3, tile in x/y, random attributes."""
    assert scene_grid % 2 == 1, "scene_grid must be odd"
    data = np.load(os.path.join(ROOT, "tests", "golden", "test_garden.npz"))
    height, width = int(data["height"]), int(data["width"])
    viewmats = torch.from_numpy(data["viewmats"]).float().to(device)
    Ks = torch.from_numpy(data["Ks"]).float().to(device)
    means = torch.from_numpy(data["means3d"]).float().to(device)
    colors = torch.from_numpy(data["colors"] / 255.0).float().to(device)
    keep = ((means >= -crop) & (means <= crop)).all(dim=-1)
    means, colors = means[keep], colors[keep]
    half = scene_grid // 2
    ticks = torch.arange(-half, half + 1, device=device, dtype=torch.float32)
    gx, gy = torch.meshgrid(ticks, ticks, indexing="ij")
    shift = torch.stack([gx, gy, torch.zeros_like(gx)], dim=-1).reshape(-1, 1, 3) * (2 * crop)
    means = (means[None] + shift).reshape(-1, 3)
    colors = colors.repeat(scene_grid**2, 1)
    g = torch.Generator(device="cpu").manual_seed(seed)
    n = means.shape[0]
    scales = (torch.rand((n, 3), generator=g) * 0.02).to(device)
    quats = torch.nn.functional.normalize(torch.randn((n, 4), generator=g), dim=-1).to(device)
    opacities = torch.rand((n,), generator=g).to(device)
    return means, quats, scales, opacities, colors, viewmats, Ks, width, height


def timeit(repeats, fn, *args, **kwargs):
    for _ in range(5):
        fn(*args, **kwargs)
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(repeats):
        out = fn(*args, **kwargs)
    torch.cuda.synchronize()
    return (time.time() - t0) / repeats, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch-size", type=int, default=1)
    ap.add_argument("--channels", type=int, default=3)
    ap.add_argument("--reso", default="1080p", choices=list(RESOLUTIONS))
    ap.add_argument("--scene-grid", type=int, default=5)
    ap.add_argument("--packed", action=argparse.BooleanOptionalAction, default=True)
    ap.add_argument("--sparse-grad", action="store_true")
    ap.add_argument("--repeats", type=int, default=100)
    args = ap.parse_args()

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=device)
    from gsplat.rendering import rasterization

    means, quats, scales, opacities, colors, viewmats, Ks, width, height = tiled_garden(device, args.scene_grid)
    viewmats = viewmats[:1].repeat(args.batch_size, 1, 1)
    Ks = Ks[:1].repeat(args.batch_size, 1, 1)
    colors = colors[:, :1].repeat(1, args.channels)
    params = [t[rank::world].contiguous().requires_grad_(True) for t in (means, quats, scales, opacities, colors)]
    rw, rh = RESOLUTIONS[args.reso]
    Ks[..., 0, :] *= rw / width
    Ks[..., 1, :] *= rh / height

    torch.cuda.reset_peak_memory_stats()
    base = torch.cuda.max_memory_allocated() / 1024**3
    t_fwd, out = timeit(args.repeats, rasterization, *params, viewmats, Ks, rw, rh, packed=args.packed,
                        near_plane=0.01, far_plane=100.0, radius_clip=3.0, sparse_grad=args.sparse_grad,
                        distributed=world > 1)
    mem_fwd = torch.cuda.max_memory_allocated() / 1024**3 - base
    loss = out[0].sum()

    def backward():
        loss.backward(retain_graph=True)
        for v in params:
            v.grad = None

    t_bwd, _ = timeit(args.repeats, backward)
    mem_all = torch.cuda.max_memory_allocated() / 1024**3 - base
    print(f"[rank {rank}] Rasterization Mem Allocation: [FWD]{mem_fwd:.2f} GB, [All]{mem_all:.2f} GB "
          f"Time: [FWD]{t_fwd:.4f}s, [BWD]{t_bwd:.4f}s N Gaussians: {params[0].shape[0]} "
          f"n_isects: {out[2]['flatten_ids'].numel()}", flush=True)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
