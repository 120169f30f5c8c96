"""BASELINE.json configs[4] ("cfg5"): rasterize_to_pixels forward and backward alone on synthetic,
already-projected Gaussians of radius ~12 px: resolutions 1080p / 1440p / 4K x target densities 16 / 64 / 256 / 1024
Gaussians per tile (N scaled), D = 3. Prints one JSON line per point: kernel ms (CUDA events around the C-ABI
launches), intersections per tile, and the reference-algorithm evaluation rate I*256/ms against the FP32 figure of
DESIGN section 4 (E*14 / E*40 issue slots, an upper bound on E because of early termination).

    python tools/sweep_cfg5.py [--reps 10] [--max-isects 40000000]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gsplat-triton_b200"))
from gsplat.b200_impl import isect_offset_encode, isect_tiles, rasterize_to_pixels  # noqa: E402
from gsplat.b200_impl._lib import timing  # noqa: E402

RESOLUTIONS = {"1080p": (1920, 1080), "1440p": (2560, 1440), "4k": (3840, 2160)}
DENSITIES = (16, 64, 256, 1024)
TILE = 16


def scene(W, H, density, dev, seed=0):
    g = torch.Generator().manual_seed(seed)
    tw, th = -(-W // TILE), -(-H // TILE)
    # a disc of radius r px touches about (2r/16 + 1)^2 tiles
    radius = 12.0
    per_gauss = (2 * radius / TILE + 1) ** 2
    N = max(int(density * tw * th / per_gauss), 1)
    means2d = torch.rand(N, 2, generator=g) * torch.tensor([float(W), float(H)])
    sigma = radius / 3.0 * (0.9 + 0.2 * torch.rand(N, generator=g))  # 3 sigma ~ 12 px
    rho = torch.rand(N, generator=g) * 0.6 - 0.3
    sx, sy = sigma, sigma * (0.8 + 0.4 * torch.rand(N, generator=g))
    cxx, cyy, cxy = sx * sx, sy * sy, rho * sx * sy
    det = cxx * cyy - cxy * cxy
    conics = torch.stack([cyy / det, -cxy / det, cxx / det], dim=-1)
    radii = torch.ceil(3.0 * torch.sqrt(torch.maximum(cxx, cyy))).int()
    depths = torch.rand(N, generator=g) * 9.0 + 1.0
    opacities = torch.rand(N, generator=g) * 0.5 + 0.02
    colors = torch.rand(N, 3, generator=g)
    return [t[None].contiguous().to(dev) for t in (means2d, conics, colors, opacities, radii, depths)], tw, th


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--max-isects", type=int, default=40_000_000)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    sm_clock_mhz = 1965.0
    fp32_peak = 148 * 128 * 2 * sm_clock_mhz * 1e6  # FLOP/s, as bench.py's fallback peak
    for name, (W, H) in RESOLUTIONS.items():
        for density in DENSITIES:
            (means2d, conics, colors, opacities, radii, depths), tw, th = scene(W, H, density, dev)
            _, isect_ids, flatten_ids = isect_tiles(means2d, radii, depths, TILE, tw, th)
            n_isects = flatten_ids.numel()
            if n_isects > args.max_isects:
                print(json.dumps({"reso": name, "density": density, "skipped": f"{n_isects} intersections"}))
                continue
            offsets = isect_offset_encode(isect_ids, 1, tw, th)
            leaves = [t.clone().requires_grad_(True) for t in (means2d, conics, colors, opacities)]
            gw = torch.Generator().manual_seed(1)
            w_c = torch.rand(1, H, W, 3, generator=gw).to(dev)
            w_a = torch.rand(1, H, W, 1, generator=gw).to(dev)

            def step():
                rc, ra = rasterize_to_pixels(leaves[0], leaves[1], leaves[2], leaves[3], W, H, TILE, offsets,
                                             flatten_ids)
                torch.autograd.grad((rc * w_c).sum() + (ra * w_a).sum(), leaves)
                return ra

            for _ in range(3):
                ra = step()
            timing.reset(True)
            for _ in range(args.reps):
                step()
            summary = timing.summary()
            timing.reset(False)
            fwd = summary["gsb_rasterize_fwd_cap"][1] / args.reps
            bwd = summary["gsb_rasterize_bwd"][1] / args.reps
            evals = n_isects * TILE * TILE  # every pixel of a tile x every intersection of the tile (upper bound on E)
            print(json.dumps({
                "reso": name, "density_target": density, "n_gaussians": means2d.shape[1], "n_isects": n_isects,
                "isects_per_tile": round(n_isects / (tw * th), 1), "mean_alpha": round(float(ra.detach().mean()), 3),
                "fwd_ms": round(fwd, 4), "bwd_ms": round(bwd, 4),
                "fwd_gevals_per_s": round(evals / fwd / 1e6, 1), "bwd_gevals_per_s": round(evals / bwd / 1e6, 1),
                "fwd_frac_fp32_upper": round(evals * 14 * 2 / (fwd * 1e-3) / fp32_peak, 3),
                "bwd_frac_fp32_upper": round(evals * 40 * 2 / (bwd * 1e-3) / fp32_peak, 3),
            }), flush=True)
            del leaves, offsets, isect_ids, flatten_ids
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
