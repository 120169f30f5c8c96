"""BASELINE.json configs[2] ("cfg3"): 3M Gaussians, C=8 cameras on a ring at 1297x840, packed=True,
rasterize_mode="antialiased", SH degree 3, fwd+bwd. Prints ms/step and the per-entry-point table (CUDA events).
Not the bench line (bench.py measures cfg2); a measured data point for the packed/batched path."""
import json
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gsplat-triton_b200"))
os.environ.setdefault("GSPLAT_BACKEND", "b200")
from gsplat.b200_impl._lib import timing  # noqa: E402
from gsplat.rendering import rasterization  # noqa: E402

N, C, W, H = int(os.environ.get("CFG3_N", 3_000_000)), 8, 1297, 840
dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(0)
means = (torch.rand(N, 3, generator=g) * 6 - 3).to(dev)
quats = torch.randn(N, 4, generator=g).to(dev)
lo, hi = math.log(0.002), math.log(0.02)
scales = torch.exp(torch.rand(N, 3, generator=g) * (hi - lo) + lo).to(dev)
opac = torch.rand(N, generator=g).to(dev)
shs = torch.randn(N, 16, 3, generator=g) * 0.3
shs[:, 0] += 0.5
shs = shs.to(dev)
f = W / (2 * math.tan(math.radians(30)))
vms = []
for i in range(C):
    az = 2 * math.pi * i / C
    eye = torch.tensor([6 * math.cos(az), 0.0, 6 * math.sin(az)])
    fwd = -eye / eye.norm()
    right = torch.linalg.cross(torch.tensor([0.0, 1.0, 0.0]), fwd)
    right = right / right.norm()
    up = torch.linalg.cross(fwd, right)
    R = torch.stack([right, up, fwd])
    vm = torch.eye(4)
    vm[:3, :3], vm[:3, 3] = R, -R @ eye
    vms.append(vm)
viewmats = torch.stack(vms).to(dev)
Ks = torch.tensor([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1.0]]).expand(C, 3, 3).contiguous().to(dev)
gw = torch.Generator().manual_seed(1)
G = torch.rand(C, H, W, 3, generator=gw).to(dev)
gA = torch.rand(C, H, W, 1, generator=gw).to(dev)
params = [t.requires_grad_(True) for t in (means, quats, scales, opac, shs)]
packed = os.environ.get("CFG3_PACKED", "1") == "1"


def step():
    for p in params:
        p.grad = None
    rc, ra, meta = rasterization(*params, viewmats, Ks, W, H, sh_degree=3, packed=packed,
                                 rasterize_mode="antialiased")
    ((rc * G).sum() + (ra * gA).sum()).backward()
    return meta


for _ in range(3):
    meta = step()
torch.cuda.synchronize()
steps = 10
evs = []
for _ in range(steps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    step()
    b.record()
    evs.append((a, b))
torch.cuda.synchronize()
ms = sum(a.elapsed_time(b) for a, b in evs) / steps
timing.reset(True)
for _ in range(5):
    step()
table = {k: round(v[1] / 5, 4) for k, v in timing.summary().items()}
timing.reset(False)
print(json.dumps({"workload": f"cfg3: {N} Gaussians, C=8 at 1297x840, packed={packed}, antialiased, SH deg 3",
                  "ms_per_step": round(ms, 3), "camera_renders_per_s": round(C * 1e3 / ms, 1),
                  "nnz": int(meta["radii"].numel()), "n_isects": int(meta["flatten_ids"].numel()),
                  "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 2), "stages_ms": table}))
