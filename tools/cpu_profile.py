"""cProfile of the launching thread over resident cfg2 steps (where does the Python side spend its time)."""
import cProfile
import os
import pstats
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

from gsplat.rendering import rasterization  # noqa: E402

dev = torch.device("cuda", 0)
host = bench.make_cfg2("cpu")
params = [host[k].to(dev).requires_grad_(True) for k in ("means", "quats", "scales", "opacities", "shs")]
viewmats, Ks = host["viewmats"].to(dev), host["Ks"].to(dev)
G, gA = bench.loss_weights(dev)


def step():
    for p in params:
        p.grad = None
    rc, ra, meta = rasterization(*params, viewmats, Ks, bench.WIDTH, bench.HEIGHT, sh_degree=3, packed=False)
    ((rc * G).sum() + (ra * gA).sum()).backward()


for _ in range(5):
    step()
torch.cuda.synchronize()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
pr = cProfile.Profile()
pr.enable()
for _ in range(n):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(45)
st.sort_stats("tottime").print_stats(30)
