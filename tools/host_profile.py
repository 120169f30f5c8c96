"""Host-side (Python) cost of one rasterization() fwd+bwd call: cProfile over back-to-back iterations of cfg1 (the
GPU work of cfg1 is ~0.5 ms, so the loop is launch-bound and the profile is the wrapper's own overhead)."""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gsplat-triton_b200"), ROOT]
import torch

import bench
from gsplat.rendering import rasterization

cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg1"
host, names, kw = bench.make_scene(cfg, "cpu")
dev = torch.device("cuda")
params = [host[k].to(dev).requires_grad_(True) for k in names]
vm, Ks = host["viewmats"].to(dev), host["Ks"].to(dev)
G, gA = bench.loss_weights(dev, width=host["width"], height=host["height"])


def it():
    rc, ra, _ = rasterization(*params, vm, Ks, host["width"], host["height"], **kw)
    bench.weighted_loss(rc, ra, G, gA).backward()
    for p in params:
        p.grad = None


for _ in range(20):
    it()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200):
    it()
t_launch = time.perf_counter() - t0
torch.cuda.synchronize()
t_all = time.perf_counter() - t0
print(f"HOST {cfg}: {t_launch / 200 * 1e3:.3f} ms of host time per iteration, {t_all / 200 * 1e3:.3f} ms wall")
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    it()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
