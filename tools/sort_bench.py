"""A/B of the hand-written onesweep radix sort (csrc/radix_sort.cu) against cub::DeviceRadixSort on the two sorts
of a cfg2 step: 1 M (depth bits, row) pairs over 32 bits and 6.35 M (tile word, row) pairs over 13 bits."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gsplat-triton_b200"))
from gsplat.b200_impl import _wrapper as Wp  # noqa: E402

dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def time_sort(n, end_bit, impl, reps=30):
    os.environ["GSPLAT_B200_SORT_IMPL"] = impl
    if end_bit == 32:
        keys = (torch.rand(n, generator=g) * 8 + 2).view(torch.int32).to(dev)
    else:
        keys = torch.randint(0, 8160, (n,), generator=g, dtype=torch.int32).to(dev)
    vals = torch.arange(n, dtype=torch.int32, device=dev)
    ts = []
    for _ in range(reps + 3):
        k, v = keys.clone(), vals.clone()
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        Wp._sort_pairs_u32(k, v, end_bit)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts = sorted(ts[3:])
    return ts[len(ts) // 2]


for n, bits in ((1_000_000, 32), (6_347_458, 13), (27_000_000, 15)):
    own, cub = time_sort(n, bits, "onesweep"), time_sort(n, bits, "cub")
    print(f"n={n} bits={bits}: onesweep {own * 1e3:.1f} us, cub {cub * 1e3:.1f} us", flush=True)
