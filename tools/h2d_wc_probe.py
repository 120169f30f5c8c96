"""Host->device copy rate of the bench's 236 MB parameter set from (a) torch pinned memory, (b) write-combined pinned
memory (cudaHostAlloc, cudaHostAllocWriteCombined), (c) two copy streams. One GPU. Prints one JSON line."""
import ctypes
import json

import torch

N = 236_000_000 // 4
dev = torch.device("cuda", 0)
dst = torch.empty(N, dtype=torch.float32, device=dev)
src = torch.randn(N).pin_memory()


def rate(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return round(N * 4 * reps / (a.elapsed_time(b) * 1e-3) / 1e9, 2)


out = {"pinned_GBps": rate(lambda: dst.copy_(src, non_blocking=True))}
rt = ctypes.CDLL("libcudart.so.12")
p = ctypes.c_void_p()
rc = rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(N * 4), ctypes.c_uint(0x04))
out["cudaHostAlloc_wc_rc"] = rc
if rc == 0:
    buf = (ctypes.c_float * N).from_address(p.value)
    wc = torch.frombuffer(buf, dtype=torch.float32)
    wc.copy_(src)
    out["wc_is_pinned"] = wc.is_pinned()
    out["write_combined_GBps"] = rate(lambda: dst.copy_(wc, non_blocking=True))
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
h = N // 2


def two():
    with torch.cuda.stream(s1):
        dst[:h].copy_(src[:h], non_blocking=True)
    with torch.cuda.stream(s2):
        dst[h:].copy_(src[h:], non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)


out["two_streams_GBps"] = rate(two)
print("H2D_PROBE " + json.dumps(out))
