"""Pinned host<->device copy bandwidth per rank, alone and all ranks at once (the end-to-end leg
at 8 GPUs moved 11.6 GB/s per GPU). Run under torchrun, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/h2d_bandwidth.py

GSPLAT_BENCH_NUMA=0 leaves the processes unbound (torchrun's default); 1 (default) binds each to its GPU's NUMA node
before the pinned buffers are allocated. Prints one `H2D {...}` JSON line (rank 0)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from numa_bind import bind_to_gpu  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    bind = os.environ.get("GSPLAT_BENCH_NUMA", "1") == "1"
    info = bind_to_gpu(local) if bind else {"bound": False}
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = 256 << 20
    host = torch.empty(n, dtype=torch.uint8).pin_memory()
    host.fill_(1)  # first touch after binding
    devbuf = torch.empty(n, dtype=torch.uint8, device=dev)

    def measure(direction, reps=8):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(reps):
            if direction == "h2d":
                devbuf.copy_(host, non_blocking=True)
            else:
                host.copy_(devbuf, non_blocking=True)
        b.record()
        torch.cuda.synchronize()
        return n * reps / (a.elapsed_time(b) * 1e-3) / 1e9

    def barrier():
        if world > 1:
            dist.barrier()

    measure("h2d", 2)
    res = {}
    for direction in ("h2d", "d2h"):
        alone = torch.zeros(world, device=dev, dtype=torch.float64)
        for r in range(world):  # one rank at a time
            barrier()
            if r == rank:
                alone[r] = measure(direction)
        barrier()
        together = torch.zeros(world, device=dev, dtype=torch.float64)
        together[rank] = measure(direction)  # every rank at once
        if world > 1:
            dist.all_reduce(alone)
            dist.all_reduce(together)
        res[direction] = {"alone_GBps": [round(v, 1) for v in alone.tolist()],
                          "concurrent_GBps": [round(v, 1) for v in together.tolist()],
                          "concurrent_aggregate_GBps": round(float(together.sum()), 1)}
    infos = [None] * world
    if world > 1:
        dist.all_gather_object(infos, info)
    else:
        infos = [info]
    if rank == 0:
        print("H2D " + json.dumps({"world": world, "numa_binding": bind, "buffer_MB": n >> 20, "ranks": infos, **res}),
              flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
