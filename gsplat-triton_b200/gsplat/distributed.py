"""Collectives used inside `rasterization(distributed=True)`: one process per GPU, NCCL over NVLink.

Same public helpers as the reference's gsplat/distributed.py (all_gather_int32 :10,
all_to_all_int32 :55, all_gather_tensor_list :102, all_to_all_tensor_list :170, cli :304) so
callers can switch, re-built on single-buffer collectives (`all_gather_into_tensor`,
`all_to_all_single`) instead of per-peer tensor lists, plus the fused row exchange the b200
`rasterization()` uses: ONE all-to-all-v per direction carrying a packed (8+D)-word row per
projected Gaussian (means2d, depth, conic, opacity, colour, radius) written by a CUDA pack
kernel, where the reference does cat/split copies and a second collective for the int32 radii
(rendering.py:413-494, distributed.py:229-257).

The helpers also run on CPU tensors with the `gloo` backend (used by the CPU tests of the
multi-rank bookkeeping); the pack/unpack kernels are CUDA-only.
"""
import os
from typing import Any, Callable, List, Optional, Sequence, Tuple, Union

import torch
import torch.distributed as dist
from torch import Tensor


def _as_int_list(values: Sequence[Union[int, Tensor]]) -> List[int]:
    return [int(v.item()) if isinstance(v, Tensor) else int(v) for v in values]


def all_gather_int32(world_size: int, value: Union[int, Tensor], device: Optional[torch.device] = None):
    """Gathers one 32-bit integer from every rank (not differentiable)."""
    if world_size == 1:
        return [value]
    scalar = isinstance(value, int)
    if scalar:
        assert device is not None, "device is required for scalar input"
        t = torch.tensor([value], dtype=torch.int32, device=device)
    else:
        t = value.reshape(1)
    out = torch.empty(world_size, dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, t)
    return out.tolist() if scalar else list(out.unbind())


def all_to_all_int32(world_size: int, values: List[Union[int, Tensor]], device: Optional[torch.device] = None):
    """Rank r sends values[i] to rank i and receives one integer from each (not differentiable)."""
    if world_size == 1:
        return values
    assert len(values) == world_size, "The length of values should be equal to world_size"
    scalar = [isinstance(v, int) for v in values]
    if any(scalar):
        assert device is not None, "device is required for scalar input"
    dev = device if device is not None else next(v.device for v in values if isinstance(v, Tensor))
    send = torch.stack([
        torch.tensor(v, dtype=torch.int32, device=dev) if isinstance(v, int) else v.to(torch.int32).reshape(())
        for v in values
    ])
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send)
    host = recv.tolist() if any(scalar) else None
    return [host[i] if scalar[i] else recv[i] for i in range(world_size)]


class _AllGatherRows(torch.autograd.Function):
    """all_gather along dim 0 with equal shapes; backward = reduce_scatter of the gradient."""

    @staticmethod
    def forward(ctx, data: Tensor, world_size: int):
        out = torch.empty((world_size * data.shape[0],) + tuple(data.shape[1:]), dtype=data.dtype,
                          device=data.device)
        dist.all_gather_into_tensor(out, data.contiguous())
        ctx.world_size = world_size
        return out

    @staticmethod
    def backward(ctx, grad: Tensor):
        grad = grad.contiguous()
        n = grad.shape[0] // ctx.world_size
        # sum over ranks of the slice that belongs to this rank (gloo has no reduce_scatter)
        if dist.get_backend() == "gloo":
            dist.all_reduce(grad)
            r = dist.get_rank()
            return grad[r * n:(r + 1) * n].clone(), None
        out = torch.empty((n,) + tuple(grad.shape[1:]), dtype=grad.dtype, device=grad.device)
        dist.reduce_scatter_tensor(out, grad)
        return out, None


def all_gather_tensor_list(world_size: int, tensor_list: List[Tensor]) -> List[Tensor]:
    """Gathers tensors with a common first dimension from all ranks (differentiable)."""
    if world_size == 1:
        return tensor_list
    n = len(tensor_list[0])
    for t in tensor_list:
        assert len(t) == n, "All tensors should have the same first dimension size"
    widths = [t.numel() // max(n, 1) for t in tensor_list]
    data = torch.cat([t.reshape(n, -1) for t in tensor_list], dim=-1)
    out = _AllGatherRows.apply(data, world_size)
    return [o.reshape(-1, *t.shape[1:]) for o, t in zip(out.split(widths, dim=-1), tensor_list)]


class _AllToAllRows(torch.autograd.Function):
    """all_to_all_single over dim 0 with per-peer row counts; backward = the reverse exchange."""

    @staticmethod
    def forward(ctx, data: Tensor, in_splits: Tuple[int, ...], out_splits: Tuple[int, ...]):
        data = data.contiguous()
        out = torch.empty((sum(out_splits),) + tuple(data.shape[1:]), dtype=data.dtype, device=data.device)
        dist.all_to_all_single(out, data, output_split_sizes=list(out_splits), input_split_sizes=list(in_splits))
        ctx.in_splits, ctx.out_splits = in_splits, out_splits
        return out

    @staticmethod
    def backward(ctx, grad: Tensor):
        grad = grad.contiguous()
        out = torch.empty((sum(ctx.in_splits),) + tuple(grad.shape[1:]), dtype=grad.dtype, device=grad.device)
        dist.all_to_all_single(out, grad, output_split_sizes=list(ctx.in_splits),
                               input_split_sizes=list(ctx.out_splits))
        return out, None, None


def all_to_all_rows(data: Tensor, in_splits: Sequence[int], out_splits: Sequence[int]) -> Tensor:
    """Differentiable all-to-all-v of the rows of `data` (one contiguous buffer per direction)."""
    return _AllToAllRows.apply(data, tuple(int(s) for s in in_splits), tuple(int(s) for s in out_splits))


def all_to_all_tensor_list(
    world_size: int,
    tensor_list: List[Tensor],
    splits: List[Union[int, Tensor]],
    output_splits: Optional[List[Union[int, Tensor]]] = None,
) -> List[Tensor]:
    """Splits tensors along dim 0 and exchanges the pieces between all ranks (differentiable)."""
    if world_size == 1:
        return tensor_list
    n = len(tensor_list[0])
    for t in tensor_list:
        assert len(t) == n, "All tensors should have the same first dimension size"
    assert len(splits) == world_size, "The length of splits should be equal to world_size"
    if output_splits is None:
        output_splits = all_to_all_int32(world_size, list(splits), device=tensor_list[0].device)
    in_s, out_s = _as_int_list(splits), _as_int_list(output_splits)
    widths = [t.numel() // max(n, 1) for t in tensor_list]
    data = torch.cat([t.reshape(n, -1) for t in tensor_list], dim=-1)
    out = all_to_all_rows(data, in_s, out_s)
    return [o.reshape(-1, *t.shape[1:]) for o, t in zip(out.split(widths, dim=-1), tensor_list)]


# ---------------------------------------------------------------------------------------------
# fused row exchange used by the b200 rasterization()
# ---------------------------------------------------------------------------------------------
class _PackRows(torch.autograd.Function):
    """(means2d, depths, conics, opacities, colors, radii) -> rows [P, 8+D]; CUDA kernel both ways."""

    @staticmethod
    def forward(ctx, means2d, depths, conics, opacities, colors, radii):
        from .b200_impl._lib import lib, ptr, require_cuda, stream_of

        require_cuda(means2d)
        P, D = depths.numel(), colors.shape[-1]
        rows = torch.empty((P, 8 + D), dtype=torch.float32, device=means2d.device)
        lib.gsb_a2a_pack(P, D, ptr(means2d.contiguous()), ptr(depths.contiguous()), ptr(conics.contiguous()),
                         ptr(opacities.contiguous()), ptr(colors.contiguous()), ptr(radii.contiguous()),
                         ptr(rows), stream_of(means2d))
        ctx.D = D
        return rows

    @staticmethod
    def backward(ctx, v_rows):
        return _unpack(v_rows.contiguous(), ctx.D, want_radii=False)[:5] + (None,)


def _unpack(rows: Tensor, D: int, want_radii: bool):
    from .b200_impl._lib import lib, ptr, stream_of

    P, dev = rows.shape[0], rows.device
    means2d = torch.empty((P, 2), dtype=torch.float32, device=dev)
    depths = torch.empty((P,), dtype=torch.float32, device=dev)
    conics = torch.empty((P, 3), dtype=torch.float32, device=dev)
    opacities = torch.empty((P,), dtype=torch.float32, device=dev)
    colors = torch.empty((P, D), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev) if want_radii else None
    lib.gsb_a2a_unpack(P, D, ptr(rows), ptr(means2d), ptr(depths), ptr(conics), ptr(opacities), ptr(colors),
                       ptr(radii), stream_of(rows))
    return means2d, depths, conics, opacities, colors, radii


class _UnpackRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rows, D):
        out = _unpack(rows.contiguous(), D, want_radii=True)
        ctx.mark_non_differentiable(out[5])
        ctx.D = D
        return out

    @staticmethod
    def backward(ctx, v_means2d, v_depths, v_conics, v_opacities, v_colors, v_radii):
        from .b200_impl._lib import lib, ptr, stream_of

        P, D = v_depths.numel(), ctx.D
        rows = torch.empty((P, 8 + D), dtype=torch.float32, device=v_depths.device)
        zeros = torch.zeros((max(P, 1),), dtype=torch.int32, device=v_depths.device)
        lib.gsb_a2a_pack(P, D, ptr(v_means2d.contiguous()), ptr(v_depths.contiguous()), ptr(v_conics.contiguous()),
                         ptr(v_opacities.contiguous()), ptr(v_colors.contiguous()), ptr(zeros), ptr(rows),
                         stream_of(rows))
        return rows, None


def pack_rows(means2d, depths, conics, opacities, colors, radii) -> Tensor:
    return _PackRows.apply(means2d, depths, conics, opacities, colors, radii)


def unpack_rows(rows: Tensor, D: int):
    return _UnpackRows.apply(rows, D)


def interleave_camera_major(rows: Tensor, C_local: int, N_world: Sequence[int]) -> Tensor:
    """[sum_j C_local*N_j, w] received peer-major (peer j, camera c, Gaussian n) -> [C_local, sum_j N_j, w].

    Same re-layout as the reference's `reshape_view` (rendering.py:260-267).
    """
    if C_local == 1:
        return rows.reshape(1, rows.shape[0], *rows.shape[1:])
    per_peer = rows.split([C_local * n for n in N_world], dim=0)
    return torch.cat([p.reshape(C_local, n, *rows.shape[1:]) for p, n in zip(per_peer, N_world)], dim=1)


# ---------------------------------------------------------------------------------------------
# process launcher (one process per GPU), cf. reference distributed.py:260-357
# ---------------------------------------------------------------------------------------------
def _find_free_port() -> int:
    import socket

    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(local_rank: int, fn: Callable, args: Any, world_size: int, port: int, backend: str):
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ["MASTER_PORT"] = str(port)
    if backend == "nccl":
        torch.cuda.set_device(local_rank)
    dist.init_process_group(backend=backend, rank=local_rank, world_size=world_size)
    try:
        fn(local_rank, local_rank, world_size, args)
    finally:
        dist.barrier()
        dist.destroy_process_group()


def cli(fn: Callable, args: Any, verbose: bool = False, world_size: Optional[int] = None,
        backend: Optional[str] = None) -> Optional[Any]:
    """Runs `fn(local_rank, world_rank, world_size, args)` on every visible GPU of this node."""
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if world_size is None:
        world_size = torch.cuda.device_count() if backend == "nccl" else 1
    if world_size <= 1:
        return fn(0, 0, 1, args)
    port = _find_free_port()
    torch.multiprocessing.spawn(_worker, args=(fn, args, world_size, port, backend), nprocs=world_size, join=True)
    return None
