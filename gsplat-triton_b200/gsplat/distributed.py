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


class _HostTrace:
    """Optional CPU-side phase timer (GSPLAT_B200_TRACE=1): where the launching thread spends its time."""

    def __init__(self):
        self.enabled = os.environ.get("GSPLAT_B200_TRACE", "0") == "1"
        self.acc, self.t = {}, None

    def mark(self, name=None):
        if not self.enabled:
            return
        import time

        now = time.perf_counter()
        if name is not None and self.t is not None:
            c, tot = self.acc.get(name, (0, 0.0))
            self.acc[name] = (c + 1, tot + now - self.t)
        self.t = now

    def summary(self):
        return {k: round(1e6 * tot / max(c, 1), 1) for k, (c, tot) in self.acc.items()}


host_trace = _HostTrace()


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


class HostBoard:
    """Host-side all-gather of one integer per rank through a shared-memory file per rank (one node).

    `all_gather_int32` on a CUDA tensor costs a device-to-host read-back: the CPU waits for every queued kernel
    of the previous step and the next forward's launches are exposed. The count of Gaussians is host data
    already, so the ranks post it on a board in /dev/shm instead: microseconds, no GPU involvement, the CPU
    keeps running ahead of the GPU. Slot layout per rank: int64[2][2 + 16] = (sequence, check, values...) for even /
    odd sequence numbers (a rank can be at most one gather ahead of a peer that has not read it yet); up to 16
    values per gather (the packed exchange posts one row count per destination rank).
    """

    _MAGIC = 0x5DEECE66D
    _WIDTH = 16

    def __init__(self, group=None):
        import mmap
        import tempfile
        import uuid

        import numpy as np

        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        root = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
        name = [f"gsb_board_{os.getpid()}_{uuid.uuid4().hex[:12]}"]
        dist.broadcast_object_list(name, src=0, group=group)
        self.paths = [os.path.join(root, f"{name[0]}_{r}") for r in range(self.world)]
        nbytes = 2 * (2 + self._WIDTH) * 8
        with open(self.paths[self.rank], "wb") as f:
            f.write(b"\0" * nbytes)
        dist.barrier(group)  # every rank's file exists
        self.maps, self.slots = [], []
        for path in self.paths:
            f = open(path, "r+b")
            m = mmap.mmap(f.fileno(), nbytes)
            f.close()
            self.maps.append(m)
            self.slots.append(np.frombuffer(m, dtype=np.int64, count=2 * (2 + self._WIDTH)).reshape(2, 2 + self._WIDTH))
        dist.barrier(group)  # every rank has mapped every file: the names can go
        try:
            os.unlink(self.paths[self.rank])
        except OSError:
            pass
        self.seq = 0

    def all_gather(self, value: int, timeout_s: float = 300.0) -> List[int]:
        return [v[0] for v in self.all_gather_vec([int(value)], timeout_s)]

    def all_gather_vec(self, values: Sequence[int], timeout_s: float = 300.0) -> List[List[int]]:
        """Every rank posts `values` (same length on every rank, <= 16); returns the rows of all ranks."""
        import time

        n = len(values)
        assert 1 <= n <= self._WIDTH
        self.seq += 1
        seq, p = self.seq, self.seq & 1
        mask = 0x7FFFFFFFFFFFFFFF

        def check_of(vals):
            c = seq * self._MAGIC
            for k, v in enumerate(vals):
                c ^= (int(v) + 1) * (2 * k + 3)
            return c & mask

        mine = self.slots[self.rank]
        mine[p, 2:2 + n] = [int(v) for v in values]
        mine[p, 1] = check_of(values)
        mine[p, 0] = seq
        out = []
        for r in range(self.world):
            slot = self.slots[r]
            spins, t0 = 0, None
            while True:
                s_, c_ = int(slot[p, 0]), int(slot[p, 1])
                vals = [int(v) for v in slot[p, 2:2 + n]]
                if s_ == seq and c_ == check_of(vals):
                    out.append(vals)
                    break
                spins += 1
                if spins > 2000:
                    if t0 is None:
                        t0 = time.monotonic()
                    elif time.monotonic() - t0 > timeout_s:
                        raise RuntimeError(f"gsplat b200 host board: rank {r} did not post gather {seq}")
                    time.sleep(20e-6)
        return out


_HOST_BOARD = None


def _group_key():
    """Identity of the current default process group (boards / mailboxes do not outlive it)."""
    return (id(dist.distributed_c10d._get_default_group()), dist.get_world_size(), dist.get_rank())


def all_gather_int_host(world_size: int, value: int) -> List[int]:
    """Gathers one host integer from every rank of this node without touching the GPU (see HostBoard).
    Collective: every rank must call it."""
    global _HOST_BOARD
    if world_size == 1:
        return [int(value)]
    key = _group_key()
    if _HOST_BOARD is None or _HOST_BOARD[0] != key:  # a new process group (init/destroy cycle): new board
        _HOST_BOARD = (key, HostBoard())
    return _HOST_BOARD[1].all_gather(int(value))


def all_gather_ints_host(world_size: int, values: Sequence[int]) -> List[List[int]]:
    """Every rank posts up to 16 host integers; returns all ranks' rows (row r = what rank r posted) without touching
    the GPU (see HostBoard). Collective: every rank must call it with the same number of values."""
    global _HOST_BOARD
    if world_size == 1:
        return [[int(v) for v in values]]
    key = _group_key()
    if _HOST_BOARD is None or _HOST_BOARD[0] != key:
        _HOST_BOARD = (key, HostBoard())
    return _HOST_BOARD[1].all_gather_vec(values)


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
        # contiguous copies stay bound to locals until the launch is queued (never `ptr(x.contiguous())`: the
        # temporary would be freed, and possibly reused by the next copy, before the kernel reads it)
        means2d, depths, conics, opacities, colors, radii = (
            t.contiguous() for t in (means2d, depths, conics, opacities, colors, radii))
        lib.gsb_a2a_pack(P, D, ptr(means2d), ptr(depths), ptr(conics), ptr(opacities), ptr(colors), ptr(radii),
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
        # e.g. RGB+ED hands v_colors / v_depths over as narrow views of one cat() gradient: two copies back to back
        v_means2d, v_depths, v_conics, v_opacities, v_colors = (
            t.float().contiguous() for t in (v_means2d, v_depths, v_conics, v_opacities, v_colors))
        lib.gsb_a2a_pack(P, D, ptr(v_means2d), ptr(v_depths), ptr(v_conics), ptr(v_opacities), ptr(v_colors),
                         ptr(zeros), ptr(rows), stream_of(rows))
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
# peer-memory exchange (NVLink): fused pack + put kernels of csrc/peer.cu instead of NCCL all-to-all
# ---------------------------------------------------------------------------------------------
class PeerMailbox:
    """Per-process mailbox that the other ranks of the node write into over NVLink (CUDA IPC).

    One instance per process group. `ensure(bytes)` (collective: every rank must call it with the
    same argument) grows the four regions; handles are exchanged with one all_gather. The device
    protocol (epochs, flags, double-buffered regions) is described in csrc/peer.cu.
    """

    def __init__(self, group=None):
        import ctypes

        from .b200_impl._lib import lib

        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        assert self.world <= 16, "peer exchange supports up to 16 ranks (GSB_MAX_PEERS)"
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.header = int(lib.gsb_peer_header_bytes())
        self.capacity = 0  # bytes per region
        self.local = None  # own mailbox (device pointer, int)
        self.peers = [None] * self.world  # mapped peer mailboxes (own pointer at [rank])
        self.peer_array = (ctypes.c_void_p * self.world)()
        self.epoch = [0, 0, 0, 0]  # per channel: forward, backward, forward colours, backward colours
        self.side_stream = torch.cuda.Stream(device=self.device)  # the split exchange's puts run here
        flag = ctypes.c_void_p()
        _check(lib.gsb_peer_flag_alloc(ctypes.byref(flag)), "gsb_peer_flag_alloc")
        self.timeout_flag = flag.value
        self._flag_view = ctypes.cast(flag.value, ctypes.POINTER(ctypes.c_int))

    def ensure(self, region_bytes: int) -> None:
        if region_bytes <= self.capacity:
            return
        import ctypes

        from .b200_impl._lib import lib

        cap = max(int(region_bytes * 1.25), 1 << 20)
        cap = (cap + 1023) // 1024 * 1024
        # everything in flight (our puts into peers, their puts into us) must have landed before remapping
        torch.cuda.synchronize(self.device)
        dist.barrier(self.group)
        self._release()
        local = ctypes.c_void_p()
        _check(lib.gsb_peer_alloc(self.header + 4 * cap, ctypes.byref(local)), "gsb_peer_alloc")
        self.local = local.value
        hb = int(lib.gsb_peer_handle_bytes())
        handle = (ctypes.c_ubyte * hb)()
        _check(lib.gsb_peer_get_handle(self.local, handle), "gsb_peer_get_handle")
        mine = torch.tensor(list(handle), dtype=torch.uint8, device=self.device)
        everyone = torch.empty((self.world, hb), dtype=torch.uint8, device=self.device)
        dist.all_gather_into_tensor(everyone, mine, group=self.group)
        everyone = everyone.cpu()
        for r in range(self.world):
            if r == self.rank:
                self.peers[r] = self.local
            else:
                raw = (ctypes.c_ubyte * hb)(*everyone[r].tolist())
                mapped = ctypes.c_void_p()
                _check(lib.gsb_peer_open_handle(raw, ctypes.byref(mapped)), "gsb_peer_open_handle")
                self.peers[r] = mapped.value
            self.peer_array[r] = self.peers[r]
        self.capacity = cap
        self.epoch = [0, 0, 0, 0]  # fresh, zeroed flags on every rank
        dist.barrier(self.group)  # nobody writes before every rank has mapped every mailbox

    def region(self, direction: int, epoch: int) -> int:
        return self.header + (2 * (direction & 1) + (epoch & 1)) * self.capacity

    def next_epoch(self, channel: int) -> int:
        self.check()
        self.epoch[channel] += 1
        return self.epoch[channel]

    def on_side_stream(self, launch, tensors) -> "torch.cuda.Event":
        """Runs `launch(cuda_stream)` on the side stream, ordered after everything queued so far on the current
        stream; `tensors` are the buffers it reads (kept alive for the side stream). Returns the completion event."""
        main = torch.cuda.current_stream(self.device)
        ready = torch.cuda.Event()
        ready.record(main)
        self.side_stream.wait_event(ready)
        for t in tensors:
            if t is not None:
                t.record_stream(self.side_stream)
        with torch.cuda.stream(self.side_stream):
            launch(self.side_stream.cuda_stream)
        done = torch.cuda.Event()
        done.record(self.side_stream)
        return done

    def check(self) -> None:
        """Raises if an earlier wait kernel gave up on a peer (it writes 1 + rank into host-mapped memory)."""
        v = self._flag_view[0]
        if v != 0:
            raise RuntimeError(f"gsplat b200 peer exchange: rank {v - 1} never published its rows (timeout)")

    def _release(self):
        from .b200_impl._lib import lib

        for r, p in enumerate(self.peers):
            if p is not None and r != self.rank:
                lib.gsb_peer_close_handle(p)
            self.peers[r] = None
        if self.local is not None:
            lib.gsb_peer_free(self.local)
            self.local = None
        self.capacity = 0

    def close(self):
        from .b200_impl._lib import lib

        torch.cuda.synchronize(self.device)
        self._release()
        if self.timeout_flag is not None:
            lib.gsb_peer_flag_free(self.timeout_flag)
            self.timeout_flag = None


def _check(rc: int, what: str) -> None:
    if rc != 0:
        from .b200_impl._lib import lib

        raise RuntimeError(f"{what} failed ({rc}): {lib.gsb_last_error().decode()}")


_MAILBOXES = {}


_PEER_OK = {}


def peer_exchange_available() -> bool:
    """True when the row exchange can go through peer memory: NCCL process group, GSPLAT_B200_EXCHANGE != "nccl",
    and EVERY rank agrees that all ranks sit on one node with peer access between their GPUs (decided once per
    process group by a collective: every rank must reach the first call, which `rasterization(distributed=True)`
    guarantees). Anything else selects the NCCL all-to-all."""
    if os.environ.get("GSPLAT_B200_EXCHANGE", "peer") == "nccl":
        return False
    if not (dist.is_available() and dist.is_initialized() and torch.cuda.is_available()):
        return False
    if dist.get_backend() != "nccl":
        return False
    world = dist.get_world_size()
    if not 1 < world <= 16:
        return False
    key = _group_key()
    if key not in _PEER_OK:
        import socket

        try:
            boot = open("/proc/sys/kernel/random/boot_id").read().strip()
        except OSError:
            boot = ""
        mine = (socket.gethostname(), boot, torch.cuda.current_device(), torch.cuda.device_count())
        everyone = [None] * world
        dist.all_gather_object(everyone, mine)
        same_node = all(e[:2] == mine[:2] for e in everyone)
        devices = [e[2] for e in everyone]
        ok = same_node and len(set(devices)) == world and world <= mine[3]
        if ok:
            me = torch.cuda.current_device()
            ok = all(d == me or torch.cuda.can_device_access_peer(me, d) for d in devices)
        verdict = torch.tensor([1 if ok else 0], device=torch.device("cuda", torch.cuda.current_device()))
        dist.all_reduce(verdict, op=dist.ReduceOp.MIN)  # unanimous, or nobody uses the mailboxes
        _PEER_OK.clear()
        _PEER_OK[key] = bool(verdict.item())
    return _PEER_OK[key]


def get_mailbox(group=None) -> PeerMailbox:
    key = (id(group),) if group is not None else _group_key()
    box = _MAILBOXES.get(key)
    if box is None:
        close_mailboxes()  # mailboxes of a process group that no longer exists
        box = _MAILBOXES[key] = PeerMailbox(group)
    return box


def check_mailboxes() -> None:
    """Raises if a wait kernel of an exchange already executed gave up on a peer (`PeerMailbox.check`). Called by
    `rasterization(distributed=True)` right after its one host synchronisation of the step, so rows that never
    arrived are reported in the step that would have composited them, not in the next one."""
    for box in _MAILBOXES.values():
        box.check()


def close_mailboxes() -> None:
    for box in _MAILBOXES.values():
        box.close()
    _MAILBOXES.clear()


def _region_views(rows: int, D: int, with_radii: bool, device):
    """One flat float32 buffer laid out like a mailbox region (csrc/peer.cu) + its per-field views."""
    from .b200_impl._lib import lib

    n_fields = 6 if with_radii else 5
    offs = [int(lib.gsb_a2a_field_offset(rows, D, f)) for f in range(n_fields + 1)]
    widths = (2, 1, 3, 1, D, 1)
    out = torch.empty((offs[-1],), dtype=torch.float32, device=device)
    return out, [out[offs[f]:offs[f] + rows * widths[f]] for f in range(n_fields)]


class _PeerExchangeRows(torch.autograd.Function):
    """Projected rows [C, N_local, .] of all C = world*C_local cameras -> [C_local, N_total, .] of the local
    cameras, through the peers' mailboxes; backward sends the gradients back the same way."""

    @staticmethod
    def forward(ctx, means2d, depths, conics, opacities, colors, radii, C_local, N_world):
        from .b200_impl._lib import lib, ptr, stream_of

        box = get_mailbox()
        world, rank = box.world, box.rank
        C, N_local = depths.shape
        D = colors.shape[-1]
        assert C == world * C_local and N_world[rank] == N_local
        N_total = sum(N_world)
        bases = [0]
        for n in N_world:
            bases.append(bases[-1] + n)
        box.ensure(4 * max(int(lib.gsb_a2a_field_offset(C_local * N_total, D, 6)),
                           int(lib.gsb_a2a_field_offset(C * max(N_world), D, 5))))
        dev = means2d.device
        st = stream_of(means2d)
        epoch = box.next_epoch(0)
        region = box.region(0, epoch)
        means2d, depths, conics, opacities, colors, radii = (
            t.contiguous() for t in (means2d, depths, conics, opacities, colors, radii))  # bound past the launch
        lib.gsb_a2a_put_fwd(world, rank, C_local, N_local, N_total, bases[rank], D, ptr(means2d), ptr(depths),
                            ptr(conics), ptr(opacities), ptr(colors), ptr(radii), box.peer_array, region, epoch, st)
        rows = C_local * N_total
        out, views = _region_views(rows, D, True, dev)
        lib.gsb_a2a_wait_copy(world, 0, epoch, box.local, region, out.numel(), ptr(out), box.timeout_flag, st)
        o_means2d, o_depths, o_conics, o_opac, o_colors, o_radii = (
            views[0].view(C_local, N_total, 2), views[1].view(C_local, N_total), views[2].view(C_local, N_total, 3),
            views[3].view(C_local, N_total), views[4].view(C_local, N_total, D),
            views[5].view(torch.int32).view(C_local, N_total))
        ctx.geom = (C_local, N_local, N_total, D, bases)
        ctx.mark_non_differentiable(o_radii)
        return o_means2d, o_depths, o_conics, o_opac, o_colors, o_radii

    @staticmethod
    def backward(ctx, v_means2d, v_depths, v_conics, v_opac, v_colors, v_radii):
        import ctypes

        from .b200_impl._lib import lib, ptr, stream_of

        C_local, N_local, N_total, D, bases = ctx.geom
        box = get_mailbox()
        world, rank = box.world, box.rank
        C = world * C_local
        dev = box.device
        st = torch.cuda.current_stream(dev).cuda_stream

        def f32(t):
            return None if t is None else t.float().contiguous()

        epoch = box.next_epoch(1)
        region = box.region(1, epoch)
        grads = [f32(t) for t in (v_means2d, v_depths, v_conics, v_opac, v_colors)]
        bases_arr = (ctypes.c_int64 * (world + 1))(*bases)
        lib.gsb_a2a_put_bwd(world, rank, C_local, N_total, bases_arr, D, *(ptr(g) for g in grads), box.peer_array,
                            region, epoch, st)
        out, views = _region_views(C * N_local, D, False, dev)
        lib.gsb_a2a_wait_copy(world, 1, epoch, box.local, region, out.numel(), ptr(out), box.timeout_flag, st)
        g_means2d, g_depths, g_conics, g_opac, g_colors = (
            views[0].view(C, N_local, 2), views[1].view(C, N_local), views[2].view(C, N_local, 3),
            views[3].view(C, N_local), views[4].view(C, N_local, D))
        return g_means2d, g_depths, g_conics, g_opac, g_colors, None, None, None


class SplitExchange:
    """The unpacked peer exchange of one `rasterization()` call, its forward half split in two so that the transfers
    hide under compute:

      start_geometry()   put of means2d / depths / conics / opacities / radii on the side stream, right after the
                         projection: it travels while the SH colour kernel runs;
      start_colors()     put of the colours on the side stream, right after the colour kernel: it travels while the
                         receiver sorts its intersections (which need the geometry only);
      finish()           waits for every peer's geometry + copies it out (main stream) and returns ALL exchanged
                         tensors; the colour tensor's contents are valid only after
      complete_colors()  waits for every peer's colours + copies them out; called right before the compositor (stream
                         order makes the data visible to everything queued after it).
    Geometry and colours use separate epoch flags (channels 0 and 2) and disjoint fields of the same mailbox region.
    The backward exchange is one put + one wait for all gradient fields (measured at 2 GPUs: splitting it as well
    costs more in launches and stream hand-overs than the overlap returns)."""

    GEOMETRY_FWD = 0b101111  # means2d, depths, conics, opacities, radii
    COLORS = 0b010000

    def __init__(self, C_local: int, N_world: Sequence[int], D: int):
        from .b200_impl._lib import lib

        self.box = get_mailbox()
        self.C_local, self.N_world, self.D = int(C_local), tuple(int(n) for n in N_world), int(D)
        world, rank = self.box.world, self.box.rank
        self.N_local, self.N_total = self.N_world[rank], sum(self.N_world)
        self.bases = [0]
        for n in self.N_world:
            self.bases.append(self.bases[-1] + n)
        C = world * self.C_local
        self.rows_fwd, self.rows_bwd = self.C_local * self.N_total, C * self.N_local
        self.box.ensure(4 * max(int(lib.gsb_a2a_field_offset(self.rows_fwd, D, 6)),
                                int(lib.gsb_a2a_field_offset(C * max(self.N_world), D, 5))))
        self.off_fwd = [int(lib.gsb_a2a_field_offset(self.rows_fwd, D, f)) for f in range(7)]
        self.off_bwd = [int(lib.gsb_a2a_field_offset(self.rows_bwd, D, f)) for f in range(6)]
        self.epochs = {0: self.box.next_epoch(0), 2: self.box.next_epoch(2)}  # this call's epoch on each channel
        self.region = self.box.region(0, self.epochs[0])
        self.out = torch.empty((self.off_fwd[6],), dtype=torch.float32, device=self.box.device)
        self.done = {}
        self.colors_pending = False

    def _put_fwd(self, mask, channel, means2d, depths, conics, opacities, colors, radii):
        from .b200_impl._lib import lib, ptr

        box = self.box
        tensors = [None if t is None else t.contiguous() for t in (means2d, depths, conics, opacities, colors, radii)]
        epoch = self.epochs[channel]

        def launch(stream):
            lib.gsb_a2a_put_fwd_part(box.world, box.rank, self.C_local, self.N_local, self.N_total,
                                     self.bases[box.rank], self.D, *(ptr(t) for t in tensors), mask, channel,
                                     box.peer_array, self.region, epoch, stream)

        self.done[channel] = box.on_side_stream(launch, tensors)

    def start_geometry(self, means2d, depths, conics, opacities, radii):
        assert depths.shape == (self.box.world * self.C_local, self.N_local), depths.shape
        with torch.no_grad():
            self._put_fwd(self.GEOMETRY_FWD, 0, means2d, depths, conics, opacities, None, radii)

    def start_colors(self, colors):
        with torch.no_grad():
            self._put_fwd(self.COLORS, 2, None, None, None, None, colors, None)
        self.colors_pending = True

    def _wait_fwd(self, channel, ranges):
        from .b200_impl._lib import lib, ptr, stream_of

        box = self.box
        torch.cuda.current_stream(box.device).wait_event(self.done[channel])  # own put first: never spin on ourselves
        (b0, e0), (b1, e1) = ranges
        lib.gsb_a2a_wait_copy_part(box.world, channel, self.epochs[channel], box.local, self.region, b0, e0, b1, e1,
                                   ptr(self.out), box.timeout_flag, stream_of(self.out))

    def finish(self, means2d, depths, conics, opacities, colors, radii):
        """-> (means2d, depths, conics, opacities, colors, radii) of the local cameras, [C_local, N_total, .];
        `colors` becomes valid with complete_colors()."""
        return _SplitFinish.apply(means2d, depths, conics, opacities, colors, radii, self)

    def complete_colors(self):
        if self.colors_pending:
            o = self.off_fwd
            self._wait_fwd(2, ((o[4], o[5]), (0, 0)))
            self.colors_pending = False


class _SplitFinish(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means2d, depths, conics, opacities, colors, radii, ex):
        o = ex.off_fwd
        ex._wait_fwd(0, ((o[0], o[4]), (o[5], o[6])))
        C_l, N_t, out, rows, D = ex.C_local, ex.N_total, ex.out, ex.rows_fwd, ex.D
        outs = (out[o[0]:o[0] + 2 * rows].view(C_l, N_t, 2), out[o[1]:o[1] + rows].view(C_l, N_t),
                out[o[2]:o[2] + 3 * rows].view(C_l, N_t, 3), out[o[3]:o[3] + rows].view(C_l, N_t),
                out[o[4]:o[4] + D * rows].view(C_l, N_t, D),
                out[o[5]:o[5] + rows].view(torch.int32).view(C_l, N_t))
        ctx.ex = ex
        ctx.mark_non_differentiable(outs[5])
        return outs

    @staticmethod
    def backward(ctx, v_means2d, v_depths, v_conics, v_opac, v_colors, _v_radii):
        import ctypes

        from .b200_impl._lib import lib, ptr, stream_of

        ex = ctx.ex
        box = ex.box
        world, rank = box.world, box.rank
        o = ex.off_bwd
        epoch = box.next_epoch(1)
        region = box.region(1, epoch)
        out = torch.empty((o[5],), dtype=torch.float32, device=box.device)
        # depths carry no gradient unless a depth channel is rendered: the field is then neither sent nor copied
        # (every rank runs the same graph, so every rank skips it)
        has_depth = v_depths is not None
        mask = 0b11101 | (0b00010 if has_depth else 0)
        ranges = ((o[0], o[5]), (0, 0)) if has_depth else ((o[0], o[1]), (o[2], o[5]))
        grads = [None if g is None else g.float().contiguous() for g in (v_means2d, v_depths, v_conics, v_opac, v_colors)]
        bases_arr = (ctypes.c_int64 * (world + 1))(*ex.bases)
        st = stream_of(out)
        lib.gsb_a2a_put_bwd_part(world, rank, ex.C_local, ex.N_total, bases_arr, ex.D, *(ptr(g) for g in grads), mask, 1,
                                 box.peer_array, region, epoch, st)
        (b0, e0), (b1, e1) = ranges
        lib.gsb_a2a_wait_copy_part(world, 1, epoch, box.local, region, b0, e0, b1, e1, ptr(out), box.timeout_flag, st)
        C, N_l, rows, D = world * ex.C_local, ex.N_local, ex.rows_bwd, ex.D

        def view(f, w):
            v = out[o[f]:o[f] + rows * w]
            return v.view(C, N_l, w) if w > 1 else v.view(C, N_l)

        return view(0, 2), (view(1, 1) if has_depth else None), view(2, 3), view(3, 1), view(4, D), None, None


class PushExchange(SplitExchange):
    """The unpacked peer exchange without a put kernel in the forward direction: `fully_fused_projection` and the SH
    colour kernel store every row they produce ALSO into the mailbox of the rank that owns its camera
    (`gsb_projection_fwd_push`, `gsb_sh_colors_fwd_push`: NVLink peer stores from the producing kernels), a one-warp
    kernel raises the epoch flags, and `finish()` waits + copies out as before. The backward exchange is the usual
    put + wait. `rasterization()` installs the object as the wrappers' push context around the two producer calls."""

    def __init__(self, C_local: int, N_world: Sequence[int], D: int, opacities: Tensor):
        super().__init__(C_local, N_world, D)
        self.opacities = opacities.detach().float().contiguous()
        self.pushed = set()

    def push_args(self):
        """(world, C_local, N_total, n_base, peer table, region offset) for the *_push entry points."""
        box = self.box
        return box.world, self.C_local, self.N_total, self.bases[box.rank], box.peer_array, self.region

    def finish(self, means2d, depths, conics, opacities, colors, radii):
        from .b200_impl._lib import lib, stream_of

        assert self.pushed == {"geometry", "colors"}, f"push exchange: producers did not push ({self.pushed})"
        box = self.box
        lib.gsb_a2a_publish(box.world, box.rank, 0, self.epochs[0], box.peer_array, stream_of(means2d))
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(box.device))
        self.done[0] = ev
        return _PushFinish.apply(means2d, depths, conics, opacities, colors, radii, self)


class _PushFinish(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means2d, depths, conics, opacities, colors, radii, ex):
        o = ex.off_fwd
        ex._wait_fwd(0, ((o[0], o[6]), (0, 0)))
        C_l, N_t, out, rows, D = ex.C_local, ex.N_total, ex.out, ex.rows_fwd, ex.D
        outs = (out[o[0]:o[0] + 2 * rows].view(C_l, N_t, 2), out[o[1]:o[1] + rows].view(C_l, N_t),
                out[o[2]:o[2] + 3 * rows].view(C_l, N_t, 3), out[o[3]:o[3] + rows].view(C_l, N_t),
                out[o[4]:o[4] + D * rows].view(C_l, N_t, D),
                out[o[5]:o[5] + rows].view(torch.int32).view(C_l, N_t))
        ctx.ex = ex
        ctx.mark_non_differentiable(outs[5])
        return outs

    backward = _SplitFinish.backward


def peer_exchange_rows(means2d, depths, conics, opacities, colors, radii, C_local: int, N_world: Sequence[int]):
    """Fused exchange of the projected rows (unpacked mode): see `_PeerExchangeRows`."""
    return _PeerExchangeRows.apply(means2d, depths, conics, opacities, colors, radii, int(C_local),
                                   tuple(int(n) for n in N_world))


def _region_views_packed(rows: int, D: int, n_fields: int, device):
    """Flat float32 buffer laid out like a packed mailbox region (csrc/peer.cu) + per-field views."""
    from .b200_impl._lib import lib

    offs = [int(lib.gsb_a2a_field_offset(rows, D, f)) for f in range(n_fields + 1)]
    widths = (2, 1, 3, 1, D, 1, 1, 1)
    out = torch.empty((max(offs[-1], 4),), dtype=torch.float32, device=device)
    return out, [out[offs[f]:offs[f] + rows * widths[f]] for f in range(n_fields)]


class _PeerExchangePacked(torch.autograd.Function):
    """Packed (COO, camera-major) rows -> the ranks owning their cameras, through the peers' mailboxes.

    `matrix[r][d]` = rows rank r sends to rank d (host data, identical on every rank: HostBoard). Returns the rows of
    the local cameras grouped by sender, with camera ids local to this rank and Gaussian ids global, exactly what the
    reference builds with two all_to_all calls (rendering.py:413-451); backward returns the gradients the same way."""

    @staticmethod
    def forward(ctx, means2d, depths, conics, opacities, colors, radii, camera_ids, gaussian_ids, matrix, C_local,
                gauss_base):
        import ctypes

        from .b200_impl._lib import lib, ptr, stream_of

        box = get_mailbox()
        world, rank = box.world, box.rank
        D = colors.shape[-1]
        send = [int(v) for v in matrix[rank]]
        assert sum(send) == depths.numel()
        rows_in = [sum(matrix[r][d] for r in range(world)) for d in range(world)]  # table size of every rank
        nnz = [sum(matrix[r]) for r in range(world)]
        box.ensure(4 * max(int(lib.gsb_a2a_field_offset(max(max(rows_in), max(nnz), 1), D, 8)), 4))
        dev, st = means2d.device, stream_of(means2d)
        arr = lambda v: (ctypes.c_int64 * world)(*[int(x) for x in v])  # noqa: E731
        dst_base = [sum(matrix[r][d] for r in range(rank)) for d in range(world)]
        cam_base = [d * C_local for d in range(world)]
        epoch = box.next_epoch(0)
        region = box.region(0, epoch)
        means2d, depths, conics, opacities, colors, radii, camera_ids, gaussian_ids = (
            t.contiguous() for t in (means2d, depths, conics, opacities, colors, radii, camera_ids, gaussian_ids))
        lib.gsb_a2a_put_packed_fwd(world, rank, D, arr(send), arr(dst_base), arr(rows_in), arr(cam_base),
                                   int(gauss_base), ptr(means2d), ptr(depths), ptr(conics), ptr(opacities), ptr(colors),
                                   ptr(radii), ptr(camera_ids), ptr(gaussian_ids), box.peer_array, region, epoch, st)
        rows = rows_in[rank]
        out, v = _region_views_packed(rows, D, 8, dev)
        lib.gsb_a2a_wait_copy(world, 0, epoch, box.local, region, out.numel() // 4 * 4, ptr(out), box.timeout_flag, st)
        outs = (v[0].view(rows, 2), v[1], v[2].view(rows, 3), v[3], v[4].view(rows, D), v[5].view(torch.int32),
                v[6].view(torch.int32).long(), v[7].view(torch.int32).long())
        ctx.plan = (matrix, D, nnz)
        ctx.mark_non_differentiable(outs[5], outs[6], outs[7])
        return outs

    @staticmethod
    def backward(ctx, v_means2d, v_depths, v_conics, v_opac, v_colors, _v_radii, _v_cam, _v_gauss):
        import ctypes

        from .b200_impl._lib import lib, ptr

        matrix, D, nnz = ctx.plan
        box = get_mailbox()
        world, rank = box.world, box.rank
        dev = box.device
        st = torch.cuda.current_stream(dev).cuda_stream
        arr = lambda v: (ctypes.c_int64 * world)(*[int(x) for x in v])  # noqa: E731
        recv = [matrix[r][rank] for r in range(world)]  # rows held for sender r, in sender order
        dst_base = [sum(matrix[r][:rank]) for r in range(world)]  # where they sit in r's packed order
        epoch = box.next_epoch(1)
        region = box.region(1, epoch)
        grads = [None if t is None else t.float().contiguous() for t in (v_means2d, v_depths, v_conics, v_opac, v_colors)]
        lib.gsb_a2a_put_packed_bwd(world, rank, D, arr(recv), arr(dst_base), arr(nnz), *(ptr(g) for g in grads),
                                   box.peer_array, region, epoch, st)
        rows = nnz[rank]
        out, v = _region_views_packed(rows, D, 5, dev)
        lib.gsb_a2a_wait_copy(world, 1, epoch, box.local, region, out.numel() // 4 * 4, ptr(out), box.timeout_flag, st)
        return (v[0].view(rows, 2), v[1], v[2].view(rows, 3), v[3], v[4].view(rows, D)) + (None,) * 6


def peer_exchange_packed(means2d, depths, conics, opacities, colors, radii, camera_ids, gaussian_ids, matrix,
                         C_local: int, gauss_base: int):
    """Fused exchange of packed rows: see `_PeerExchangePacked`."""
    return _PeerExchangePacked.apply(means2d, depths, conics, opacities, colors, radii, camera_ids, gaussian_ids,
                                     [[int(x) for x in row] for row in matrix], int(C_local), int(gauss_base))


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
