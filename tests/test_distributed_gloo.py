"""CPU, world_size=2 over gloo: the collectives `rasterization(distributed=True)` is built on
(reference gsplat/distributed.py docstring examples :130-138, :199-214 and tests/_test_distributed.py)."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import PKG


def _port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fn_name):
    sys.path.insert(0, PKG)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        globals()[fn_name](rank, world)
    finally:
        dist.barrier()
        dist.destroy_process_group()


def _spawn(fn_name, world=2):
    mp.spawn(_worker, args=(world, _port(), fn_name), nprocs=world, join=True)


def _ints(rank, world):
    from gsplat.distributed import all_gather_int32, all_to_all_int32

    dev = torch.device("cpu")
    assert all_gather_int32(world, rank + 1, device=dev) == [1, 2]
    got = all_gather_int32(world, torch.tensor(rank + 1, dtype=torch.int32))
    assert [int(x) for x in got] == [1, 2]
    # rank r sends values[i] to rank i
    sent = [10 * rank + i for i in range(world)]
    assert all_to_all_int32(world, sent, device=dev) == [10 * i + rank for i in range(world)]


def _gather_list(rank, world):
    from gsplat.distributed import all_gather_tensor_list

    a = torch.full((2, 3), float(rank), requires_grad=True)
    b = torch.full((2, 4, 4), float(rank + 10), requires_grad=True)
    ga, gb = all_gather_tensor_list(world, [a, b])
    assert ga.shape == (4, 3) and gb.shape == (4, 4, 4)
    assert ga[:, 0].tolist() == [0.0, 0.0, 1.0, 1.0] and gb[:, 0, 0].tolist() == [10.0, 10.0, 11.0, 11.0]
    # backward: every rank contributes (rank+1) * ones for every gathered row -> sum over ranks = 3
    ((ga.sum() + gb.sum()) * (rank + 1)).backward()
    assert torch.allclose(a.grad, torch.full_like(a, 3.0)) and torch.allclose(b.grad, torch.full_like(b, 3.0))


def _a2a_list(rank, world):
    from gsplat.distributed import all_to_all_rows, all_to_all_tensor_list, interleave_camera_major

    # reference docstring example (distributed.py:199-214), world 2: rank r holds rows tagged r
    n_rows = [3, 5][rank]
    splits = [[1, 2], [4, 1]][rank]
    x = (torch.arange(n_rows, dtype=torch.float32)[:, None] + 100 * rank).repeat(1, 2).requires_grad_(True)
    y = torch.arange(n_rows, dtype=torch.float32) + 100 * rank
    ox, oy = all_to_all_tensor_list(world, [x, y], splits)
    expect = {0: [0.0, 100.0, 101.0, 102.0, 103.0], 1: [1.0, 2.0, 104.0]}[rank]
    assert oy.tolist() == expect and ox[:, 0].tolist() == expect
    (ox.sum() * (rank + 1)).backward()  # gradient travels back through the reverse exchange
    dest = torch.repeat_interleave(torch.arange(world), torch.tensor(splits)).float()
    assert torch.allclose(x.grad, (dest + 1)[:, None].expand(-1, 2))
    # unpacked exchange layout: C_local cameras x N_j Gaussians per peer -> [C_local, sum N_j]
    C_local, N_world = 2, [3, 2]
    N = N_world[rank]
    rows = torch.stack([torch.full((N,), float(c)) * 10 + torch.arange(N) + 100 * rank
                        for c in range(world * C_local)]).reshape(-1, 1)  # [C_global * N, 1], camera-major
    recv = all_to_all_rows(rows, [C_local * N] * world, [C_local * n for n in N_world])
    out = interleave_camera_major(recv, C_local, N_world)
    assert out.shape == (C_local, sum(N_world), 1)
    for c in range(C_local):
        cam = rank * C_local + c
        want = [cam * 10 + n for n in range(3)] + [cam * 10 + n + 100 for n in range(2)]
        assert out[c, :, 0].tolist() == [float(v) for v in want]


def test_int_collectives():
    _spawn("_ints")


def test_all_gather_tensor_list_is_differentiable():
    _spawn("_gather_list")


def test_all_to_all_rows_and_relayout():
    _spawn("_a2a_list")


def test_world_size_one_is_identity():
    sys.path.insert(0, PKG)
    from gsplat.distributed import all_gather_int32, all_gather_tensor_list, all_to_all_tensor_list, cli

    t = [torch.ones(2, 3)]
    assert all_gather_int32(1, 7) == [7]
    assert all_gather_tensor_list(1, t) is t and all_to_all_tensor_list(1, t, [2]) is t
    assert cli(lambda lr, wr, ws, a: (lr, wr, ws, a), "x", world_size=1) == (0, 0, 1, "x")


def _host_board(rank, world):
    from gsplat.distributed import all_gather_int_host

    # shared-memory board: many gathers in a row (slot parity reuse), values change every call
    for i in range(200):
        assert all_gather_int_host(world, 1000 * rank + i) == [1000 * r + i for r in range(world)]
    assert all_gather_int_host(world, (1 << 40) + rank) == [(1 << 40) + r for r in range(world)]
    # vector form (the packed exchange's row-count matrix), interleaved with scalar gathers
    from gsplat.distributed import all_gather_ints_host

    for i in range(50):
        rows = all_gather_ints_host(world, [rank * 100 + i + d for d in range(world)])
        assert rows == [[r * 100 + i + d for d in range(world)] for r in range(world)]
        assert all_gather_int_host(world, i - rank) == [i - r for r in range(world)]
    assert all_gather_ints_host(world, list(range(16))) == [list(range(16))] * world


def test_host_board_all_gather():
    _spawn("_host_board")
