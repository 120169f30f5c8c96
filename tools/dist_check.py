"""Multi-GPU parity check of `rasterization(distributed=True)` (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/dist_check.py

Every rank renders its own cameras from Gaussians sharded over the ranks, once through the fused
peer-store exchange (csrc/peer.cu) and once through the NCCL all-to-all, unpacked and packed rows, RGB and
RGB+ED (antialiased), and compares images, alphas and the gradients of its Gaussian shard with a plain
single-GPU render of the whole scene (all cameras, all Gaussians) computed locally. Also runs forward-only
steps between training steps (mailbox double buffering) and uneven shard sizes. Prints one line per rank and
(rank 0) a final JSON summary line `DIST_CHECK {...}`; exit code 1 on mismatch.
"""
import math
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gsplat-triton_b200"))
os.environ.setdefault("GSPLAT_BACKEND", "b200")


def scene(n, n_cams, W, H, seed=0):
    g = torch.Generator().manual_seed(seed)
    means = (torch.rand(n, 3, generator=g) - 0.5) * 4.0
    quats = torch.randn(n, 4, generator=g)
    scales = torch.exp(torch.rand(n, 3, generator=g) * 1.5 - 3.5)
    opac = torch.rand(n, generator=g)
    shs = torch.randn(n, 9, 3, generator=g) * 0.3
    viewmats = []
    for i in range(n_cams):
        az = 2 * math.pi * i / n_cams
        eye = torch.tensor([6 * math.cos(az), 0.5, 6 * math.sin(az)])
        fwd = -eye / eye.norm()
        right = torch.linalg.cross(torch.tensor([0.0, 1.0, 0.0]), fwd)
        right = right / right.norm()
        up = torch.linalg.cross(fwd, right)
        R = torch.stack([right, up, fwd])
        vm = torch.eye(4)
        vm[:3, :3] = R
        vm[:3, 3] = -R @ eye
        viewmats.append(vm)
    f = 0.5 * W / math.tan(math.radians(30))
    Ks = torch.tensor([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1.0]]).expand(n_cams, 3, 3).contiguous()
    return [means, quats, scales, opac, shs], torch.stack(viewmats), Ks


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev)
    from gsplat.distributed import close_mailboxes
    from gsplat.rendering import rasterization

    W, H, C_local = 320, 200, 2
    ok, worst, n_cases = True, [0.0, 0.0, 0.0], 0
    for n, uneven in ((20000, False), (30011, True)):
        params, viewmats, Ks = scene(n, world * C_local, W, H)
        # shard: contiguous blocks (uneven sizes when asked) so global Gaussian order = concatenation over ranks
        cuts = [0]
        for r in range(world):
            share = n // world + (37 * (r + 1) if uneven and r + 1 < world else 0)
            cuts.append(min(n, cuts[-1] + share) if r + 1 < world else n)
        lo, hi = cuts[rank], cuts[rank + 1]
        my_cams = slice(rank * C_local, (rank + 1) * C_local)
        g = torch.Generator().manual_seed(7)
        Gw = torch.rand(world * C_local, H, W, 3, generator=g).to(dev)
        Aw = torch.rand(world * C_local, H, W, 1, generator=g).to(dev)

        for packed, mode, raster in ((False, "RGB", "classic"), (False, "RGB+ED", "antialiased"),
                                     (True, "RGB", "classic"), (True, "RGB+ED", "antialiased")):
            Dch = 4 if mode == "RGB+ED" else 3
            Gc = Gw if Dch == 3 else torch.cat([Gw, Gw[..., :1] * 0.1], dim=-1)
            kw = dict(sh_degree=2, render_mode=mode, rasterize_mode=raster)
            # single-GPU truth: all cameras, all Gaussians
            full = [p.to(dev).requires_grad_(True) for p in params]
            rc, ra, _ = rasterization(*full, viewmats.to(dev), Ks.to(dev), W, H, packed=False, **kw)
            ((rc * Gc).sum() + (ra * Aw).sum()).backward()
            truth = dict(rc=rc[my_cams].detach(), ra=ra[my_cams].detach(),
                         grads=[p.grad[lo:hi].clone() for p in full])
            del full, rc, ra
            for transport in ("peer", "peer_push", "peer_split", "nccl"):
                if transport in ("peer_split", "peer_push") and packed:
                    continue  # packed rows have a single peer path (one fused put per direction)
                os.environ["GSPLAT_B200_EXCHANGE"] = "nccl" if transport == "nccl" else "peer"
                os.environ["GSPLAT_B200_EXCHANGE_SPLIT"] = "1" if transport == "peer_split" else "0"
                os.environ["GSPLAT_B200_EXCHANGE_PUSH"] = "1" if transport == "peer_push" else "0"
                for step in range(3):
                    shard = [p[lo:hi].to(dev).requires_grad_(True) for p in params]
                    args = (viewmats[my_cams].to(dev), Ks[my_cams].to(dev), W, H)
                    if step == 1:  # forward-only steps in between (mailbox reuse without a backward exchange)
                        with torch.no_grad():
                            for _ in range(3):
                                rasterization(*shard, *args, packed=packed, distributed=True, **kw)
                    rc, ra, meta = rasterization(*shard, *args, packed=packed, distributed=True, **kw)
                    ((rc * Gc[my_cams]).sum() + (ra * Aw[my_cams]).sum()).backward()
                    e_img = (rc - truth["rc"]).abs().max().item()
                    e_a = (ra - truth["ra"]).abs().max().item()
                    e_g = max(((p.grad - t).abs().max() / (t.abs().max() + 1e-12)).item()
                              for p, t in zip(shard, truth["grads"]))
                    good = e_img < 2e-4 and e_a < 2e-4 and e_g < 2e-3
                    ok &= good
                    worst = [max(a, b) for a, b in zip(worst, (e_img, e_a, e_g))]
                    n_cases += 1
                    print(f"rank {rank} n={n} uneven={uneven} packed={packed} {mode} {transport} step {step}: "
                          f"img {e_img:.2e} alpha {e_a:.2e} grad {e_g:.2e} {'ok' if good else 'MISMATCH'}", flush=True)
    torch.cuda.synchronize()
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    w = torch.tensor(worst, device=dev, dtype=torch.float64)
    dist.all_reduce(w, op=dist.ReduceOp.MAX)
    if rank == 0:
        import json

        print("DIST_CHECK " + json.dumps({
            "world": world, "cases_per_rank": n_cases, "ok": flag.item() == 0, "img_max_err": w[0].item(),
            "alpha_max_err": w[1].item(), "grad_max_rel": w[2].item(),
            "covers": "packed x {RGB classic, RGB+ED antialiased} x {peer put, peer push from the producer kernels, peer with split forward exchange, nccl} x {even, uneven shards} x 3 steps "
                      "(forward-only steps in between), vs a single-GPU render of the whole scene"}), flush=True)
    dist.barrier()
    close_mailboxes()
    dist.destroy_process_group()
    sys.exit(1 if flag.item() else 0)


if __name__ == "__main__":
    main()
