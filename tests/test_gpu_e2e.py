"""GPU: `gsplat.rendering.rasterization()` end to end under GSPLAT_BACKEND=b200 against the oracle
restatement of the reference's rendering.py, plus size-independent properties at BASELINE cfg2 size.
Protocol of the reference's tests/test_rasterization.py:22-89 (rtol = atol = 1e-4 on images/alphas).
"""
import math
import os

import pytest
import torch

import gs_oracle as O
from conftest import garden_scene

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _scene(step=25, cams=2):
    means, quats, scales, opac, colors, viewmats, Ks, W, H = garden_scene(step=step)
    return means, quats, scales * 3 + 0.002, opac, colors, viewmats[:cams], Ks[:cams], W, H


@pytest.mark.parametrize("sh_degree", [None, 3])
@pytest.mark.parametrize("render_mode", ["RGB", "RGB+D", "D", "RGB+ED"])
@pytest.mark.parametrize("packed", [False, True])
def test_rasterization_matches_oracle(sh_degree, render_mode, packed):
    from gsplat.rendering import rasterization

    means, quats, scales, opac, colors, viewmats, Ks, W, H = _scene()
    if sh_degree is not None:
        g = torch.Generator().manual_seed(42)
        colors = torch.cat([colors[:, None, :], torch.randn(len(means), 15, 3, generator=g) * 0.2], dim=1)
    bg = torch.rand(len(viewmats), 3, generator=torch.Generator().manual_seed(1))
    cpu = (means, quats, scales, opac, colors, viewmats, Ks)
    from parity import check_pair, run_pair

    r = run_pair(cpu[:5], viewmats, Ks, W, H, dict(sh_degree=sh_degree, render_mode=render_mode, backgrounds=bg),
                 packed=packed)
    rc, ra, meta, o_meta = r["rc"], r["ra"], r["meta"], r["o_meta"]
    assert ra.shape == (len(viewmats), H, W, 1)
    check_pair(r, f"e2e {render_mode} sh={sh_degree} packed={packed}")
    for key in ("camera_ids", "gaussian_ids", "radii", "means2d", "depths", "conics", "opacities", "tile_width",
                "tile_height", "tiles_per_gauss", "isect_ids", "flatten_ids", "isect_offsets", "width", "height",
                "tile_size", "n_cameras"):
        assert key in meta, key
    if not packed:
        assert meta["radii"].shape == o_meta["radii"].shape
        assert float((meta["radii"].cpu() != o_meta["radii"]).float().mean()) < 2e-3


@pytest.mark.parametrize("antialiased", [False, True])
def test_rasterization_gradients_match_oracle(antialiased):
    from gsplat.rendering import rasterization

    means, quats, scales, opac, colors, viewmats, Ks, W, H = _scene(step=60, cams=2)
    g = torch.Generator().manual_seed(42)
    shs = torch.cat([colors[:, None, :], torch.randn(len(means), 8, 3, generator=g) * 0.2], dim=1)
    w_c = torch.randn(2, H, W, 3, generator=g)
    w_a = torch.randn(2, H, W, 1, generator=g)
    mode = "antialiased" if antialiased else "classic"

    def run(fn, dev, **kw):
        p = [t.clone().to(dev).requires_grad_(True) for t in (means, quats, scales, opac, shs)]
        rc, ra, meta = fn(p[0], p[1], p[2], p[3], p[4], viewmats.to(dev), Ks.to(dev), W, H, sh_degree=2,
                          rasterize_mode=mode, **kw)
        if dev == DEV:
            meta["means2d"].retain_grad()  # the densification strategy's contract (strategy/default.py:150)
        loss = (rc * w_c.to(dev)).sum() + (ra * w_a.to(dev)).sum()
        loss.backward()
        return [t.grad for t in p], meta

    from parity import check_pair, run_pair

    grads, meta = run(rasterization, DEV, packed=False, absgrad=True)
    assert meta["means2d"].grad is not None and meta["means2d"].absgrad.shape == meta["means2d"].shape
    tol = {"means": 1e-3, "quats": 5e-3, "scales": 5e-2, "opacities": 1e-3, "colors": 1e-3}
    for packed in (False, True):
        check_pair(run_pair((means, quats, scales, opac, shs), viewmats, Ks, W, H,
                            dict(sh_degree=2, rasterize_mode=mode), packed=packed), f"grads {mode} packed={packed}")
    p_grads, _ = run(rasterization, DEV, packed=True)
    from parity import assert_grads_close

    for name, a, b in zip(tol, p_grads, grads):  # packed vs unpacked: same kernels, same lists, other summation order
        assert_grads_close(a, b, tol[name], 1e-4, f"packed vs unpacked v_{name}")


def test_rasterization_covars_channels_and_ortho():
    from gsplat.rendering import rasterization

    means, quats, scales, opac, colors, viewmats, Ks, W, H = _scene(step=50, cams=1)
    feats = torch.rand(len(means), 40, generator=torch.Generator().manual_seed(5))  # > channel_chunk
    covars = O.quat_scale_to_covar(quats, scales)
    cpu = (means, quats, scales, opac, feats, viewmats, Ks)
    from parity import assert_close_classified, check_pair, run_pair

    r = run_pair(cpu[:5], viewmats, Ks, W, H, dict(covars=covars))
    assert r["rc"].shape == (1, H, W, 40)
    assert_close_classified(r["rc"], r["o_rc"], r["flag"], 1e-4, 1e-4, "40 channels + covars")
    Ko = Ks.clone()
    Ko[:, 0, 0] = Ko[:, 1, 1] = 150.0
    cpu = (means, quats, scales, opac, colors, viewmats, Ko)
    check_pair(run_pair(cpu[:5], viewmats, Ko, W, H, dict(camera_model="ortho")), "ortho")
    with pytest.raises(NotImplementedError):
        rasterization(*[t.to(DEV) for t in cpu], W, H, camera_model="equirectangular")


@pytest.mark.parametrize("tile_size,packed", [(8, False), (4, True)])
def test_rasterization_other_tile_sizes(tile_size, packed):
    """rasterization(tile_size != 16) end to end: intersection keys, offsets and the plain compositor."""
    from parity import check_pair, run_pair

    means, quats, scales, opac, colors, viewmats, Ks, W, H = _scene(step=60, cams=2)
    r = run_pair((means, quats, scales, opac, colors), viewmats, Ks, W, H, dict(tile_size=tile_size), packed=packed)
    assert r["meta"]["tile_size"] == tile_size and r["meta"]["isect_offsets"].shape[1:] == (
        -(-H // tile_size), -(-W // tile_size))
    check_pair(r, f"tile_size {tile_size} packed={packed}")


def test_full_size_properties_cfg2():
    """BASELINE cfg2 (1M Gaussians, SH deg 3, 1920x1080): size-independent properties."""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from bench import make_cfg2
    from gsplat.rendering import rasterization

    s = make_cfg2(DEV)
    p = [s[k].requires_grad_(True) for k in ("means", "quats", "scales", "opacities", "shs")]
    rc, ra, meta = rasterization(*p, s["viewmats"], s["Ks"], s["width"], s["height"], sh_degree=3, packed=False)
    ids, flat, offs = meta["isect_ids"], meta["flatten_ids"], meta["isect_offsets"]
    n_isects = ids.numel()
    assert n_isects == int(meta["tiles_per_gauss"].sum())  # count == emitted
    assert bool((ids[1:] >= ids[:-1]).all())  # sorted
    tb = (meta["tile_width"] * meta["tile_height"] - 1).bit_length()
    tiles = (ids >> 32) & ((1 << tb) - 1)
    counts = torch.bincount(tiles, minlength=offs.numel())
    assert torch.equal((torch.cumsum(counts, 0) - counts).int(), offs.flatten())  # offsets == histogram scan
    assert torch.equal(torch.sort(flat.long()).values,
                       torch.repeat_interleave(torch.arange(flat.max() + 1, device=DEV),
                                               meta["tiles_per_gauss"].flatten()[: flat.max() + 1]))
    depth_bits = (ids & 0xFFFFFFFF).int().view(torch.float32)
    assert torch.equal(depth_bits, meta["depths"].flatten()[flat.long()])  # key carries the depth bit pattern
    assert float(ra.min()) >= 0.0 and float(ra.max()) <= 1.0 and bool(torch.isfinite(rc).all())
    (rc.sum() + ra.sum()).backward()
    for t in p:
        assert t.grad is not None and bool(torch.isfinite(t.grad).all())
    # linearity of the compositor in the colours: render(2c) == 2 render(c)
    with torch.no_grad():
        rc2, _, _ = rasterization(p[0], p[1], p[2], p[3], torch.rand(len(p[0]), 3, device=DEV) * 0 + 0.5,
                                  s["viewmats"], s["Ks"], s["width"], s["height"], packed=False)
        rc4, ra4, _ = rasterization(p[0], p[1], p[2], p[3], torch.ones(len(p[0]), 3, device=DEV), s["viewmats"],
                                    s["Ks"], s["width"], s["height"], packed=False)
    torch.testing.assert_close(rc2 * 2, rc4, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(rc4[..., :1], ra4, rtol=1e-4, atol=1e-5)  # unit colours render the alpha


def _ring_scene(N, C, W, H, seed=0):
    """BASELINE cfg3/cfg4 protocol: cameras on a ring of radius 6 looking at the origin."""
    g = torch.Generator().manual_seed(seed)
    means = torch.rand(N, 3, generator=g) * 6 - 3
    quats = torch.randn(N, 4, generator=g)
    scales = torch.exp(torch.rand(N, 3, generator=g) * (math.log(0.02) - math.log(0.002)) + math.log(0.002)) * 4
    opac = torch.rand(N, generator=g)
    shs = torch.randn(N, 16, 3, generator=g) * 0.3
    shs[:, 0] += 0.5
    fx = W / (2 * math.tan(math.radians(30)))
    Ks = torch.tensor([[fx, 0, W / 2], [0, fx, H / 2], [0, 0, 1.0]]).repeat(C, 1, 1)
    viewmats = []
    for i in range(C):
        a = 2 * math.pi * i / C
        eye = torch.tensor([6 * math.cos(a), 0.5, 6 * math.sin(a)])
        fwd = -eye / eye.norm()
        right = torch.linalg.cross(torch.tensor([0.0, 1.0, 0.0]), fwd)
        right = right / right.norm()
        up = torch.linalg.cross(fwd, right)
        R = torch.stack([right, up, fwd])  # world -> camera rows
        m = torch.eye(4)
        m[:3, :3] = R
        m[:3, 3] = -R @ eye
        viewmats.append(m)
    return means, quats, scales, opac, shs, torch.stack(viewmats), Ks


def test_cfg3_like_packed_antialiased_matches_oracle():
    """cfg3 shape (C=8 ring cameras, packed + antialiased, SH deg 3) at a size the CPU oracle finishes in seconds."""
    from gsplat.rendering import rasterization

    W, H, C = 324, 210, 8
    cpu = _ring_scene(6000, C, W, H)
    from parity import check_pair, run_pair

    r = run_pair(cpu[:5], cpu[5], cpu[6], W, H, dict(sh_degree=3, rasterize_mode="antialiased"), packed=True)
    check_pair(r, "cfg3-like packed antialiased")
    meta = r["meta"]
    assert meta["camera_ids"].dtype == torch.int64 and int(meta["camera_ids"].max()) == C - 1
    assert meta["isect_offsets"].shape == (C, 14, 21)


def test_cfg3_full_size_packed_equals_unpacked_and_properties():
    """cfg3 at full camera count / resolution (C=8, 1297x840): packed == unpacked images and gradients at 300k
    Gaussians; intersection keys sorted with the camera in the top bits."""
    from gsplat.rendering import rasterization

    W, H, C = 1297, 840, 8
    cpu = _ring_scene(300_000, C, W, H, seed=1)
    outs = []
    for packed in (True, False):
        p = [t.to(DEV).requires_grad_(True) for t in cpu[:5]]
        rc, ra, meta = rasterization(*p, cpu[5].to(DEV), cpu[6].to(DEV), W, H, sh_degree=3, packed=packed,
                                     rasterize_mode="antialiased")
        (rc.square().sum() + ra.sum()).backward()
        outs.append((rc.detach(), ra.detach(), [t.grad for t in p], meta))
    torch.testing.assert_close(outs[0][0], outs[1][0], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(outs[0][1], outs[1][1], rtol=1e-5, atol=1e-6)
    from parity import assert_grads_close

    for name, a, b in zip(("means", "quats", "scales", "opacities", "colors"), outs[0][2], outs[1][2]):
        assert_grads_close(a, b, 1e-3, 1e-4, f"cfg3 packed vs unpacked v_{name}")
    ids = outs[0][3]["isect_ids"]
    assert bool((ids[1:] >= ids[:-1]).all())
    tb = (82 * 53 - 1).bit_length()
    assert int((ids >> (32 + tb)).max()) == C - 1
    assert outs[0][3]["flatten_ids"].numel() == outs[1][3]["flatten_ids"].numel()


def test_cfg5_4k_tile_grid_properties():
    """cfg5's largest grid (3840x2160 -> 240x135 = 32400 tiles): offsets == histogram scan, alpha in [0,1]."""
    from gsplat.rendering import rasterization

    W, H = 3840, 2160
    g = torch.Generator().manual_seed(2)
    N = 400_000
    fx = 0.5 * W / math.tan(math.radians(30))
    z = torch.rand(N, generator=g) * 8 + 2
    means = torch.stack([(torch.rand(N, generator=g) * 2.1 - 1.05) * z * W / (2 * fx),
                         (torch.rand(N, generator=g) * 2.1 - 1.05) * z * H / (2 * fx), z], dim=-1)
    quats = torch.randn(N, 4, generator=g)
    scales = torch.exp(torch.rand(N, 3, generator=g) * 2.3 + math.log(0.002))
    opac, colors = torch.rand(N, generator=g), torch.rand(N, 3, generator=g)
    Ks = torch.tensor([[[fx, 0, W / 2], [0, fx, H / 2], [0, 0, 1.0]]])
    rc, ra, meta = rasterization(*[t.to(DEV) for t in (means, quats, scales, opac, colors)],
                                 torch.eye(4)[None].to(DEV), Ks.to(DEV), W, H, packed=False)
    offs, ids = meta["isect_offsets"], meta["isect_ids"]
    assert offs.shape == (1, 135, 240)
    tiles = (ids >> 32) & ((1 << 15) - 1)
    counts = torch.bincount(tiles, minlength=offs.numel())
    assert torch.equal((torch.cumsum(counts, 0) - counts).int(), offs.flatten())
    assert float(ra.min()) >= 0 and float(ra.max()) <= 1 and bool(torch.isfinite(rc).all())


@pytest.mark.gpu
def test_distributed_peer_exchange_two_gpus():
    """rasterization(distributed=True) on 2 GPUs, fused peer-store exchange and NCCL all-to-all, vs a
    single-GPU render of the whole scene (tools/dist_check.py). Needs two visible GPUs."""
    import subprocess
    import sys

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
         "127.0.0.1", "--master-port", "29533", os.path.join(root, "tools", "dist_check.py")],
        capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MISMATCH" not in r.stdout


@pytest.mark.parametrize("packed", [False, True])
def test_speculative_intersection_buffers_match_exact(packed, monkeypatch):
    """rasterization() sizes the intersection buffers from the previous call's count and reads the real count
    back after the compositor is queued. Same ids / offsets / images / gradients as the exact path, also when the
    estimate is too small (overflow -> exact redo) and when nothing is visible."""
    from gsplat.b200_impl import _wrapper as Wp
    from gsplat.rendering import rasterization

    means, quats, scales, opac, colors, viewmats, Ks, W, H = _scene(step=40, cams=2)
    args = [t.to(DEV) for t in (means, quats, scales, opac, colors, viewmats, Ks)]

    def run(speculate, estimate=None, shift=0.0):
        monkeypatch.setenv("GSPLAT_B200_SPECULATE", "1" if speculate else "0")
        Wp._ISECT_ESTIMATE.clear()
        if estimate is not None:
            Wp._ISECT_ESTIMATE[(torch.cuda.current_device(), 2, -(-W // 16), -(-H // 16), 16, packed)] = estimate
        p = [t.clone().requires_grad_(True) for t in args[:5]]
        p[0].data[:, 2] += shift
        rc, ra, meta = rasterization(*p, args[5], args[6], W, H, packed=packed)
        (rc.sum() + 2 * ra.sum()).backward()
        return rc, ra, meta, [t.grad for t in p]

    rc0, ra0, m0, g0 = run(False)
    n = m0["flatten_ids"].numel()
    assert n > 0
    for estimate in (n, n * 2, n // 3, 0):  # exact fit (+margin), oversized, overflow -> redo, tiny
        rc, ra, m, g = run(True, estimate)
        for k in ("isect_ids", "flatten_ids", "isect_offsets", "tiles_per_gauss"):
            assert torch.equal(m[k], m0[k]), (estimate, k)
        assert torch.equal(rc, rc0) and torch.equal(ra, ra0), estimate
        for a, b in zip(g, g0):
            assert torch.allclose(a, b, rtol=1e-4, atol=1e-6 + 1e-4 * b.abs().max().item()), estimate
        assert Wp._ISECT_ESTIMATE, "the count must be remembered for the next call"
    # nothing visible (all Gaussians far behind the cameras) with a stale, large estimate
    rc, ra, m, _ = run(True, n, shift=-1e4)
    assert m["flatten_ids"].numel() == 0 and m["isect_ids"].numel() == 0
    assert float(ra.abs().max()) == 0.0 and int(m["isect_offsets"].abs().max()) == 0


@pytest.mark.parametrize("packed", [False, True])
def test_launch_and_stream_variants_give_the_same_render(packed, monkeypatch):
    """The scheduling switches of the round-2 path change WHEN kernels run, never what they compute: chained launches
    (programmatic dependent launch) off / on for every kernel group, one stream instead of two, the gradient buffer
    zero-filled in the backward pass, plain / cp.async.bulk / cp.async staging of the SH colour kernels. Images, ids and
    offsets bit-identical (images within 2e-6 across the separately compiled SH kernels); gradients to the last few
    ulps (atomics accumulate in scheduling order)."""
    from gsplat.b200_impl._lib import lib
    from gsplat.rendering import rasterization

    torch.manual_seed(11)
    N, W, H, C = 30_000, 320, 200, 2
    means = torch.rand(N, 3) * torch.tensor([6.0, 4.0, 4.0]) + torch.tensor([-3.0, -2.0, 2.0])
    quats, scales, opac = torch.randn(N, 4), torch.rand(N, 3) * 0.08 + 0.01, torch.rand(N)
    shs = torch.randn(N, 16, 3) * 0.3
    viewmats = torch.eye(4).repeat(C, 1, 1)
    viewmats[1, 0, 3] = 0.3
    Ks = torch.tensor([[200.0, 0.0, W / 2], [0.0, 200.0, H / 2], [0.0, 0.0, 1.0]]).repeat(C, 1, 1)
    args = [t.to(DEV) for t in (means, quats, scales, opac, shs, viewmats, Ks)]
    CHAIN, SH_BWD, SH_FWD = 8, 3, 6  # GSB_OPT_CHAINED_LAUNCH, GSB_OPT_SH_STAGING, GSB_OPT_SH_FWD_STAGING
    defaults = {o: lib.gsb_get_option(o) for o in (CHAIN, SH_BWD, SH_FWD)}

    def run(options=(), env=()):
        for k, v in env:
            monkeypatch.setenv(k, v)
        for o, v in options:
            lib.gsb_set_option(o, v)
        try:
            out = None
            for _ in range(2):  # the second call runs with speculatively sized intersection buffers
                p = [t.clone().requires_grad_(True) for t in args[:5]]
                rc, ra, meta = rasterization(*p, args[5], args[6], W, H, sh_degree=3, packed=packed)
                (rc * rc).sum().backward()
                torch.cuda.synchronize()
                out = (rc.detach(), ra.detach(), meta["isect_ids"], meta["flatten_ids"], meta["isect_offsets"],
                       [t.grad for t in p])
            return out
        finally:
            for o, v in defaults.items():
                lib.gsb_set_option(o, v)
            for k, _ in env:
                monkeypatch.delenv(k, raising=False)

    ref = run()
    variants = {
        "no chained launches": dict(options=[(CHAIN, 0)]),
        "every group chained": dict(options=[(CHAIN, 63)]),
        "one stream": dict(env=[("GSPLAT_B200_OVERLAP", "0")]),
        "gradient buffer filled in the backward pass": dict(env=[("GSPLAT_B200_PREZERO", "0")]),
        "plain SH staging": dict(options=[(SH_BWD, 0), (SH_FWD, 0)]),
        "cp.async.bulk SH staging": dict(options=[(SH_BWD, 1), (SH_FWD, 1)]),
    }
    for name, kw in variants.items():
        out = run(**kw)
        for a, b, what in zip(out[:5], ref[:5], ("colours", "alphas", "isect_ids", "flatten_ids", "offsets")):
            if "SH staging" in name and what == "colours":  # other kernels: same arithmetic, compiled separately
                assert torch.allclose(a, b, rtol=0.0, atol=2e-6), (name, what)
            else:
                assert torch.equal(a, b), (name, what)
        for a, b in zip(out[5], ref[5]):
            assert torch.allclose(a, b, rtol=1e-4, atol=1e-6 + 1e-5 * b.abs().max().item()), name


def test_rasterization_fisheye_matches_oracle():
    """camera_model="fisheye" end to end (forward image + gradients) against the oracle restatement."""
    from gsplat.rendering import rasterization

    means, quats, scales, opac, colors, viewmats, Ks, W, H = _scene(step=60, cams=2)
    g = torch.Generator().manual_seed(3)
    w_c = torch.randn(2, H, W, 3, generator=g)

    def run(fn, dev, **kw):
        p = [t.clone().to(dev).requires_grad_(True) for t in (means, quats, scales, opac, colors)]
        rc, ra, _ = fn(*p, viewmats.to(dev), Ks.to(dev), W, H, camera_model="fisheye", **kw)
        ((rc * w_c.to(dev)).sum() + ra.sum()).backward()
        return rc, ra, [t.grad for t in p]

    from parity import check_pair, run_pair

    r = run_pair((means, quats, scales, opac, colors), viewmats, Ks, W, H, dict(camera_model="fisheye"))
    assert float(r["o_ra"].mean()) > 0.01
    check_pair(r, "fisheye")


@pytest.mark.parametrize("packed,absgrad", [(False, False), (True, True)])
def test_training_loop_contract(packed, absgrad):
    """The trainer's use of rasterization() (examples/simple_trainer.py:483-560 + strategy/default.py:140-262):
    activations outside, `info["means2d"].retain_grad()` before backward, `.grad` / `.absgrad`, `radii`,
    `gaussian_ids`, `width/height/n_cameras` after it, Adam steps on the raw parameters. The L1 loss against
    a target rendered from perturbed parameters must drop, the densification statistics must be sane, and a
    changing Gaussian count (prune) must be accepted from one step to the next."""
    from gsplat.rendering import rasterization

    means, quats, scales, opac, colors, viewmats, Ks, W, H = _scene(step=80, cams=2)
    dev = DEV
    viewmats, Ks = viewmats.to(dev), Ks.to(dev)
    g = torch.Generator().manual_seed(11)
    sh0 = ((colors - 0.5) / 0.2820947917738781)[:, None, :]
    shN = torch.zeros(len(means), 3, 3)

    def render(p, **kw):
        return rasterization(p["means"], p["quats"], torch.exp(p["scales"]), torch.sigmoid(p["opacities"]),
                             torch.cat([p["sh0"], p["shN"]], 1), viewmats, Ks, W, H, sh_degree=1, packed=packed,
                             absgrad=absgrad, **kw)

    truth = {"means": means, "quats": quats, "scales": torch.log(scales), "opacities": torch.logit(opac * 0.9 + 0.05),
             "sh0": sh0, "shN": shN}
    truth = {k: v.to(dev) for k, v in truth.items()}
    with torch.no_grad():
        target, _, _ = render(truth)
    params = {k: v.clone() for k, v in truth.items()}
    params["means"] += 0.01 * torch.randn(means.shape, generator=g).to(dev)
    params["sh0"] += 0.5 * torch.randn(sh0.shape, generator=g).to(dev)
    params["opacities"] -= 1.0
    params = {k: v.requires_grad_(True) for k, v in params.items()}
    opt = torch.optim.Adam([{"params": [v], "lr": lr} for v, lr in zip(
        params.values(), (1e-4, 1e-3, 5e-3, 5e-2, 2.5e-2, 1e-3))])
    state = {"grad2d": torch.zeros(len(means), device=dev), "count": torch.zeros(len(means), device=dev)}
    losses = []
    for step in range(40):
        rc, ra, info = render(params)
        info["means2d"].retain_grad()  # strategy.step_pre_backward
        loss = (rc - target).abs().mean()
        loss.backward()
        losses.append(float(loss))
        for key in ("width", "height", "n_cameras", "radii", "gaussian_ids", "means2d"):  # _update_state
            assert key in info
        grads = (info["means2d"].absgrad if absgrad else info["means2d"].grad).clone()
        assert grads.shape == info["means2d"].shape and bool(torch.isfinite(grads).all())
        grads[..., 0] *= info["width"] / 2.0 * info["n_cameras"]
        grads[..., 1] *= info["height"] / 2.0 * info["n_cameras"]
        if packed:
            ids = info["gaussian_ids"]
        else:
            sel = info["radii"] > 0
            ids, grads = torch.where(sel)[1], grads[sel]
        state["grad2d"].index_add_(0, ids, grads.norm(dim=-1))
        state["count"].index_add_(0, ids, torch.ones_like(ids, dtype=torch.float32))
        opt.step()
        opt.zero_grad(set_to_none=True)
    assert losses[-1] < 0.6 * losses[0], (losses[0], losses[-1])
    assert float(state["count"].max()) == 2 * 40 and float(state["grad2d"].sum()) > 0
    # prune (a new Gaussian count from one call to the next) and render again
    keepers = torch.sigmoid(params["opacities"]) > 0.2
    pruned = {k: v.detach()[keepers].requires_grad_(True) for k, v in params.items()}
    rc, ra, info = render(pruned)
    rc.mean().backward()
    assert pruned["means"].grad.shape == pruned["means"].shape and bool(torch.isfinite(rc).all())


def test_selective_adam_matches_the_formula():
    """gsplat.optimizers.SelectiveAdam (reference optimizers/selective_adam.py + AdamCUDA.cu:13-52): visible rows
    follow m, v, p -= lr m / (sqrt(v) + eps); invisible rows keep parameter and moments, for [N], [N,3], [N,K,3]."""
    from gsplat.optimizers import SelectiveAdam

    g = torch.Generator().manual_seed(9)
    N = 1000
    vis = torch.rand(N, generator=g) > 0.4
    for shape, lr in (((N,), 1e-2), ((N, 3), 1e-3), ((N, 16, 3), 2.5e-3)):
        p0 = torch.randn(shape, generator=g)
        param = p0.clone().to(DEV).requires_grad_(True)
        opt = SelectiveAdam([{"params": [param], "lr": lr}], eps=1e-15, betas=(0.9, 0.999))
        m, v, ref = torch.zeros(shape), torch.zeros(shape), p0.clone()
        sel = vis.reshape((N,) + (1,) * (len(shape) - 1)).expand(shape)
        for _ in range(3):
            grad = torch.randn(shape, generator=g)
            param.grad = grad.to(DEV)
            opt.step(visibility=vis.to(DEV))
            # the kernel forms 1 - beta in float32, as the reference's does (1 - 0.999f = 0.00100005)
            b1, b2 = torch.tensor(0.9), torch.tensor(0.999)
            m_new, v_new = b1 * m + (1 - b1) * grad, b2 * v + (1 - b2) * grad * grad
            m, v = torch.where(sel, m_new, m), torch.where(sel, v_new, v)
            ref = torch.where(sel, ref - lr * m / (v.sqrt() + 1e-15), ref)
        torch.testing.assert_close(param.detach().cpu(), ref, rtol=1e-5, atol=1e-6)
        st = opt.state[param]
        torch.testing.assert_close(st["exp_avg"].cpu(), m, rtol=1e-5, atol=1e-7)
        torch.testing.assert_close(st["exp_avg_sq"].cpu(), v, rtol=1e-5, atol=1e-10)
        assert torch.equal(param.detach().cpu()[~vis], p0[~vis])


def test_compute_relocation_matches_equation_9():
    """gsplat.relocation.compute_relocation (reference relocation.py:10-50, RelocationCUDA.cu:13-45) against a
    float64 restatement of eq. 9; ratios beyond the table are clamped in place as the reference does."""
    import math

    from gsplat.relocation import compute_relocation

    g = torch.Generator().manual_seed(11)
    N, n_max = 4096, 51
    binoms = torch.zeros(n_max, n_max)
    for n in range(n_max):
        for k in range(n + 1):
            binoms[n, k] = math.comb(n, k)
    opac = torch.rand(N, generator=g) * 0.98 + 0.01
    scales = torch.rand(N, 3, generator=g) * 0.1
    ratios = torch.randint(0, 60, (N,), generator=g)
    r_dev = ratios.to(DEV)
    new_o, new_s = compute_relocation(opac.to(DEV), scales.to(DEV), r_dev, binoms.to(DEV))
    assert int(r_dev.min()) >= 1 and int(r_dev.max()) <= n_max
    n = ratios.clamp(1, n_max).double()
    o = opac.double()
    ref_o = 1.0 - (1.0 - o) ** (1.0 / n)
    denom = torch.zeros(N, dtype=torch.float64)
    B = binoms.double()
    for i in range(1, n_max + 1):
        for k in range(i):
            term = B[i - 1, k] * ((-1.0) ** k / math.sqrt(k + 1)) * ref_o ** (k + 1)
            denom += torch.where(n >= i, term, torch.zeros_like(term))
    ref_s = (o / denom)[:, None] * scales.double()
    # float32 cancellation in 1 - (1 - o)^(1/n) for small o and large n costs ~1e-4 relative (emulated on the CPU)
    torch.testing.assert_close(new_o.cpu().double(), ref_o, rtol=2e-3, atol=1e-7)
    torch.testing.assert_close(new_s.cpu().double(), ref_s, rtol=2e-3, atol=1e-8)
    assert torch.isfinite(new_s).all()


@pytest.mark.parametrize("case", ["sh", "rgb"])
def test_rasterization_matches_the_reference_end_to_end_vectors(golden, case):
    """The CUDA path against tests/golden/e2e.npz: outputs and autograd gradients of the REFERENCE's own pure-torch
    `_rasterization` (rendering.py:601-760, frozen by oracle/make_golden.py::case_e2e)."""
    from gsplat.rendering import rasterization

    g = golden("e2e")
    W, H = int(g["width"]), int(g["height"])
    colors = g["shs"] if case == "sh" else g["rgb"]
    kw = dict(sh_degree=3, render_mode="RGB+ED") if case == "sh" else dict(rasterize_mode="antialiased",
                                                                            backgrounds=g["backgrounds"])
    from parity import assert_close_classified, assert_grads_close, check_pair, run_pair

    cpu = [g[k] for k in ("means", "quats", "scales", "opacities")] + [colors]
    r = run_pair(cpu, g["viewmats"], g["Ks"], W, H, kw)
    check_pair(r, f"e2e vectors {case} (vs oracle)")
    # the REFERENCE's frozen image / alpha: beyond-tolerance pixels only where a decision sits on a threshold
    assert_close_classified(r["rc"], g[f"{case}_render_colors"], r["flag"], 1e-4, 1e-4, f"{case} colours vs reference")
    assert_close_classified(r["ra"], g[f"{case}_render_alphas"], r["flag"], 1e-4, 1e-4, f"{case} alphas vs reference")
    # the REFERENCE's frozen autograd gradients under the full loss (flips stay in): zero outliers at rtol + 1e-3 max|g|
    leaves = [t.clone().to(DEV).requires_grad_(True) for t in cpu]
    dkw = {k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in kw.items()}
    rc, ra, _ = rasterization(*leaves, g["viewmats"].to(DEV), g["Ks"].to(DEV), W, H, **dkw)
    loss = (rc * g[f"{case}_w_colors"].to(DEV)).sum() + (ra * g[f"{case}_w_alphas"].to(DEV)).sum()
    grads = torch.autograd.grad(loss, leaves)
    tol = {"means": 1e-3, "quats": 5e-3, "scales": 5e-2, "opacities": 1e-3, "colors": 1e-3}
    for name, got in zip(tol, grads):
        assert_grads_close(got, g[f"{case}_v_{name}"], tol[name], 1e-3, f"{case} v_{name} vs reference (full loss)")
