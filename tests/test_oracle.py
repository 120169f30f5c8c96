"""CPU: the oracle (oracle/gs_oracle.py) against the golden vectors frozen from the reference's
own `_torch_impl` by oracle/make_golden.py, plus the literal known-answer vectors of the reference."""
import numpy as np
import pytest
import torch

import gs_oracle as O


def _close(a, b, rtol=1e-4, atol=1e-4):
    torch.testing.assert_close(a, b, rtol=rtol, atol=atol)


def test_projection_matches_reference(golden):
    g = golden("projection")
    covars = O.quat_scale_to_covar(g["quats"], g["scales"])
    for model in ("pinhole", "ortho"):
        radii, means2d, depths, conics, comps = O.fully_fused_projection(
            g["means"], covars, g["viewmats"], g["Ks"], g["width"], g["height"], eps2d=0.3,
            calc_compensations=True, camera_model=model)
        # integer output: equal to the frozen reference values, except that a radius sitting exactly on a
        # ceil() boundary may flip by one when the CPU matmuls reduce in another order (thread count / load):
        # the reference's own tolerance (triton_tests/test_fused_proj.py:116, atol=1)
        diff = (radii - g[f"{model}_radii"]).abs()
        assert int(diff.max()) <= 1 and float((diff > 0).float().mean()) < 1e-3
        v = radii > 0
        assert v.sum() > 100
        _close(means2d[v], g[f"{model}_means2d"][v])
        _close(depths[v], g[f"{model}_depths"][v])
        _close(conics[v], g[f"{model}_conics"][v])
        _close(comps[v], g[f"{model}_comps"][v])


def test_projection_fisheye_matches_reference(golden):
    g = golden("projection_fisheye")
    p = [g[k].clone().requires_grad_(True) for k in ("means", "quats", "scales", "viewmats")]
    radii, means2d, depths, conics, comps = O.fully_fused_projection(
        p[0], O.quat_scale_to_covar(p[1], p[2]), p[3], g["Ks"], g["width"], g["height"], calc_compensations=True,
        camera_model="fisheye")
    diff = (radii - g["radii"]).abs()
    assert int(diff.max()) <= 1 and float((diff > 0).float().mean()) < 1e-3
    v = (radii > 0) & (g["radii"] > 0)
    for got, name in ((means2d, "means2d"), (depths, "depths"), (conics, "conics"), (comps, "comps")):
        _close(got[v], g[name][v])
    m = (g["radii"] > 0).float()
    loss = ((means2d * g["w_means2d"]).sum(-1) * m + depths * g["w_depths"] * m
            + (conics * g["w_conics"]).sum(-1) * m + comps * g["w_comps"] * m).sum()
    grads = torch.autograd.grad(loss, p)
    for got, name in zip(grads, ("v_means", "v_quats", "v_scales", "v_viewmats")):
        _close(got, g[name], rtol=1e-3, atol=1e-3 * float(g[name].abs().max()))


def test_projection_gradients_match_reference(golden):
    g = golden("projection")
    p = [g[k].clone().requires_grad_(True) for k in ("means", "quats", "scales", "viewmats")]
    radii, means2d, depths, conics, comps = O.fully_fused_projection(
        p[0], O.quat_scale_to_covar(p[1], p[2]), p[3], g["Ks"], g["width"], g["height"], calc_compensations=True)
    m = (radii > 0).float()
    loss = ((means2d * g["pinhole_w_means2d"]).sum(-1) * m + depths * g["pinhole_w_depths"] * m
            + (conics * g["pinhole_w_conics"]).sum(-1) * m + comps * g["pinhole_w_comps"] * m).sum()
    grads = torch.autograd.grad(loss, p)
    for got, name in zip(grads, ("v_means", "v_quats", "v_scales", "v_viewmats")):
        ref = g["pinhole_" + name]
        _close(got, ref, rtol=1e-3, atol=1e-3 * float(ref.abs().max()))


def test_sh_matches_reference(golden):
    g = golden("sh")
    for deg in range(5):
        d = g["dirs"].clone().requires_grad_(True)
        c = g["coeffs"].clone().requires_grad_(True)
        col = O.spherical_harmonics(deg, d, c)
        _close(col, g[f"colors_{deg}"], atol=1e-5)
        vd, vc = torch.autograd.grad((col * g["w"]).sum(), (d, c), allow_unused=True)
        _close(vc, g[f"v_coeffs_{deg}"], atol=1e-5)
        if deg > 0:
            _close(vd, g[f"v_dirs_{deg}"], atol=1e-4)


def test_sh_masks_zero_the_output():
    torch.manual_seed(0)
    dirs, coeffs = torch.randn(2, 33, 3), torch.randn(2, 33, 16, 3)
    masks = torch.rand(2, 33) > 0.5
    col = O.spherical_harmonics(3, dirs, coeffs, masks)
    assert (col[~masks] == 0).all() and (col[masks] != 0).any()


def test_isect_matches_reference(golden):
    p, g = golden("projection"), golden("isect")
    tpg, ids, flat = O.isect_tiles(p["pinhole_means2d"], p["pinhole_radii"], p["pinhole_depths"], g["tile_size"],
                                   g["tile_width"], g["tile_height"], bits_rule="cuda")
    assert torch.equal(tpg, g["tiles_per_gauss"])
    assert torch.equal(ids, g["isect_ids"])
    assert torch.equal(flat, g["flatten_ids"])
    offs = O.isect_offset_encode(ids, 3, g["tile_width"], g["tile_height"], bits_rule="cuda")
    assert torch.equal(offs, g["offsets"])
    assert bool((ids[1:] >= ids[:-1]).all())  # sortedness


def test_offsets_docstring_vector():
    # reference triton_impl/isect_offset.py:17-20 / IntersectTile.cu:211-214
    ids = torch.tensor([1, 1, 1, 3, 3, 5, 5], dtype=torch.int64) << 32
    assert O.isect_offset_encode(ids, 1, 7, 1).flatten().tolist() == [0, 0, 3, 3, 5, 5, 7]


def test_tile_bits_rules():
    # the two reference rules differ only at powers of two (SURVEY 8a hazard 3)
    assert O.tile_n_bits(1107, "triton") == O.tile_n_bits(1107, "cuda") == 11
    assert O.tile_n_bits(256, "triton") == 8 and O.tile_n_bits(256, "cuda") == 9


def test_isect_empty_and_packed():
    means2d = torch.tensor([[8.0, 8.0], [100.0, 40.0], [30.0, 30.0]])
    radii = torch.tensor([0, 0, 0], dtype=torch.int32)
    depths = torch.tensor([1.0, 2.0, 3.0])
    tpg, ids, flat = O.isect_tiles(means2d[None], radii[None], depths[None], 16, 8, 4)
    assert ids.numel() == 0 and flat.numel() == 0 and tpg.sum() == 0
    assert O.isect_offset_encode(ids, 1, 8, 4).sum() == 0
    radii = torch.tensor([4, 20, 1], dtype=torch.int32)
    cam = torch.tensor([0, 1, 1])
    tpg_p, ids_p, flat_p = O.isect_tiles(means2d, radii, depths, 16, 8, 4, camera_ids=cam, n_cameras=2)
    assert tpg_p.tolist() == [1, 9, 1]
    tb = O.tile_n_bits(32)
    assert ((ids_p >> (32 + tb)).tolist()) == [0] + [1] * 10


def test_raster_matches_reference(golden):
    g = golden("raster")
    p = [g[k].clone().requires_grad_(True) for k in ("means2d", "conics", "colors", "opacities", "backgrounds")]
    rc, ra, last = O.rasterize_to_pixels(p[0], p[1], p[2], p[3], g["width"], g["height"], g["tile_size"],
                                         g["offsets"], g["flatten_ids"], backgrounds=p[4], return_last_ids=True)
    _close(rc, g["render_colors"], atol=1e-5)
    _close(ra, g["render_alphas"], atol=1e-5)
    grads = torch.autograd.grad((rc * g["w_colors"]).sum() + (ra * g["w_alphas"]).sum(), p)
    for got, name in zip(grads, ("v_means2d", "v_conics", "v_colors", "v_opacities", "v_backgrounds")):
        _close(got, g[name], rtol=1e-3, atol=1e-5)
    assert last.shape == (2, g["height"], g["width"]) and int(last.max()) < g["flatten_ids"].numel()


def test_raster_masks_and_empty():
    # a masked-out tile shows only the background; an empty scene renders background with alpha 0
    offs = torch.zeros((1, 2, 2), dtype=torch.int32)
    flat = torch.zeros((0,), dtype=torch.int32)
    z = torch.zeros
    bg = torch.tensor([[0.25, 0.5, 0.75]])
    rc, ra = O.rasterize_to_pixels(z(1, 1, 2), z(1, 1, 3), z(1, 1, 3), z(1, 1), 20, 18, 16, offs, flat, backgrounds=bg)
    assert rc.shape == (1, 18, 20, 3) and torch.equal(rc[0, 3, 4], bg[0]) and float(ra.abs().max()) == 0.0


def test_whole_path_runs_on_garden_subset():
    from conftest import garden_scene

    means, quats, scales, opac, colors, viewmats, Ks, W, H = garden_scene(step=400)
    rc, ra, meta = O.rasterization(means, quats, scales * 4, opac, colors, viewmats[:1], Ks[:1], W, H)
    assert rc.shape == (1, H, W, 3) and ra.shape == (1, H, W, 1)
    assert float(ra.max()) <= 1.0 and float(ra.max()) > 0.0
    assert int(meta["isect_offsets"][-1, -1, -1]) <= meta["flatten_ids"].numel()


def test_indices_in_range_agree_with_the_compositor(golden):
    """rasterize_to_indices_in_range restatement (RasterizeToIndices3DGS.cu:14-170): the pairs it lists, blended
    with w = alpha * prod(1 - alpha), reproduce the oracle compositor's alpha image; batches partition the list."""
    g = golden("raster")
    C, H, W, ts = g["means2d"].shape[0], g["height"], g["width"], g["tile_size"]
    T0 = torch.ones(C, H, W)
    gs, px, cam = O.rasterize_to_indices_in_range(0, 1 << 20, T0, g["means2d"], g["conics"], g["opacities"], W, H, ts,
                                                  g["offsets"], g["flatten_ids"])
    assert gs.numel() > 1000 and int(gs.max()) < g["means2d"].shape[1]
    key = cam * H * W + px
    assert bool((key[1:] >= key[:-1]).all())  # pixel-major
    d = g["means2d"][cam, gs] - torch.stack([px % W + 0.5, px // W + 0.5], -1)
    cn = g["conics"][cam, gs]
    sigma = 0.5 * (cn[:, 0] * d[:, 0] ** 2 + cn[:, 2] * d[:, 1] ** 2) + cn[:, 1] * d[:, 0] * d[:, 1]
    alpha = torch.clamp_max(g["opacities"][cam, gs] * torch.exp(-sigma), 0.999)
    logT = torch.zeros(C * H * W).index_add_(0, key, torch.log1p(-alpha))
    _close(1.0 - torch.exp(logT).reshape(C, H, W, 1), g["render_alphas"], atol=1e-4)
    first = O.rasterize_to_indices_in_range(0, 1, T0, g["means2d"], g["conics"], g["opacities"], W, H, ts,
                                            g["offsets"], g["flatten_ids"])
    assert 0 < first[0].numel() <= gs.numel()


def test_product_accumulate_reproduces_the_reference_image_and_gradients(golden):
    """gsplat.accumulate (b200_impl/_accumulate.py, the reference's pure-PyTorch compositor _torch_impl.py:432-519
    without nerfacc) on the pair lists of the whole range reproduces the REFERENCE's frozen image, alpha and autograd
    gradients (the golden was made by the reference's own _rasterize_to_pixels + accumulate)."""
    from gsplat.b200_impl import accumulate

    g = golden("raster")
    C, H, W, ts = g["means2d"].shape[0], g["height"], g["width"], g["tile_size"]
    gs, px, cam = O.rasterize_to_indices_in_range(0, 1 << 20, torch.ones(C, H, W), g["means2d"], g["conics"],
                                                  g["opacities"], W, H, ts, g["offsets"], g["flatten_ids"])
    leaves = [g[k].clone().requires_grad_(True) for k in ("means2d", "conics", "opacities", "colors")]
    bg = g["backgrounds"].clone().requires_grad_(True)
    renders, alphas = accumulate(*leaves, gs, px, cam, W, H)
    renders = renders + bg[:, None, None, :] * (1.0 - alphas)
    _close(renders, g["render_colors"], atol=1e-4)
    _close(alphas, g["render_alphas"], atol=1e-4)
    ((renders * g["w_colors"]).sum() + (alphas * g["w_alphas"]).sum()).backward()
    for leaf, name in zip(leaves + [bg], ("v_means2d", "v_conics", "v_opacities", "v_colors", "v_backgrounds")):
        _close(leaf.grad, g[name], rtol=1e-3, atol=1e-4)
    # an empty list composites nothing
    e = torch.zeros(0, dtype=torch.int64)
    r0, a0 = accumulate(g["means2d"], g["conics"], g["opacities"], g["colors"], e, e, e, W, H)
    assert float(r0.abs().sum()) == 0.0 and float(a0.abs().sum()) == 0.0


def test_torch_compositor_loop_batches(monkeypatch):
    """Host logic of b200_impl.rasterize_to_pixels_torch (the reference's _rasterize_to_pixels :522-617): the batch
    loop with the transmittance carried over, here driven by the CPU restatement of the index op, equals the oracle
    compositor for one and for many batches per iteration (the GPU test runs it on the CUDA index kernel)."""
    import sys

    from gsplat.b200_impl import _wrapper, rasterize_to_pixels_torch

    sys.path.insert(0, __import__("os").path.dirname(__file__))
    from test_gpu_raster import _random_splats

    monkeypatch.setattr(_wrapper, "rasterize_to_indices_in_range", O.rasterize_to_indices_in_range)
    C, N, W, H, ts, D = 2, 2500, 70, 50, 16, 3
    means2d, conics, colors, opac, radii, depths = _random_splats(C, N, W, H, D, seed=5)
    tw, th = -(-W // ts), -(-H // ts)
    _, ids, flat = O.isect_tiles(means2d, radii, depths, ts, tw, th)
    offs = O.isect_offset_encode(ids, C, tw, th)
    per_tile = torch.diff(torch.cat([offs.flatten(), torch.tensor([flat.numel()])]))
    assert int(per_tile.max()) > 2 * ts * ts
    backgrounds = torch.rand(C, D)
    # The reference's loop does not carry "this pixel has stopped" from one iteration to the next (the stop rule
    # T (1 - alpha) <= 1e-4 is re-armed with every call of the index op), so it equals the fused compositor when one
    # iteration covers the whole tile (the default, 100 batches) or when no pixel stops (faint scene, any batching).
    for scale, batch_per_iter in ((1.0, 100), (0.02, 1), (0.02, 2)):
        o = opac * scale
        o_rc, o_ra = O.rasterize_to_pixels(means2d, conics, colors, o, W, H, ts, offs, flat, backgrounds=backgrounds)
        rc, ra = rasterize_to_pixels_torch(means2d, conics, colors, o, W, H, ts, offs, flat,
                                           backgrounds=backgrounds, batch_per_iter=batch_per_iter)
        bad = ((rc - o_rc).abs() > 1e-5 + 1e-4 * o_rc.abs()).float().mean().item()
        assert bad < 3e-3, (batch_per_iter, bad)
        bad = ((ra - o_ra).abs() > 1e-5 + 1e-4 * o_ra.abs()).float().mean().item()
        assert bad < 3e-3, (batch_per_iter, bad)


@pytest.mark.parametrize("case", ["sh", "rgb"])
def test_oracle_pipeline_reproduces_the_reference_end_to_end(golden, case):
    """tests/golden/e2e.npz holds the outputs and autograd gradients of the REFERENCE's own pure-torch pipeline
    (`gsplat.rendering._rasterization`, rendering.py:601-760, run by oracle/make_golden.py::case_e2e): SH degree 3 with
    RGB+ED, and plain colours antialiased over a background. The oracle's rasterization() -- what the GPU tests hold
    the CUDA path against -- must reproduce them."""
    g = golden("e2e")
    W, H = int(g["width"]), int(g["height"])
    colors = g["shs"] if case == "sh" else g["rgb"]
    kw = dict(sh_degree=3, render_mode="RGB+ED") if case == "sh" else dict(rasterize_mode="antialiased",
                                                                            backgrounds=g["backgrounds"])
    leaves = [g[k].clone().requires_grad_(True) for k in ("means", "quats", "scales", "opacities")]
    leaves.append(colors.clone().requires_grad_(True))
    rc, ra, _ = O.rasterization(*leaves, g["viewmats"], g["Ks"], W, H, **kw)
    _close(rc, g[f"{case}_render_colors"])
    _close(ra, g[f"{case}_render_alphas"])
    grads = torch.autograd.grad((rc * g[f"{case}_w_colors"]).sum() + (ra * g[f"{case}_w_alphas"]).sum(), leaves)
    for name, got in zip(("means", "quats", "scales", "opacities", "colors"), grads):
        want = g[f"{case}_v_{name}"]
        _close(got, want, rtol=1e-3, atol=1e-4 * float(want.abs().max()) + 1e-6)


# ---------------------------------------------------------------------------------------------
# the C compositor oracle (oracle/gs_compositor.c) and the parity bookkeeping built on it
# ---------------------------------------------------------------------------------------------
def test_c_compositor_matches_the_reference_vectors(golden):
    """oracle/gs_compositor.c against tests/golden/raster.npz (the reference's `_rasterize_to_pixels` + `accumulate`
    and their autograd gradients) and against the torch restatement, last_ids included."""
    import c_oracle as CO

    g = golden("raster")
    names = ("means2d", "conics", "colors", "opacities", "backgrounds")
    p = [g[k].clone().requires_grad_(True) for k in names]
    rc, ra, last, margin = CO.rasterize_to_pixels(p[0], p[1], p[2], p[3], g["width"], g["height"], g["tile_size"],
                                                  g["offsets"], g["flatten_ids"], backgrounds=p[4],
                                                  return_last_ids=True, return_margin=True)
    _close(rc, g["render_colors"], atol=1e-5)
    _close(ra, g["render_alphas"], atol=1e-5)
    grads = torch.autograd.grad((rc * g["w_colors"]).sum() + (ra * g["w_alphas"]).sum(), p)
    for got, name in zip(grads, ("v_means2d", "v_conics", "v_colors", "v_opacities", "v_backgrounds")):
        _close(got, g[name], rtol=1e-3, atol=1e-4 * float(g[name].abs().max()))
    q = [g[k].clone() for k in names]
    _, _, t_last = O.rasterize_to_pixels(q[0], q[1], q[2], q[3], g["width"], g["height"], g["tile_size"], g["offsets"],
                                         g["flatten_ids"], backgrounds=q[4], return_last_ids=True)
    assert torch.equal(last, t_last)
    assert margin.shape == last.shape and float(margin.min()) >= 0.0


@pytest.mark.parametrize("case", ["sh", "rgb"])
def test_c_compositor_pipeline_reproduces_the_reference_end_to_end(golden, case):
    """O.rasterization(compositor="c") against the reference's frozen `_rasterization` outputs and gradients."""
    g = golden("e2e")
    W, H = int(g["width"]), int(g["height"])
    colors = g["shs"] if case == "sh" else g["rgb"]
    kw = dict(sh_degree=3, render_mode="RGB+ED") if case == "sh" else dict(rasterize_mode="antialiased",
                                                                            backgrounds=g["backgrounds"])
    leaves = [g[k].clone().requires_grad_(True) for k in ("means", "quats", "scales", "opacities")]
    leaves.append(colors.clone().requires_grad_(True))
    rc, ra, meta = O.rasterization(*leaves, g["viewmats"], g["Ks"], W, H, compositor="c", bits_rule="cuda", **kw)
    _close(rc, g[f"{case}_render_colors"], rtol=1e-4, atol=1e-4)
    _close(ra, g[f"{case}_render_alphas"], rtol=1e-4, atol=1e-4)
    grads = torch.autograd.grad((rc * g[f"{case}_w_colors"]).sum() + (ra * g[f"{case}_w_alphas"]).sum(), leaves)
    for name, got in zip(("means", "quats", "scales", "opacities", "colors"), grads):
        want = g[f"{case}_v_{name}"]
        _close(got, want, rtol=1e-3, atol=5e-4 * float(want.abs().max()))


def test_c_compositor_margin_flags_a_threshold_pair():
    """One Gaussian whose alpha at a pixel centre is 1/255 up to rounding: the margin of that pixel is ~0, a pixel
    where every decision is far from its threshold reports a large margin."""
    import c_oracle as CO

    o = 0.5
    sigma = float(np.log(255.0 * o))  # alpha == 1/255 exactly at this sigma
    d = float(np.sqrt(2.0 * sigma))  # conic = identity: sigma = d^2 / 2 along x
    means2d = torch.tensor([[[8.5 + d, 8.5]]])
    conics = torch.tensor([[[1.0, 0.0, 1.0]]])
    offs = torch.zeros(1, 1, 1, dtype=torch.int32)
    flat = torch.zeros(1, dtype=torch.int32)
    _, _, _, margin = CO.rasterize_to_pixels(means2d, conics, torch.ones(1, 1, 3), torch.tensor([[o]]), 16, 16, 16,
                                             offs, flat, return_last_ids=True, return_margin=True)
    assert float(margin[0, 8, 8]) < 1e-5
    assert float(margin[0, 8, 11]) > 1e-2


def test_parity_tile_hash_sees_reorders_and_changed_lists():
    from parity import flagged_pixels, tile_list_hash

    offs = torch.tensor([[[0, 3], [3, 5]]], dtype=torch.int32)  # 2x2 tiles
    flat = torch.tensor([4, 7, 9, 1, 2, 8], dtype=torch.int32)
    h = tile_list_hash(flat, offs)
    swapped = flat.clone()
    swapped[0], swapped[1] = flat[1], flat[0]
    h2 = tile_list_hash(swapped, offs)
    assert h2[0] != h[0] and torch.equal(h2[1:], h[1:])
    longer = torch.cat([flat, torch.tensor([3], dtype=torch.int32)])
    h3 = tile_list_hash(longer, offs)
    assert h3[3] != h[3] and torch.equal(h3[:3], h[:3])
    margin = torch.full((1, 20, 20), 1.0)
    margin[0, 0, 19] = 1e-6
    flag, info = flagged_pixels(margin, 20, 20, 16, h, h3)
    assert info["reordered_tiles"] == 1 and bool(flag[0, 16:, 16:].all()) and bool(flag[0, 0, 19])
    assert info["flagged"] == 16 + 1


def test_c_compositor_pipeline_matches_the_reference_at_full_cfg1_size(golden):
    """BASELINE configs[0] at full size (N = 111 785, 648x420): the oracle pipeline on the C compositor against the
    reference's frozen `_rasterization` output (tests/golden/cfg1.npz); every pixel beyond rtol = atol = 1e-4 must be
    a threshold flip (margin < 1e-4), gradients (every 16th row) within 2e-4 of the largest entry."""
    from conftest import garden_scene
    from parity import assert_close_classified, assert_grads_close, flagged_pixels

    g = golden("cfg1")
    means, quats, scales, opac, colors, viewmats, Ks, W, H = garden_scene()
    leaves = [t.clone().requires_grad_(True) for t in (means, quats, scales, opac, colors)]
    assert torch.allclose(torch.stack([t.detach().double().sum() for t in leaves]), g["checksum"], rtol=0, atol=1e-9)
    rc, ra, meta = O.rasterization(*leaves, viewmats[:1], Ks[:1], W, H, compositor="c", bits_rule="cuda")
    assert meta["flatten_ids"].numel() == int(g["n_isects"])
    flag, _ = flagged_pixels(meta["margin"], H, W, 16)
    assert_close_classified(rc, g["render_colors"], flag, 1e-4, 1e-4, "cfg1 oracle colours vs reference")
    assert_close_classified(ra, g["render_alphas"], flag, 1e-4, 1e-4, "cfg1 oracle alphas vs reference")
    gen = torch.Generator().manual_seed(1)
    w_c, w_a = torch.rand(rc.shape, generator=gen), torch.rand(ra.shape, generator=gen)
    grads = torch.autograd.grad((rc * w_c).sum() + (ra * w_a).sum(), leaves)
    for name, got in zip(("means", "quats", "scales", "opacities", "colors"), grads):
        assert_grads_close(got[::16], g["v_" + name], 1e-3, 2e-4, f"cfg1 oracle v_{name} vs reference")
