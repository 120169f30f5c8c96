"""GPU parity tests: SH, tile intersection (bit-exact), compositor fwd/bwd, against golden + oracle."""
import pytest
import torch

import gs_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _close(a, b, rtol=1e-4, atol=1e-4):
    torch.testing.assert_close(a.cpu(), b.cpu(), rtol=rtol, atol=atol)


def _close_frac(a, b, rtol, atol, max_bad=1e-3):
    """Only for comparing two b200 paths with each other (both sides CUDA); comparisons with the oracle / the
    reference's vectors go through tests/parity.py (classified remainder, no blanket allowance)."""
    a, b = a.cpu().double(), b.cpu().double()
    bad = ((a - b).abs() > atol + rtol * b.abs()).double().mean().item()
    assert bad <= max_bad, f"{bad:.2e} of the elements differ beyond rtol={rtol} atol={atol}"


# ---------------------------------------------------------------------------------------------
# spherical harmonics
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("deg", [0, 1, 2, 3, 4])
def test_sh_forward_backward(golden, deg):
    from gsplat.b200_impl import spherical_harmonics

    g = golden("sh")
    d = g["dirs"].to(DEV).requires_grad_(True)
    c = g["coeffs"].to(DEV).requires_grad_(True)
    col = spherical_harmonics(deg, d, c)
    _close(col, g[f"colors_{deg}"], rtol=1e-4, atol=1e-5)
    vd, vc = torch.autograd.grad((col * g["w"].to(DEV)).sum(), (d, c))
    _close(vc, g[f"v_coeffs_{deg}"], rtol=1e-4, atol=1e-5)
    _close(vd, g[f"v_dirs_{deg}"], rtol=1e-3, atol=1e-4)


def test_sh_broadcast_masks_and_shapes():
    from gsplat.b200_impl import spherical_harmonics

    torch.manual_seed(43)
    C, N, K = 3, 1031, 16
    dirs = torch.randn(C, N, 3, device=DEV, requires_grad=True)
    coeffs = torch.randn(N, K, 3, device=DEV, requires_grad=True)
    masks = torch.rand(C, N, device=DEV) > 0.3
    w = torch.randn(C, N, 3, device=DEV)
    outs = []
    for form in ("broadcast", "expand", "materialised"):
        cf = {"broadcast": coeffs, "expand": coeffs.expand(C, -1, -1, -1),
              "materialised": coeffs.expand(C, -1, -1, -1).contiguous()}[form]
        col = spherical_harmonics(3, dirs, cf, masks=masks)
        vd, vc = torch.autograd.grad((col * w).sum(), (dirs, coeffs))
        outs.append((col, vd, vc))
    ref_d = dirs.detach().cpu().requires_grad_(True)
    ref_c = coeffs.detach().cpu().requires_grad_(True)
    ref = O.spherical_harmonics(3, ref_d, ref_c.expand(C, -1, -1, -1), masks.cpu())
    rvd, rvc = torch.autograd.grad((ref * w.cpu()).sum(), (ref_d, ref_c))
    for col, vd, vc in outs:
        _close(col, ref, atol=1e-5)
        _close(vd, rvd, rtol=1e-3, atol=1e-4)
        _close(vc, rvc, rtol=1e-4, atol=1e-4)
        assert float(col[~masks].abs().max()) == 0.0 and float(vd[~masks].abs().max()) == 0.0
    # flat [..., 3] / [..., K, 3] inputs of any rank (triton_tests/test_sh.py shape (129, 31, ...))
    d2 = torch.randn(129, 31, 3, device=DEV)
    c2 = torch.randn(129, 31, 25, 3, device=DEV)
    _close(spherical_harmonics(4, d2, c2), O.spherical_harmonics(4, d2.cpu(), c2.cpu()), atol=1e-5)


def test_fused_view_dependent_colors_equals_the_unfused_glue():
    """b200_impl.view_dependent_colors == clamp_min(spherical_harmonics(means - campos, masks=radii>0) + .5, 0)."""
    from gsplat.b200_impl import spherical_harmonics
    from gsplat.b200_impl._wrapper import view_dependent_colors

    torch.manual_seed(7)
    C, N = 3, 2049
    means = torch.randn(N, 3, device=DEV, requires_grad=True)
    coeffs = (torch.randn(N, 16, 3, device=DEV) * 0.4).requires_grad_(True)
    radii = torch.randint(0, 3, (C, N), dtype=torch.int32, device=DEV)
    A = torch.linalg.qr(torch.randn(C, 3, 3))[0] * (1 + 0.1 * torch.rand(C, 1, 1))  # rotation x scale: affine
    viewmats = torch.eye(4).repeat(C, 1, 1)
    viewmats[:, :3, :3] = A
    viewmats[:, :3, 3] = torch.randn(C, 3) * 3
    viewmats = viewmats.to(DEV)
    w = torch.randn(C, N, 3, device=DEV)
    for deg in (0, 2, 3):
        fused = view_dependent_colors(deg, means, viewmats, coeffs, radii)
        g_f = torch.autograd.grad((fused * w).sum(), (means, coeffs))
        dirs = means[None] - torch.inverse(viewmats)[:, None, :3, 3]
        ref = torch.clamp_min(spherical_harmonics(deg, dirs, coeffs, masks=radii > 0) + 0.5, 0.0)
        g_r = torch.autograd.grad((ref * w).sum(), (means, coeffs))
        _close(fused, ref, rtol=1e-5, atol=1e-5)
        _close(g_f[0], g_r[0], rtol=1e-3, atol=1e-4)
        _close(g_f[1], g_r[1], rtol=1e-4, atol=1e-5)
        assert bool((fused[radii == 0] == 0.5).all())


@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("C,N", [(1, 1), (1, 37), (3, 128), (1, 4099), (2, 4099), (1, 300_001), (3, 100_003)])
def test_staged_colour_kernels_equal_the_plain_ones(mode, C, N):
    """gsb_sh_colors_fwd / _bwd with asynchronous staging of the coefficient rows (GSB_OPT_SH_STAGING 1 = cp.async.bulk
    + mbarrier, 2 = cp.async; persistent CTAs, two stages) against the plain-load kernels (0) on the same inputs,
    through the C ABI: same arithmetic in the same order. Sizes: a single row, partial warps / tiles, more tiles than
    resident CTAs (several tiles per CTA, both stages reused), one camera (outer-product path) and several."""
    from gsplat.b200_impl._lib import lib, ptr, stream_of

    OPTS = (3, 6)  # GSB_OPT_SH_STAGING (backward kernel), GSB_OPT_SH_FWD_STAGING (forward kernel)
    g = torch.Generator().manual_seed(100 * C + N % 97)
    means = torch.randn(N, 3, generator=g).to(DEV)
    coeffs = (torch.randn(N, 16, 3, generator=g) * 0.4).to(DEV)
    radii = torch.randint(0, 3, (C, N), generator=g, dtype=torch.int32).to(DEV)
    campos = (torch.randn(C, 3, generator=g) * 3).to(DEV)
    v_colors = torch.randn(C, N, 3, generator=g).to(DEV)

    def run(option, need_means):
        for opt in OPTS:
            lib.gsb_set_option(opt, option)
        colors = torch.full((C, N, 3), float("nan"), device=DEV)
        v_coeffs = torch.full((N, 16, 3), float("nan"), device=DEV)
        v_means = torch.full((N, 3), float("nan"), device=DEV) if need_means else None
        st = stream_of(means)
        lib.gsb_sh_colors_fwd(3, N, C, 16, ptr(means), ptr(campos), ptr(coeffs), ptr(radii), ptr(colors), st)
        lib.gsb_sh_colors_bwd(3, N, C, 16, ptr(means), ptr(campos), ptr(coeffs), ptr(radii), ptr(v_colors),
                              ptr(v_coeffs), ptr(v_means), st)
        torch.cuda.synchronize()
        return colors, v_coeffs, v_means

    defaults = [lib.gsb_get_option(opt) for opt in OPTS]
    try:
        for need_means in (True, False):
            ref = run(0, need_means)
            out = run(mode, need_means)
            for name, a, b in zip(("colors", "v_coeffs", "v_means"), out, ref):
                if b is None:
                    assert a is None
                    continue
                assert bool(torch.isfinite(a).all()), f"{name}: rows left unwritten"
                torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6, msg=lambda m: f"{name}: {m}")
    finally:
        for opt, value in zip(OPTS, defaults):
            lib.gsb_set_option(opt, value)


# ---------------------------------------------------------------------------------------------
# isect_tiles / isect_offset_encode: bit-exact given the reference's projection outputs
# ---------------------------------------------------------------------------------------------
def test_isect_bit_exact_against_reference(golden, monkeypatch):
    from gsplat.b200_impl import isect_offset_encode, isect_tiles

    monkeypatch.setenv("GSPLAT_B200_TILE_BITS", "cuda")  # the golden was made with _torch_impl's rule
    p, g = golden("projection"), golden("isect")
    means2d, radii, depths = (p[k].to(DEV) for k in ("pinhole_means2d", "pinhole_radii", "pinhole_depths"))
    tpg, ids, flat = isect_tiles(means2d, radii, depths, g["tile_size"], g["tile_width"], g["tile_height"])
    assert tpg.dtype == torch.int32 and ids.dtype == torch.int64 and flat.dtype == torch.int32
    assert torch.equal(tpg.cpu(), g["tiles_per_gauss"])
    assert torch.equal(ids.cpu(), g["isect_ids"])
    assert torch.equal(flat.cpu(), g["flatten_ids"])
    offs = isect_offset_encode(ids, 3, g["tile_width"], g["tile_height"])
    assert offs.dtype == torch.int32 and torch.equal(offs.cpu(), g["offsets"])


@pytest.mark.parametrize("C,N,W,H", [(3, 1000, 40, 60), (1, 5000, 256, 256), (2, 1, 33, 17), (4, 3000, 640, 360)])
def test_isect_random_against_oracle(C, N, W, H):
    # tests/test_basic.py:410 protocol (random data); 256x256 -> 256 tiles is the power-of-two bit-rule case
    from gsplat.b200_impl import isect_offset_encode, isect_tiles

    torch.manual_seed(50)
    ts = 16
    tw, th = -(-W // ts), -(-H // ts)
    means2d = torch.rand(C, N, 2) * torch.tensor([W + 40.0, H + 40.0]) - 20.0
    radii = torch.randint(0, 40, (C, N), dtype=torch.int32)
    depths = torch.rand(C, N) * 10 + 0.01
    depths[:, ::5] = 1.0  # equal depths: ties must keep emission order (stable sort)
    tpg, ids, flat = isect_tiles(means2d.to(DEV), radii.to(DEV), depths.to(DEV), ts, tw, th)
    o_tpg, o_ids, o_flat = O.isect_tiles(means2d, radii, depths, ts, tw, th)
    assert torch.equal(tpg.cpu(), o_tpg) and torch.equal(ids.cpu(), o_ids) and torch.equal(flat.cpu(), o_flat)
    offs = isect_offset_encode(ids, C, tw, th)
    assert torch.equal(offs.cpu(), O.isect_offset_encode(o_ids, C, tw, th))
    # unsorted emission order
    _, ids_u, flat_u = isect_tiles(means2d.to(DEV), radii.to(DEV), depths.to(DEV), ts, tw, th, sort=False)
    _, o_ids_u, o_flat_u = O.isect_tiles(means2d, radii, depths, ts, tw, th, sort=False)
    assert torch.equal(ids_u.cpu(), o_ids_u) and torch.equal(flat_u.cpu(), o_flat_u)


@pytest.mark.parametrize("N", [1, 3, 5000, 250_001])
def test_isect_finalize_variants_are_identical(N):
    """gsb_isect_finalize with four intersections per thread (GSB_OPT_FINALIZE_VEC 1) against one per thread
    (0, default): ids, row ids and offsets bit-identical, also when the count is not a multiple of four."""
    from gsplat.b200_impl import isect_offset_encode, isect_tiles
    from gsplat.b200_impl._lib import lib

    OPT = 4  # GSB_OPT_FINALIZE_VEC
    g = torch.Generator().manual_seed(N)
    C, W, H, ts = 2, 200, 120, 16
    tw, th = -(-W // ts), -(-H // ts)
    means2d = (torch.rand(C, N, 2, generator=g) * torch.tensor([W * 0.7, H * 1.0])).to(DEV)  # some tiles stay empty
    radii = torch.randint(0, 20, (C, N), generator=g, dtype=torch.int32).to(DEV)
    depths = (torch.rand(C, N, generator=g) * 5 + 0.1).to(DEV)
    default = lib.gsb_get_option(OPT)
    out = {}
    try:
        for opt in (0, 1):
            lib.gsb_set_option(OPT, opt)
            tpg, ids, flat = isect_tiles(means2d, radii, depths, ts, tw, th)
            out[opt] = (tpg, ids, flat, isect_offset_encode(ids, C, tw, th))
    finally:
        lib.gsb_set_option(OPT, default)
    for a, b in zip(out[0], out[1]):
        assert torch.equal(a, b)
    o_tpg, o_ids, o_flat = O.isect_tiles(means2d.cpu(), radii.cpu(), depths.cpu(), ts, tw, th)
    assert torch.equal(out[1][1].cpu(), o_ids) and torch.equal(out[1][2].cpu(), o_flat)
    assert torch.equal(out[1][3].cpu(), O.isect_offset_encode(o_ids, C, tw, th))


@pytest.mark.parametrize("N,packed", [(3000, False), (30000, True), (400000, False)])
def test_isect_presorted_path_equals_global_radix_sort(N, packed, monkeypatch):
    """The depth-presorted path (default: row sort + 2-pass tile sort) against the global stable radix sort of
    the 64-bit keys: equal depths (ties), empty tiles, very full tiles, packed rows."""
    from gsplat.b200_impl import isect_offset_encode, isect_tiles

    torch.manual_seed(N)
    C, W, H, ts = 3, 96, 80, 16
    tw, th = 6, 5
    shape = (C * N,) if packed else (C, N)
    means2d = (torch.rand(*shape, 2) * torch.tensor([W * 0.6, H * 1.0])).to(DEV)  # right-hand tiles stay empty-ish
    radii = torch.randint(0, 12, shape, dtype=torch.int32).to(DEV)
    depths = (torch.rand(*shape) * 5 + 0.1).to(DEV)
    depths[..., ::3] = 2.5
    kw = {}
    if packed:
        cam = torch.sort(torch.randint(0, C, shape)).values.to(DEV)
        kw = dict(packed=True, n_cameras=C, camera_ids=cam, gaussian_ids=torch.arange(C * N, device=DEV))
    out = {}
    for mode, binning in (("presort", "bin"), ("presort", "radix"), ("global", "radix")):
        monkeypatch.setenv("GSPLAT_B200_SORT", mode)
        monkeypatch.setenv("GSPLAT_B200_ISECT", binning)
        tpg, ids, flat = isect_tiles(means2d, radii, depths, ts, tw, th, **kw)
        out[mode, binning] = (tpg, ids, flat, isect_offset_encode(ids, C, tw, th))
    for other in (("presort", "radix"), ("presort", "bin")):
        for a, b in zip(out[other], out["global", "radix"]):
            assert torch.equal(a, b)
    assert out["presort", "bin"][1].numel() > N


@pytest.mark.parametrize("C,N,W,H,rmax,spread", [(1, 60000, 1920, 1080, 300, 1.0), (2, 20000, 1000, 1000, 900, 1.0),
                                                 (1, 300, 1920, 1080, 2500, 1.0), (1, 200000, 648, 420, 30, 1.0),
                                                 (1, 100000, 1920, 1080, 12, 0.03)])
def test_isect_tile_binning_equals_radix_path(C, N, W, H, rmax, spread, monkeypatch):
    """csrc/isect_bin.cu (counting sort, one bin per tile) against the two-pass radix path: rows covering thousands of
    tiles (a single row larger than a chunk of 8192 intersections), ties, rows without tiles, two cameras, everything
    piled on a few tiles (spread 0.03: per-bin lists beyond the counting rank -> the kernel's radix fallback), the
    fallback forced for every chunk; and through isect_offset_encode recomputed from the ids."""
    from gsplat.b200_impl import isect_offset_encode, isect_tiles
    from gsplat.b200_impl._lib import lib

    torch.manual_seed(N + rmax)
    ts = 16
    tw, th = -(-W // ts), -(-H // ts)
    means2d = ((torch.rand(C, N, 2) * spread + (1 - spread) / 2) * torch.tensor([W + 200.0, H + 200.0]) - 100.0).to(DEV)
    radii = (torch.rand(C, N) ** 4 * rmax).int().to(DEV)  # mostly small, a few huge; zeros included
    depths = (torch.rand(C, N) * 20 + 0.05).to(DEV)
    depths[:, ::7] = 3.0
    out = {}
    for binning in ("bin", "bin_all_radix", "radix", "radix_search"):
        monkeypatch.setenv("GSPLAT_B200_ISECT", binning.split("_")[0])
        # the radix path's emission: chunked (marker max-scan) or the default search per slot
        monkeypatch.setenv("GSPLAT_B200_EMIT", "search" if binning == "radix_search" else "chunks")
        lib.gsb_set_option(2, int(binning == "bin_all_radix"))  # GSB_OPT_ISECT_BIN_RADIX
        try:
            tpg, ids, flat = isect_tiles(means2d, radii, depths, ts, tw, th)
        finally:
            lib.gsb_set_option(2, 0)
        cached = isect_offset_encode(ids, C, tw, th)
        fresh = isect_offset_encode(ids.clone(), C, tw, th)  # a clone carries no cached offsets: recomputed from ids
        assert torch.equal(cached, fresh)
        out[binning] = (tpg, ids, flat, cached)
    for other in ("bin", "bin_all_radix", "radix"):
        for a, b in zip(out[other], out["radix_search"]):
            assert torch.equal(a, b)
    assert out["bin"][1].numel() > 8192


def test_isect_packed_empty_and_docstring():
    from gsplat.b200_impl import isect_offset_encode, isect_tiles

    torch.manual_seed(1)
    nnz, C, ts, tw, th = 777, 3, 16, 9, 7
    cam = torch.sort(torch.randint(0, C, (nnz,))).values
    gid = torch.arange(nnz)
    means2d = torch.rand(nnz, 2) * 140
    radii = torch.randint(1, 30, (nnz,), dtype=torch.int32)
    depths = torch.rand(nnz) + 0.5
    tpg, ids, flat = isect_tiles(means2d.to(DEV), radii.to(DEV), depths.to(DEV), ts, tw, th, packed=True,
                                 n_cameras=C, camera_ids=cam.to(DEV), gaussian_ids=gid.to(DEV))
    o = O.isect_tiles(means2d, radii, depths, ts, tw, th, camera_ids=cam, n_cameras=C)
    assert torch.equal(tpg.cpu(), o[0]) and torch.equal(ids.cpu(), o[1]) and torch.equal(flat.cpu(), o[2])
    # empty scene
    z = torch.zeros(1, 8, dtype=torch.int32, device=DEV)
    tpg, ids, flat = isect_tiles(torch.zeros(1, 8, 2, device=DEV), z, torch.ones(1, 8, device=DEV), ts, tw, th)
    assert ids.numel() == 0 and flat.numel() == 0 and int(tpg.sum()) == 0
    assert int(isect_offset_encode(ids, 1, tw, th).abs().sum()) == 0
    # reference docstring vector (isect_offset.py:17-20)
    doc = (torch.tensor([1, 1, 1, 3, 3, 5, 5], dtype=torch.int64) << 32).to(DEV)
    assert isect_offset_encode(doc, 1, 7, 1).flatten().tolist() == [0, 0, 3, 3, 5, 5, 7]
    # a single intersection: trailing tiles must still be filled (Triton reference defect)
    one = (torch.tensor([2], dtype=torch.int64) << 32).to(DEV)
    assert isect_offset_encode(one, 1, 5, 1).flatten().tolist() == [0, 0, 0, 1, 1]


# ---------------------------------------------------------------------------------------------
# compositor
# ---------------------------------------------------------------------------------------------
def test_raster_against_reference_golden(golden):
    from gsplat.b200_impl import rasterize_to_pixels

    from parity import REFERENCE_STAGE_TOL, check_compositor_stage

    g = golden("raster")
    # (1) against the C oracle (pinned to these very vectors on the CPU): classified image, masked-loss gradients
    (rc, ra), p, flag = check_compositor_stage(
        rasterize_to_pixels, g["means2d"], g["conics"], g["colors"], g["opacities"], g["width"], g["height"],
        g["tile_size"], g["offsets"], g["flatten_ids"], backgrounds=g["backgrounds"], w_colors=g["w_colors"],
        w_alphas=g["w_alphas"], what="raster golden")
    # (2) directly against the REFERENCE's frozen outputs and autograd gradients: the reference's own tolerances
    # (triton_tests/test_ras2pix.py:132-161), zero outliers
    from parity import assert_close_classified

    assert_close_classified(rc, g["render_colors"], flag, 1e-4, 1e-5, "raster golden colours vs reference")
    assert_close_classified(ra, g["render_alphas"], flag, 1e-4, 1e-5, "raster golden alphas vs reference")
    grads = torch.autograd.grad((rc * g["w_colors"].to(DEV)).sum() + (ra * g["w_alphas"].to(DEV)).sum(), p)
    for got, name in zip(grads, ("v_means2d", "v_conics", "v_colors", "v_opacities", "v_backgrounds")):
        rtol, atol = REFERENCE_STAGE_TOL[name]
        torch.testing.assert_close(got.cpu(), g[name], rtol=rtol, atol=atol * max(1.0, float(g[name].abs().max())))


def _random_splats(C, N, W, H, D, seed):
    g = torch.Generator().manual_seed(seed)
    means2d = torch.rand(C, N, 2, generator=g) * torch.tensor([W + 8.0, H + 8.0]) - 4.0
    s = torch.exp(torch.rand(C, N, 2, generator=g) * 2.0 - 0.2)
    rho = torch.rand(C, N, generator=g) * 1.6 - 0.8
    cxx, cyy, cxy = s[..., 0] ** 2, s[..., 1] ** 2, rho * s[..., 0] * s[..., 1]
    det = cxx * cyy - cxy * cxy
    conics = torch.stack([cyy / det, -cxy / det, cxx / det], dim=-1)
    radii = torch.ceil(3.0 * torch.sqrt(torch.maximum(cxx, cyy))).int()
    depths = torch.rand(C, N, generator=g) * 9.0 + 1.0
    opacities = torch.rand(C, N, generator=g) * 0.98 + 0.01
    opacities[:, ::7] = 1.0
    colors = torch.rand(C, N, D, generator=g)
    return means2d, conics, colors, opacities, radii, depths


@pytest.mark.parametrize("D,bg,absgrad", [(1, False, False), (3, True, True), (4, False, False), (5, True, False),
                                          (8, False, True), (17, True, False), (32, False, False), (40, True, False)])
def test_raster_channels_against_oracle(D, bg, absgrad):
    from gsplat.b200_impl import rasterize_to_pixels

    C, N, W, H, ts = 2, 600, 70, 50, 16
    means2d, conics, colors, opac, radii, depths = _random_splats(C, N, W, H, D, seed=D)
    tw, th = -(-W // ts), -(-H // ts)
    _, ids, flat = O.isect_tiles(means2d, radii, depths, ts, tw, th)
    offs = O.isect_offset_encode(ids, C, tw, th)
    backgrounds = torch.rand(C, D) if bg else None
    from parity import check_compositor_stage

    (rc, ra), p, _ = check_compositor_stage(rasterize_to_pixels, means2d, conics, colors, opac, W, H, ts, offs, flat,
                                            backgrounds=backgrounds, what=f"D={D}", gpu_kw={"absgrad": absgrad})
    if absgrad:
        import c_oracle as CO

        gw = torch.Generator().manual_seed(99)
        w_c, w_a = torch.randn(C, H, W, D, generator=gw), torch.randn(C, H, W, 1, generator=gw)
        grads = torch.autograd.grad((rc * w_c.to(DEV)).sum() + (ra * w_a.to(DEV)).sum(), p[:1])
        ab = p[0].absgrad
        assert ab.shape == p[0].shape and bool((ab + 1e-6 >= grads[0].abs()).all())
        # absgrad against the C oracle's sum of |v_xy| over pixels (RasterizeToPixels3DGSBwd.cu:226-229)
        o_rc, o_ra, o_last, _ = CO.rasterize_fwd(means2d, conics, colors, opac, W, H, ts, offs, flat, backgrounds)
        o = CO.rasterize_bwd(means2d, conics, colors, opac, W, H, ts, offs, flat, o_ra, o_last, w_c, w_a,
                             backgrounds=backgrounds, absgrad=True)
        want = o["v_means2d_abs"].reshape(ab.shape)
        bad = (ab.cpu() - want).abs() > 2e-3 * want.abs() + 2e-3 * float(want.abs().max())
        assert int(bad.sum()) == 0, int(bad.sum())


@pytest.mark.parametrize("D,bg", [(1, False), (3, True), (4, True)])
def test_forward_kernel_variants_are_bit_identical(D, bg):
    """rasterize_fwd2.cu (half-warp per block, two pixels per lane, packed FFMA2) performs the same arithmetic per
    (pixel, Gaussian) pair as rasterize_fwd.cu: images, alphas and the gradients' inputs (last ids through the
    backward pass) must be identical bit for bit, also on partial tiles and with masks."""
    from gsplat.b200_impl import rasterize_to_pixels
    from gsplat.b200_impl._lib import lib

    C, N, W, H, ts = 2, 4000, 200, 117, 16
    means2d, conics, colors, opac, radii, depths = _random_splats(C, N, W, H, D, seed=11 + D)
    tw, th = -(-W // ts), -(-H // ts)
    _, ids, flat = O.isect_tiles(means2d, radii, depths, ts, tw, th)
    offs = O.isect_offset_encode(ids, C, tw, th)
    masks = torch.ones(C, th, tw, dtype=torch.bool)
    masks[1, 2, 3] = False
    backgrounds = torch.rand(C, D) if bg else None
    g = torch.Generator().manual_seed(5)
    w_c, w_a = torch.randn(C, H, W, D, generator=g).to(DEV), torch.randn(C, H, W, 1, generator=g).to(DEV)
    outs = {}
    try:
        for variant in (1, 2, 3):
            assert lib.gsb_set_option(0, variant) == 0
            p = [t.clone().to(DEV).requires_grad_(True) for t in (means2d, conics, colors, opac)]
            rc, ra = rasterize_to_pixels(p[0], p[1], p[2], p[3], W, H, ts, offs.to(DEV), flat.to(DEV),
                                         backgrounds=backgrounds.to(DEV) if bg else None, masks=masks.to(DEV))
            grads = torch.autograd.grad((rc * w_c).sum() + (ra * w_a).sum(), p)
            outs[variant] = (rc, ra, grads)
    finally:
        lib.gsb_set_option(0, 2)
    for variant in (2, 3):
        assert torch.equal(outs[1][0], outs[variant][0]) and torch.equal(outs[1][1], outs[variant][1]), variant
        for a, b in zip(outs[1][2], outs[variant][2]):  # same kernel, same inputs: only the atomics' order differs
            torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-5 * float(b.abs().max()))


@pytest.mark.parametrize("ts,D,bg,absgrad", [(4, 3, True, False), (4, 32, False, False), (8, 3, True, True),
                                                (12, 4, False, False), (32, 3, True, False)])
def test_raster_other_tile_sizes_against_oracle(ts, D, bg, absgrad):
    """`tile_size` is a public argument of rasterization(); the reference's own stage tests run tile_size = 4 with
    wide channels (triton_tests/test_ras2pix.py:88, test_isect.py:72). Every size other than 16 goes through
    csrc/rasterize_generic.cu; same classified comparison as the tuned path."""
    from gsplat.b200_impl import rasterize_to_pixels
    from parity import check_compositor_stage

    C, N, W, H = 2, 500, 70, 50
    means2d, conics, colors, opac, radii, depths = _random_splats(C, N, W, H, D, seed=100 + ts + D)
    tw, th = -(-W // ts), -(-H // ts)
    _, ids, flat = O.isect_tiles(means2d, radii, depths, ts, tw, th)
    offs = O.isect_offset_encode(ids, C, tw, th)
    backgrounds = torch.rand(C, D) if bg else None
    (rc, ra), p, _ = check_compositor_stage(rasterize_to_pixels, means2d, conics, colors, opac, W, H, ts, offs, flat,
                                            backgrounds=backgrounds, what=f"tile {ts} D={D}",
                                            gpu_kw={"absgrad": absgrad})
    if absgrad:
        torch.autograd.grad(rc.sum() + ra.sum(), p[:1])
        assert p[0].absgrad.shape == p[0].shape and bool(torch.isfinite(p[0].absgrad).all())


@pytest.mark.parametrize("D,bg,absgrad", [(1, False, True), (3, True, False), (3, False, True), (4, True, True)])
def test_backward_kernel_variants_agree_and_match_oracle(D, bg, absgrad):
    """rasterize_bwd2.cu (half-warp per block, two pixels per lane, packed FFMA2) against rasterize_bwd.cu (same
    per-pair values, different summation order) and, on its own, against the C oracle with the classified
    comparison; partial tiles, masks, absgrad."""
    from gsplat.b200_impl import rasterize_to_pixels
    from gsplat.b200_impl._lib import lib
    from parity import check_compositor_stage

    C, N, W, H, ts = 2, 4000, 200, 117, 16
    means2d, conics, colors, opac, radii, depths = _random_splats(C, N, W, H, D, seed=31 + D)
    tw, th = -(-W // ts), -(-H // ts)
    _, ids, flat = O.isect_tiles(means2d, radii, depths, ts, tw, th)
    offs = O.isect_offset_encode(ids, C, tw, th)
    masks = torch.ones(C, th, tw, dtype=torch.bool)
    masks[0, 1, 2] = False
    backgrounds = torch.rand(C, D) if bg else None
    g = torch.Generator().manual_seed(6)
    w_c, w_a = torch.randn(C, H, W, D, generator=g).to(DEV), torch.randn(C, H, W, 1, generator=g).to(DEV)
    outs = {}
    try:
        for variant in (1, 2):
            assert lib.gsb_set_option(1, variant) == 0
            p = [t.clone().to(DEV).requires_grad_(True) for t in (means2d, conics, colors, opac)]
            b = backgrounds.clone().to(DEV).requires_grad_(True) if bg else None
            rc, ra = rasterize_to_pixels(p[0], p[1], p[2], p[3], W, H, ts, offs.to(DEV), flat.to(DEV), backgrounds=b,
                                         masks=masks.to(DEV), absgrad=absgrad)
            grads = torch.autograd.grad((rc * w_c).sum() + (ra * w_a).sum(), p + ([b] if bg else []))
            outs[variant] = list(grads) + ([p[0].absgrad] if absgrad else [])
        for a, b_ in zip(outs[1], outs[2]):
            torch.testing.assert_close(a, b_, rtol=1e-4, atol=2e-5 * float(b_.abs().max()))
        check_compositor_stage(rasterize_to_pixels, means2d, conics, colors, opac, W, H, ts, offs, flat,
                               backgrounds=backgrounds, masks=masks, what=f"bwd v2 D={D}", gpu_kw={"absgrad": absgrad})
    finally:
        lib.gsb_set_option(1, 2)


def test_raster_masks_empty_and_last_tile():
    from gsplat.b200_impl import rasterize_to_pixels

    C, N, W, H, ts, D = 1, 300, 48, 48, 16, 3
    means2d, conics, colors, opac, radii, depths = _random_splats(C, N, W, H, D, seed=3)
    tw = th = 3
    _, ids, flat = O.isect_tiles(means2d, radii, depths, ts, tw, th)
    offs = O.isect_offset_encode(ids, C, tw, th)
    masks = torch.ones(C, th, tw, dtype=torch.bool)
    masks[0, 1, 1] = False
    bg = torch.rand(C, D)
    args = (means2d, conics, colors, opac, W, H, ts, offs, flat)
    rc, ra = rasterize_to_pixels(*[a.to(DEV) if torch.is_tensor(a) else a for a in args], backgrounds=bg.to(DEV),
                                 masks=masks.to(DEV))
    from parity import check_compositor_stage

    check_compositor_stage(rasterize_to_pixels, means2d, conics, colors, opac, W, H, ts, offs, flat, backgrounds=bg,
                           masks=masks, what="masked tile")
    assert torch.equal(rc[0, 20, 20].cpu(), bg[0]) and float(ra[0, 16:32, 16:32].abs().max()) == 0.0
    # no intersections at all
    e = torch.zeros((0,), dtype=torch.int32, device=DEV)
    rc, ra = rasterize_to_pixels(means2d.to(DEV), conics.to(DEV), colors.to(DEV), opac.to(DEV), W, H, ts,
                                 torch.zeros_like(offs).to(DEV), e, backgrounds=bg.to(DEV))
    assert float(ra.abs().max()) == 0.0 and torch.equal(rc[0, 0, 0].cpu(), bg[0])


def test_rasterize_to_indices_in_range_matches_oracle(golden):
    """b200 rasterize_to_indices_in_range (cuda/_wrapper.py:577) vs the CPU restatement, whole range and in two
    steps with the transmittance carried over, per-pixel counts and the lists themselves."""
    from gsplat.b200_impl import rasterize_to_indices_in_range

    g = golden("raster")
    C, H, W, ts = g["means2d"].shape[0], g["height"], g["width"], g["tile_size"]
    dev = "cuda"
    args = [g[k].to(dev) for k in ("means2d", "conics", "opacities")]
    tail = (W, H, ts, g["offsets"].to(dev), g["flatten_ids"].to(dev))
    o_tail = (W, H, ts, g["offsets"], g["flatten_ids"])
    T0 = torch.ones(C, H, W)
    for lo, hi, T in ((0, 1 << 30, T0), (0, 1, T0), (1, 3, T0 * 0.5)):
        gs, px, cam = rasterize_to_indices_in_range(lo, hi, T.to(dev), *args, *tail)
        o_gs, o_px, o_cam = O.rasterize_to_indices_in_range(lo, hi, T, g["means2d"], g["conics"], g["opacities"],
                                                            *o_tail)
        assert gs.dtype == torch.int64 and px.dtype == torch.int64 and cam.dtype == torch.int64
        key, o_key = (cam * H * W + px).cpu(), o_cam * H * W + o_px
        cnt = torch.bincount(key, minlength=C * H * W)
        o_cnt = torch.bincount(o_key, minlength=C * H * W)
        # exp() differs by an ulp between the two sides: a pair sitting on the 1/255 or 1e-4 threshold may flip
        assert float((cnt != o_cnt).float().mean()) < 5e-3, (lo, hi)
        assert abs(gs.numel() - o_gs.numel()) <= 5e-3 * max(o_gs.numel(), 1) + 2
        same = (cnt == o_cnt)
        keep, o_keep = same[key], same[o_key]
        assert torch.equal(gs.cpu()[keep], o_gs[o_keep]) and torch.equal(key[keep], o_key[o_keep]), (lo, hi)
    empty = rasterize_to_indices_in_range(5000, 5001, T0.to(dev), *args, *tail)
    assert all(t.numel() == 0 for t in empty)


def test_torch_compositor_over_our_index_kernel_equals_the_fused_compositor():
    """The reference's pure-PyTorch compositor (_torch_impl._rasterize_to_pixels :522-617 = batches of
    rasterize_to_indices_in_range + accumulate, autograd gradients), restated without nerfacc on top of OUR index
    kernel, against the fused b200 compositor: several batches per tile, one and many batches per iteration."""
    from gsplat.b200_impl import rasterize_to_pixels, rasterize_to_pixels_torch

    C, N, W, H, ts, D = 2, 2500, 70, 50, 16, 3
    means2d, conics, colors, opac, radii, depths = _random_splats(C, N, W, H, D, seed=5)
    tw, th = -(-W // ts), -(-H // ts)
    _, ids, flat = O.isect_tiles(means2d, radii, depths, ts, tw, th)
    offs = O.isect_offset_encode(ids, C, tw, th)
    per_tile = torch.diff(torch.cat([offs.flatten(), torch.tensor([flat.numel()])]))
    assert int(per_tile.max()) > 2 * ts * ts  # more than two batches somewhere
    backgrounds = torch.rand(C, D)
    gw = torch.Generator().manual_seed(7)
    w_c, w_a = torch.randn(C, H, W, D, generator=gw).to(DEV), torch.randn(C, H, W, 1, generator=gw).to(DEV)

    # The reference's loop re-arms the stop rule T (1 - alpha) <= 1e-4 with every call of the index op, so it equals
    # the fused compositor when one iteration covers the whole tile (default: 100 batches) or when no pixel stops
    # (faint scene, any batching) -- tests/test_oracle.py::test_torch_compositor_loop_batches shows the same on the CPU.
    for scale, batch_per_iter in ((1.0, 100), (0.02, 1)):
        opac_s = opac * scale

        def run(fn, **kw):
            p = [t.clone().to(DEV).requires_grad_(True) for t in (means2d, conics, colors, opac_s)]
            b = backgrounds.clone().to(DEV).requires_grad_(True)
            rc, ra = fn(p[0], p[1], p[2], p[3], W, H, ts, offs.to(DEV), flat.to(DEV), backgrounds=b, **kw)
            return rc, ra, torch.autograd.grad((rc * w_c).sum() + (ra * w_a).sum(), p + [b])

        rc, ra, grads = run(rasterize_to_pixels)
        t_rc, t_ra, t_grads = run(rasterize_to_pixels_torch, batch_per_iter=batch_per_iter)
        # two CUDA paths of ours (index kernel + torch accumulate vs the fused compositor): the exp of the index
        # kernel (expf) and of the compositor (ex2.approx) differ in the last ulps, so pairs on a threshold may flip;
        # the fraction is bounded here and classified pixel by pixel where it matters (the fused path vs the oracle)
        _close_frac(t_rc, rc, rtol=1e-4, atol=1e-5, max_bad=3e-3)
        _close_frac(t_ra, ra, rtol=1e-4, atol=1e-5, max_bad=3e-3)
        for a, b in zip(t_grads, grads):
            _close_frac(a, b, rtol=1e-3, atol=1e-5 * float(b.abs().max()) + 1e-6, max_bad=5e-3)
