"""CPU oracle for the rasterization() hot path — TEST INFRASTRUCTURE, NOT A PRODUCT PATH.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may
import this module; nothing under gsplat-triton_b200/ does (the product fails loudly without
its CUDA library instead of falling back here).

Each function restates, in plain float32 torch/numpy on the CPU, the reference algorithm it
cites (paths relative to the reference checkout, hieu1999210/gsplat-triton @ 46beece):

  quat_scale_to_covar      gsplat/cuda/_torch_impl.py:10-75   (_quat_to_rotmat, _quat_scale_to_covar_preci)
  fully_fused_projection   gsplat/cuda/_torch_impl.py:78-122 (_persp_proj), :123-182 (_fisheye_proj), :203-216 (_ortho_proj),
                           :219-247 (_world_to_cam), :250-328 (_fully_fused_projection),
                           + radius_clip from triton_impl/fused_projection_fwd.py:215-221
  sh_bases / spherical_harmonics   gsplat/cuda/_torch_impl.py:620-714
  isect_tiles              gsplat/cuda/_torch_impl.py:331-400 (vectorised; same emission order)
  isect_offset_encode      gsplat/cuda/_torch_impl.py:403-429
  rasterize_to_pixels      math: gsplat/cuda/_torch_impl.py:490-501 (accumulate);
                           ordering / thresholds / termination / last_ids:
                           gsplat/cuda/csrc/RasterizeToPixels3DGSFwd.cu:105-187
                           (== triton_impl/rasterize_to_pixels_fwd.py:105-166)

Pinning status
  projection, SH, isect, offsets: PINNED — oracle/make_golden.py runs the reference's own
  `_torch_impl` functions (imported from /root/reference) against these restatements and
  freezes the reference outputs under tests/golden/ (tests/test_oracle.py re-checks them).
  rasterize_to_pixels: the reference's `_rasterize_to_pixels` cannot run anywhere without the
  third-party `nerfacc` package (absent, un-vendored, not in setup.py) and the CUDA op
  `rasterize_to_indices_3dgs`; make_golden.py pins the alpha/weight math against the
  reference's own `accumulate()` (:432-519) driven through restated nerfacc primitives and a
  restated index enumeration. So: compositor parity is pinned to `accumulate()`, and
  "PARITY UNPINNED" at the nerfacc boundary (render_weight_from_alpha / accumulate_along_rays
  are restated from their published definition w_i = alpha_i * prod_{j<i}(1 - alpha_j)).
"""
import math
from typing import Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F
from torch import Tensor

ALPHA_MIN = 1.0 / 255.0
ALPHA_MAX = 0.999
T_MIN = 1e-4


# ------------------------------------------------------------------------------------------------
# projection
# ------------------------------------------------------------------------------------------------
def quat_to_rotmat(quats: Tensor) -> Tensor:
    """_torch_impl.py:10-30. quats (w, x, y, z), normalised inside."""
    q = F.normalize(quats, p=2, dim=-1)
    w, x, y, z = torch.unbind(q, dim=-1)
    R = torch.stack(
        [
            1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
            2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
            2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y),
        ],
        dim=-1,
    )
    return R.reshape(quats.shape[:-1] + (3, 3))


def quat_scale_to_covar(quats: Tensor, scales: Tensor) -> Tensor:
    """_torch_impl.py:41-75: Sigma = (R S)(R S)^T, [N, 3, 3]."""
    M = quat_to_rotmat(quats) * scales[..., None, :]
    return M @ M.transpose(-1, -2)


def covar_from_triu(triu: Tensor) -> Tensor:
    xx, xy, xz, yy, yz, zz = triu.unbind(-1)
    return torch.stack([xx, xy, xz, xy, yy, yz, xz, yz, zz], dim=-1).reshape(triu.shape[:-1] + (3, 3))


def fully_fused_projection(
    means: Tensor,  # [N, 3]
    covars: Tensor,  # [N, 3, 3]
    viewmats: Tensor,  # [C, 4, 4]
    Ks: Tensor,  # [C, 3, 3]
    width: int,
    height: int,
    eps2d: float = 0.3,
    near_plane: float = 0.01,
    far_plane: float = 1e10,
    radius_clip: float = 0.0,
    calc_compensations: bool = False,
    camera_model: str = "pinhole",
):
    """_torch_impl.py:250-328 -> (radii int32 [C,N], means2d, depths, conics, compensations|None)."""
    C, N = viewmats.shape[0], means.shape[0]
    R, t = viewmats[:, :3, :3], viewmats[:, :3, 3]
    means_c = torch.einsum("cij,nj->cni", R, means) + t[:, None, :]
    covars_c = torch.einsum("cij,njk,clk->cnil", R, covars, R)
    fx, fy = Ks[:, 0, 0, None], Ks[:, 1, 1, None]
    cx, cy = Ks[:, 0, 2, None], Ks[:, 1, 2, None]
    tx, ty, tz = means_c.unbind(-1)
    O = torch.zeros((C, N), dtype=means.dtype)
    if camera_model == "pinhole":  # _persp_proj :78-122
        lim_xp = (width - cx) / fx + 0.3 * (0.5 * width / fx)
        lim_xn = cx / fx + 0.3 * (0.5 * width / fx)
        lim_yp = (height - cy) / fy + 0.3 * (0.5 * height / fy)
        lim_yn = cy / fy + 0.3 * (0.5 * height / fy)
        txc = tz * torch.minimum(torch.maximum(tx / tz, -lim_xn), lim_xp)
        tyc = tz * torch.minimum(torch.maximum(ty / tz, -lim_yn), lim_yp)
        J = torch.stack([fx / tz, O, -fx * txc / tz**2, O, fy / tz, -fy * tyc / tz**2], dim=-1).reshape(C, N, 2, 3)
        means2d = torch.stack([fx * tx / tz + cx, fy * ty / tz + cy], dim=-1)
    elif camera_model == "ortho":  # _ortho_proj :203-216
        J = torch.stack([fx + O, O, O, O, fy + O, O], dim=-1).reshape(C, N, 2, 3)
        means2d = torch.stack([fx * tx + cx, fy * ty + cy], dim=-1)
    elif camera_model == "fisheye":  # _fisheye_proj :123-182 (equidistant model, the reference's eps placement)
        eps = 0.0000001
        xy_len = (tx**2 + ty**2) ** 0.5 + eps
        theta = torch.atan2(xy_len, tz + eps)
        means2d = torch.stack([tx * fx * theta / xy_len + cx, ty * fy * theta / xy_len + cy], dim=-1)
        x2, y2, xy = tx * tx + eps, ty * ty, tx * ty
        x2y2 = x2 + y2
        x2y2z2_inv = 1.0 / (x2y2 + tz * tz)
        b_ = torch.atan2(xy_len, tz) / xy_len / x2y2
        a_ = tz * x2y2z2_inv / x2y2
        J = torch.stack([fx * (x2 * a_ + y2 * b_), fx * xy * (a_ - b_), -fx * tx * x2y2z2_inv,
                         fy * xy * (a_ - b_), fy * (y2 * a_ + x2 * b_), -fy * ty * x2y2z2_inv],
                        dim=-1).reshape(C, N, 2, 3)
    else:
        raise NotImplementedError(camera_model)
    cov2d = J @ covars_c @ J.transpose(-1, -2)
    det_orig = cov2d[..., 0, 0] * cov2d[..., 1, 1] - cov2d[..., 0, 1] * cov2d[..., 1, 0]
    cov2d = cov2d + torch.eye(2, dtype=means.dtype) * eps2d
    det = cov2d[..., 0, 0] * cov2d[..., 1, 1] - cov2d[..., 0, 1] * cov2d[..., 1, 0]
    det_c = det.clamp(min=1e-10)
    compensations = torch.sqrt(torch.clamp(det_orig / det_c, min=0.0)) if calc_compensations else None
    conics = torch.stack(
        [cov2d[..., 1, 1] / det_c, -(cov2d[..., 0, 1] + cov2d[..., 1, 0]) / 2.0 / det_c, cov2d[..., 0, 0] / det_c],
        dim=-1,
    )
    depths = tz
    b = (cov2d[..., 0, 0] + cov2d[..., 1, 1]) / 2
    v1 = b + torch.sqrt(torch.clamp(b * b - det_c, min=0.01))
    radius = torch.ceil(3.0 * torch.sqrt(v1)).detach()
    valid = (det_c > 0) & (depths > near_plane) & (depths < far_plane)
    valid = valid & (radius > radius_clip)  # fused_projection_fwd.py:215 (not in _torch_impl)
    inside = (
        (means2d[..., 0] + radius > 0) & (means2d[..., 0] - radius < width)
        & (means2d[..., 1] + radius > 0) & (means2d[..., 1] - radius < height)
    )
    radius = torch.where(valid & inside, radius, torch.zeros_like(radius))
    return radius.int(), means2d, depths, conics, compensations


# ------------------------------------------------------------------------------------------------
# spherical harmonics
# ------------------------------------------------------------------------------------------------
def sh_bases(basis_dim: int, dirs: Tensor) -> Tensor:
    """_torch_impl.py:620-701 (Sloan 2013), unit `dirs` [..., 3] -> [..., basis_dim]."""
    x, y, z = dirs.unbind(-1)
    out = [torch.full_like(x, 0.2820947917738781)]
    if basis_dim > 1:
        a = -0.48860251190292
        out += [a * y, -a * z, a * x]
    if basis_dim > 4:
        z2 = z * z
        fB = -1.092548430592079 * z
        fA = 0.5462742152960395
        c1, s1 = x * x - y * y, 2 * x * y
        out += [fA * s1, fB * y, 0.9461746957575601 * z2 - 0.3153915652525201, fB * x, fA * c1]
    if basis_dim > 9:
        fC = -2.285228997322329 * z2 + 0.4570457994644658
        fB = 1.445305721320277 * z
        fA = -0.5900435899266435
        c2, s2 = x * c1 - y * s1, x * s1 + y * c1
        out += [fA * s2, fB * s1, fC * y, z * (1.865881662950577 * z2 - 1.119528997770346), fC * x, fB * c1, fA * c2]
    if basis_dim > 16:
        fD = z * (-4.683325804901025 * z2 + 2.007139630671868)
        fC = 3.31161143515146 * z2 - 0.47308734787878
        fB = -1.770130769779931 * z
        fA = 0.6258357354491763
        c3, s3 = x * c2 - y * s2, x * s2 + y * c2
        y40 = 1.984313483298443 * z2 * (1.865881662950577 * z2 - 1.119528997770346) + -1.006230589874905 * (
            0.9461746957575601 * z2 - 0.3153915652525201
        )
        out += [fA * s3, fB * s2, fC * s1, fD * y, y40, fD * x, fC * c1, fB * c2, fA * c3]
    return torch.stack(out[:basis_dim], dim=-1)


def spherical_harmonics(degree: int, dirs: Tensor, coeffs: Tensor, masks: Optional[Tensor] = None) -> Tensor:
    """_torch_impl.py:704-714; masked-out entries are 0 (triton_impl/_wrapper.py:567-568)."""
    d = F.normalize(dirs, p=2, dim=-1)
    nb = (degree + 1) ** 2
    bases = sh_bases(nb, d)  # [..., nb]
    colors = (bases[..., None] * coeffs[..., :nb, :]).sum(dim=-2)
    if masks is not None:
        colors = colors * masks[..., None].to(colors.dtype)
    return colors


# ------------------------------------------------------------------------------------------------
# tile intersection
# ------------------------------------------------------------------------------------------------
def tile_n_bits(n_tiles: int, rule: str = "triton") -> int:
    """triton: (n_tiles-1).bit_length() (isect_tiles.py:105); cuda/torch: n_tiles.bit_length() (:367)."""
    return int(n_tiles).bit_length() if rule == "cuda" else max(int(n_tiles) - 1, 0).bit_length()


@torch.no_grad()
def isect_tiles(
    means2d: Tensor,  # [C, N, 2] or [nnz, 2]
    radii: Tensor,  # [C, N] or [nnz]
    depths: Tensor,  # [C, N] or [nnz]
    tile_size: int,
    tile_width: int,
    tile_height: int,
    sort: bool = True,
    camera_ids: Optional[Tensor] = None,  # packed only
    n_cameras: Optional[int] = None,
    bits_rule: str = "triton",
):
    """_torch_impl.py:331-400, loop replaced by numpy index arithmetic with the same emission order
    (row index ascending, then tile y, then tile x) and a STABLE sort (CUB radix sort is stable)."""
    packed = camera_ids is not None
    m = means2d.detach().reshape(-1, 2).numpy().astype(np.float32)
    r = radii.reshape(-1).numpy()
    d = depths.detach().reshape(-1).numpy().astype(np.float32)
    P = r.shape[0]
    ts = np.float32(tile_size)
    tm = m / ts
    tr = (r.astype(np.float32) / ts)[:, None]
    tmin = np.floor(tm - tr).astype(np.int32)
    tmax = np.ceil(tm + tr).astype(np.int32)
    lim = np.array([tile_width, tile_height], dtype=np.int32)
    tmin = np.clip(tmin, 0, lim)
    tmax = np.clip(tmax, 0, lim)
    ext = tmax - tmin
    cnt = (ext[:, 0] * ext[:, 1]) * (r > 0)
    tiles_per_gauss = torch.from_numpy(cnt.astype(np.int32)).reshape(radii.shape)
    n_isects = int(cnt.sum())
    rows = np.repeat(np.arange(P, dtype=np.int64), cnt)
    start = np.cumsum(cnt) - cnt
    k = np.arange(n_isects, dtype=np.int64) - np.repeat(start, cnt)
    w = np.repeat(ext[:, 0].astype(np.int64), cnt)
    tx = np.repeat(tmin[:, 0].astype(np.int64), cnt) + k % np.maximum(w, 1)
    ty = np.repeat(tmin[:, 1].astype(np.int64), cnt) + k // np.maximum(w, 1)
    tile = ty * tile_width + tx
    if packed:
        cam = camera_ids.numpy().astype(np.int64)[rows]
    else:
        N = radii.shape[1]
        cam = rows // N
    tb = tile_n_bits(tile_width * tile_height, bits_rule)
    depth_bits = d.view(np.int32).astype(np.int64)[rows]  # sign-extending cast, like struct.unpack("i")
    ids = (cam << (32 + tb)) | (tile << 32) | depth_bits
    flat = rows.astype(np.int32)
    if sort:
        order = np.argsort(ids, kind="stable")
        ids, flat = ids[order], flat[order]
    return tiles_per_gauss, torch.from_numpy(ids), torch.from_numpy(flat)


@torch.no_grad()
def isect_offset_encode(isect_ids: Tensor, C: int, tile_width: int, tile_height: int, bits_rule: str = "triton"):
    """_torch_impl.py:403-429: first sorted index of every (camera, tile); empty tiles inherit the next start."""
    tb = tile_n_bits(tile_width * tile_height, bits_rule)
    n_tiles = tile_width * tile_height
    hi = isect_ids.numpy() >> 32
    flat_tile = (hi >> tb) * n_tiles + (hi & ((1 << tb) - 1))
    counts = np.bincount(flat_tile, minlength=C * n_tiles).astype(np.int64)
    offsets = np.cumsum(counts) - counts
    return torch.from_numpy(offsets.astype(np.int32)).reshape(C, tile_height, tile_width)


# ------------------------------------------------------------------------------------------------
# compositor
# ------------------------------------------------------------------------------------------------
def rasterize_to_pixels(
    means2d: Tensor,  # [C, N, 2] or [nnz, 2]
    conics: Tensor,  # [C, N, 3] or [nnz, 3]
    colors: Tensor,  # [C, N, D] or [nnz, D]
    opacities: Tensor,  # [C, N] or [nnz]
    image_width: int,
    image_height: int,
    tile_size: int,
    isect_offsets: Tensor,  # [C, th, tw]
    flatten_ids: Tensor,  # [n_isects]
    backgrounds: Optional[Tensor] = None,  # [C, D]
    masks: Optional[Tensor] = None,  # [C, th, tw] True = render
    return_last_ids: bool = False,
):
    """Dense per-tile restatement of the compositor, differentiable through torch autograd.

    Per pixel (x+.5, y+.5), Gaussians of its tile in sorted order (RasterizeToPixels3DGSFwd.cu:140-169):
    sigma = .5(a dx^2 + c dy^2) + b dx dy; alpha = min(.999, o e^-sigma); skip if sigma < 0 or
    alpha < 1/255; stop *before* the Gaussian that would make T <= 1e-4; colour += alpha T col.
    """
    C, th, tw = isect_offsets.shape
    D = colors.shape[-1]
    n_isects = flatten_ids.shape[0]
    m_f, c_f, col_f, o_f = means2d.reshape(-1, 2), conics.reshape(-1, 3), colors.reshape(-1, D), opacities.reshape(-1)
    offs = torch.cat([isect_offsets.reshape(-1).long(), torch.tensor([n_isects])]).tolist()
    fid = flatten_ids.long()
    out_c = [[None] * tw for _ in range(C * th)]
    out_a = [[None] * tw for _ in range(C * th)]
    out_l = [[None] * tw for _ in range(C * th)]
    dt = means2d.dtype
    for cam in range(C):
        for ty in range(th):
            for tx in range(tw):
                t = (cam * th + ty) * tw + tx
                x0, y0 = tx * tile_size, ty * tile_size
                w, h = min(tile_size, image_width - x0), min(tile_size, image_height - y0)
                ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
                px = (xs.reshape(-1) + x0).to(dt) + 0.5
                py = (ys.reshape(-1) + y0).to(dt) + 0.5
                lo, hi = offs[t], offs[t + 1]
                render = masks is None or bool(masks[cam, ty, tx])
                if hi > lo and render:
                    g = fid[lo:hi]
                    dx = m_f[g, 0][None, :] - px[:, None]  # [npix, G]
                    dy = m_f[g, 1][None, :] - py[:, None]
                    cn = c_f[g]
                    sigma = 0.5 * (cn[:, 0] * dx * dx + cn[:, 2] * dy * dy) + cn[:, 1] * dx * dy
                    alpha = torch.clamp_max(o_f[g][None, :] * torch.exp(-sigma), ALPHA_MAX)
                    valid = ~(sigma < 0) & ~(alpha < ALPHA_MIN)
                    a_eff = torch.where(valid, alpha, torch.zeros_like(alpha))
                    one_m = 1.0 - a_eff
                    T_incl = torch.cumprod(one_m, dim=1)
                    T_before = torch.cat([torch.ones_like(T_incl[:, :1]), T_incl[:, :-1]], dim=1)
                    stop = (valid & (T_incl.detach() <= T_MIN)).to(torch.int32).cummax(dim=1).values.bool()
                    live = valid & ~stop
                    wgt = torch.where(live, a_eff * T_before, torch.zeros_like(a_eff))
                    pix = wgt @ col_f[g]  # [npix, D]
                    acc = wgt.sum(dim=1)
                    idx = torch.arange(lo, hi)[None, :].expand_as(live)
                    last = torch.where(live, idx, torch.zeros_like(idx)).max(dim=1).values
                else:
                    pix = torch.zeros((h * w, D), dtype=dt)
                    acc = torch.zeros((h * w,), dtype=dt)
                    last = torch.zeros((h * w,), dtype=torch.long)
                if backgrounds is not None:
                    pix = pix + (1.0 - acc)[:, None] * backgrounds[cam][None, :]
                out_c[cam * th + ty][tx] = pix.reshape(h, w, D)
                out_a[cam * th + ty][tx] = acc.reshape(h, w, 1)
                out_l[cam * th + ty][tx] = last.reshape(h, w)
    rows_c = [torch.cat(r, dim=1) for r in out_c]
    rows_a = [torch.cat(r, dim=1) for r in out_a]
    rows_l = [torch.cat(r, dim=1) for r in out_l]
    render_colors = torch.stack([torch.cat(rows_c[c * th:(c + 1) * th], dim=0) for c in range(C)])
    render_alphas = torch.stack([torch.cat(rows_a[c * th:(c + 1) * th], dim=0) for c in range(C)])
    if return_last_ids:
        last_ids = torch.stack([torch.cat(rows_l[c * th:(c + 1) * th], dim=0) for c in range(C)]).int()
        return render_colors, render_alphas, last_ids
    return render_colors, render_alphas


# ------------------------------------------------------------------------------------------------
# the whole path, as gsplat/rendering.py:44-598 orchestrates it (unpacked, single process)
# ------------------------------------------------------------------------------------------------
@torch.no_grad()
def rasterize_to_indices_in_range(range_start, range_end, transmittances, means2d, conics, opacities, image_width,
                          image_height, tile_size, isect_offsets, flatten_ids):
    """RasterizeToIndices3DGS.cu:14-170 on the CPU, one pixel at a time (grouped by camera, pixel)."""
    C, N, _ = means2d.shape
    th, tw = isect_offsets.shape[1:]
    n_isects = flatten_ids.numel()
    offs = torch.cat([isect_offsets.flatten().long(), torch.tensor([n_isects])]).numpy()
    fid = flatten_ids.long().numpy()
    m = means2d.reshape(-1, 2).numpy()
    cn = conics.reshape(-1, 3).numpy()
    op = opacities.reshape(-1).numpy()
    block = tile_size * tile_size
    gs, px, cam = [], [], []
    for c in range(C):
        for i in range(image_height):
            for j in range(image_width):
                t = (c * th + i // tile_size) * tw + j // tile_size
                lo, hi = int(offs[t]), int(offs[t + 1])
                lo2 = lo + block * range_start
                hi2 = min(hi, lo + block * range_end)
                if lo2 >= hi2:
                    continue
                g = fid[lo2:hi2]
                dx = m[g, 0] - np.float32(j + 0.5)
                dy = m[g, 1] - np.float32(i + 0.5)
                sigma = np.float32(0.5) * (cn[g, 0] * dx * dx + cn[g, 2] * dy * dy) + cn[g, 1] * dx * dy
                alpha = np.minimum(np.float32(0.999), op[g] * np.exp(-sigma))
                T = np.float32(transmittances[c, i, j].item())
                for k in range(g.shape[0]):
                    if sigma[k] < 0 or alpha[k] < np.float32(1.0 / 255.0):
                        continue
                    nT = T * (np.float32(1.0) - alpha[k])
                    if nT <= np.float32(1e-4):
                        break
                    gs.append(int(g[k]) % N)
                    px.append(i * image_width + j)
                    cam.append(c)
                    T = nT
    return (torch.tensor(gs, dtype=torch.long), torch.tensor(px, dtype=torch.long),
            torch.tensor(cam, dtype=torch.long))


def rasterization(
    means, quats, scales, opacities, colors, viewmats, Ks, width, height,
    near_plane=0.01, far_plane=1e10, radius_clip=0.0, eps2d=0.3, sh_degree=None, tile_size=16,
    backgrounds=None, render_mode="RGB", rasterize_mode="classic", camera_model="pinhole", covars=None,
    bits_rule="triton", compositor="torch", pixel_weight=None, timers=None,
):
    """CPU restatement of rendering.py:44-598 for packed=False, distributed=False.

    `compositor="c"` composites with oracle/gs_compositor.c (OpenMP; full-size scenes in seconds) instead of the
    torch-autograd restatement above; `meta` then also carries `margin` (per-pixel distance of the closest
    threshold decision, see gs_compositor.c) and `last_ids`, and `pixel_weight` (bool [C,H,W]) removes pixels
    from the backward pass. `timers` (dict) receives the forward wall time of the three stage groups."""
    import time

    t_start = time.perf_counter()
    C = viewmats.shape[0]
    if covars is None:
        covars = quat_scale_to_covar(quats, scales)
    radii, means2d, depths, conics, comp = fully_fused_projection(
        means, covars, viewmats, Ks, width, height, eps2d=eps2d, near_plane=near_plane, far_plane=far_plane,
        radius_clip=radius_clip, calc_compensations=(rasterize_mode == "antialiased"), camera_model=camera_model,
    )
    opac = opacities.repeat(C, 1)
    if comp is not None:
        opac = opac * comp
    if sh_degree is None:
        cols = colors.expand(C, -1, -1) if colors.dim() == 2 else colors
    else:
        c2w = torch.inverse(viewmats)
        dirs = means[None, :, :] - c2w[:, None, :3, 3]
        shs = colors.expand(C, -1, -1, -1) if colors.dim() == 3 else colors
        cols = spherical_harmonics(sh_degree, dirs, shs, masks=radii > 0)
        cols = torch.clamp_min(cols + 0.5, 0.0)
    if render_mode in ("RGB+D", "RGB+ED"):
        cols = torch.cat((cols, depths[..., None]), dim=-1)
        if backgrounds is not None:
            backgrounds = torch.cat([backgrounds, torch.zeros(C, 1)], dim=-1)
    elif render_mode in ("D", "ED"):
        cols = depths[..., None]
        if backgrounds is not None:
            backgrounds = torch.zeros(C, 1)
    tw, th = math.ceil(width / float(tile_size)), math.ceil(height / float(tile_size))
    t_proj = time.perf_counter()
    tiles_per_gauss, isect_ids, flatten_ids = isect_tiles(means2d, radii, depths, tile_size, tw, th, bits_rule=bits_rule)
    offsets = isect_offset_encode(isect_ids, C, tw, th, bits_rule=bits_rule)
    t_isect = time.perf_counter()
    extra = {}
    if compositor == "c":
        import c_oracle

        render_colors, render_alphas, last_ids, margin = c_oracle.rasterize_to_pixels(
            means2d, conics, cols, opac, width, height, tile_size, offsets, flatten_ids, backgrounds=backgrounds,
            return_last_ids=True, return_margin=True, pixel_weight=pixel_weight)
        extra = {"margin": margin, "last_ids": last_ids}
    else:
        assert pixel_weight is None
        render_colors, render_alphas = rasterize_to_pixels(
            means2d, conics, cols, opac, width, height, tile_size, offsets, flatten_ids, backgrounds=backgrounds
        )
    if timers is not None:
        timers.update(projection_colors_fwd_s=t_proj - t_start, isect_sort_offsets_s=t_isect - t_proj,
                      compositor_fwd_s=time.perf_counter() - t_isect)
    if render_mode in ("ED", "RGB+ED"):
        render_colors = torch.cat(
            [render_colors[..., :-1], render_colors[..., -1:] / render_alphas.clamp(min=1e-10)], dim=-1
        )
    meta = {"radii": radii, "means2d": means2d, "depths": depths, "conics": conics, "opacities": opac,
            "tiles_per_gauss": tiles_per_gauss, "isect_ids": isect_ids, "flatten_ids": flatten_ids,
            "isect_offsets": offsets, "tile_width": tw, "tile_height": th, "colors": cols}
    meta.update(extra)
    return render_colors, render_alphas, meta
