"""Python side of the `b200` backend: the five functions `gsplat/rendering.py` imports.

Mirrors (same names, arguments, defaults, return values, error behaviour)
  reference gsplat/triton_impl/_wrapper.py: fully_fused_projection :429, isect_tiles :544,
  isect_offset_encode :548, spherical_harmonics :596, rasterize_to_pixels :185
and, for the arguments Triton does not implement (packed, sparse_grad, covars, ortho),
  reference gsplat/cuda/_wrapper.py :209, :349, :422, :441, :82.

Every function allocates its outputs with torch (so the caching allocator and the current
stream stay PyTorch's) and passes raw device pointers to libgsplat_b200.so through ctypes.
No Triton, no torch fallback: a missing library or a CPU tensor raises.
"""
import os
from typing import Optional, Tuple

import torch
from torch import Tensor

from ._lib import lib, ptr, require_cuda, stream_of

_CAMERA_MODELS = {"pinhole": 0, "ortho": 1, "fisheye": 2}
_RASTER_CHANNELS = (1, 2, 3, 4, 8, 16, 32)


def _camera_model_id(camera_model: str) -> int:
    if camera_model not in _CAMERA_MODELS:
        # (the Triton reference only has pinhole, triton_impl/_wrapper.py:322-323; the CUDA one has these three)
        raise NotImplementedError(f"camera_model={camera_model!r} is not implemented by the b200 backend")
    return _CAMERA_MODELS[camera_model]


def tile_n_bits(n_tiles: int) -> int:
    """Bits reserved for the tile id inside an intersection key.

    Triton rule `(n_tiles-1).bit_length()` (triton_impl/isect_tiles.py:105, isect_offset.py:23) by
    default; GSPLAT_B200_TILE_BITS=cuda selects `n_tiles.bit_length()` (IntersectTile.cu:145,
    _torch_impl.py:367). They differ only when n_tiles is a power of two.
    """
    if os.environ.get("GSPLAT_B200_TILE_BITS", "triton") == "cuda":
        return int(n_tiles).bit_length()
    return max(int(n_tiles) - 1, 0).bit_length()


# --------------------------------------------------------------------------------------------
# fully_fused_projection
# --------------------------------------------------------------------------------------------
def fully_fused_projection(
    means: Tensor,  # [N, 3]
    covars: Optional[Tensor],  # [N, 6] or None
    quats: Optional[Tensor],  # [N, 4] or None
    scales: Optional[Tensor],  # [N, 3] or None
    viewmats: Tensor,  # [C, 4, 4]
    Ks: Tensor,  # [C, 3, 3]
    width: int,
    height: int,
    eps2d: float = 0.3,
    near_plane: float = 0.01,
    far_plane: float = 1e10,
    radius_clip: float = 0.0,
    packed: bool = False,
    sparse_grad: bool = False,
    calc_compensations: bool = False,
    camera_model: str = "pinhole",
):
    """Projects Gaussians to 2D (quat/scale -> covariance, EWA projection, blur, radius, culling).

    Returns `(radii[C,N] int32, means2d[C,N,2], depths[C,N], conics[C,N,3], compensations|None)`
    or, with `packed=True`, `(camera_ids, gaussian_ids, radii, means2d, depths, conics,
    compensations)` with [nnz] rows in (camera, Gaussian) row-major order, ids int64.
    """
    C = viewmats.size(0)
    N = means.size(0)
    assert means.size() == (N, 3), means.size()
    assert viewmats.size() == (C, 4, 4), viewmats.size()
    assert Ks.size() == (C, 3, 3), Ks.size()
    means = means.contiguous()
    if covars is not None:
        assert covars.size() == (N, 6), covars.size()
        covars = covars.contiguous()
    else:
        assert quats is not None, "covars or quats is required"
        assert scales is not None, "covars or scales is required"
        assert quats.size() == (N, 4), quats.size()
        assert scales.size() == (N, 3), scales.size()
        quats = quats.contiguous()
        scales = scales.contiguous()
    if sparse_grad:
        assert packed, "sparse_grad is only supported when packed is True"
    _camera_model_id(camera_model)
    require_cuda(means, covars, quats, scales, viewmats, Ks)
    viewmats = viewmats.contiguous()
    Ks = Ks.contiguous()
    if packed:
        return _FullyFusedProjectionPacked.apply(
            means, covars, quats, scales, viewmats, Ks, width, height, eps2d, near_plane, far_plane,
            radius_clip, sparse_grad, calc_compensations, camera_model,
        )
    return _FullyFusedProjection.apply(
        means, covars, quats, scales, viewmats, Ks, width, height, eps2d, near_plane, far_plane,
        radius_clip, calc_compensations, camera_model,
    )


# Second, high-priority stream of the forward pass: rasterization() runs the intersection pipeline on it, beside the
# colour kernels of the caller's stream (the block scheduler hands free SM slots to the higher priority first, so the
# dozen small latency-bound kernels are not queued behind the colour kernel's thousands of blocks)
_SIDE_STREAMS = {}


def side_stream(device) -> "torch.cuda.Stream":
    idx = torch.device(device).index
    if idx is None:
        idx = torch.cuda.current_device()
    st = _SIDE_STREAMS.get(idx)
    if st is None:
        st = _SIDE_STREAMS[idx] = torch.cuda.Stream(device=idx, priority=-1)
    return st


def overlap_enabled() -> bool:
    return os.environ.get("GSPLAT_B200_OVERLAP", "1") == "1"


# distributed "push" exchange (gsplat.distributed.PushExchange): while set, the unpacked projection and the fused SH
# colour kernel also store their rows into the camera owners' mailboxes
_PUSH_CONTEXT = [None]


def set_push_context(ex) -> None:
    _PUSH_CONTEXT[0] = ex


def _f32(t: Optional[Tensor]) -> Optional[Tensor]:
    if t is None:
        return None
    return t.contiguous() if t.dtype == torch.float32 else t.float().contiguous()


class _FullyFusedProjection(torch.autograd.Function):
    """Unpacked projection. cf. triton_impl/_wrapper.py:300-427."""

    @staticmethod
    def forward(ctx, means, covars, quats, scales, viewmats, Ks, width, height, eps2d, near_plane,
                far_plane, radius_clip, calc_compensations, camera_model):
        C, N = viewmats.size(0), means.size(0)
        dev = means.device
        model = _camera_model_id(camera_model)
        radii = torch.empty((C, N), dtype=torch.int32, device=dev)
        means2d = torch.empty((C, N, 2), dtype=torch.float32, device=dev)
        depths = torch.empty((C, N), dtype=torch.float32, device=dev)
        conics = torch.empty((C, N, 3), dtype=torch.float32, device=dev)
        compensations = torch.empty((C, N), dtype=torch.float32, device=dev) if calc_compensations else None
        ex = _PUSH_CONTEXT[0]
        if ex is not None:
            world, C_local, N_total, n_base, peers, region = ex.push_args()
            assert C == world * C_local and N == ex.N_local, (C, N, world, C_local, ex.N_local)
            lib.gsb_projection_fwd_push(
                C, N, ptr(means), ptr(covars), ptr(quats), ptr(scales), ptr(viewmats), ptr(Ks), width, height,
                eps2d, near_plane, far_plane, radius_clip, model, ptr(radii), ptr(means2d), ptr(depths),
                ptr(conics), ptr(compensations), ptr(ex.opacities), world, C_local, N_total, n_base, ex.D, peers,
                region, stream_of(means),
            )
            ex.pushed.add("geometry")
        else:
            lib.gsb_projection_fwd(
                C, N, ptr(means), ptr(covars), ptr(quats), ptr(scales), ptr(viewmats), ptr(Ks), width, height,
                eps2d, near_plane, far_plane, radius_clip, model, ptr(radii), ptr(means2d), ptr(depths),
                ptr(conics), ptr(compensations), stream_of(means),
            )
        ctx.save_for_backward(means, covars, quats, scales, viewmats, Ks, radii, conics, compensations)
        ctx.width, ctx.height, ctx.eps2d, ctx.model = width, height, eps2d, model
        ctx.mark_non_differentiable(radii)
        # outputs nobody differentiates (radii always, depths in the RGB modes) arrive as None in backward instead of
        # zero tensors autograd would fill on the critical path (the kernel takes NULL for v_depths / v_compensations)
        ctx.set_materialize_grads(False)
        return radii, means2d, depths, conics, compensations

    @staticmethod
    def backward(ctx, v_radii, v_means2d, v_depths, v_conics, v_compensations):
        means, covars, quats, scales, viewmats, Ks, radii, conics, compensations = ctx.saved_tensors
        C, N = viewmats.size(0), means.size(0)
        need = ctx.needs_input_grad
        v_means = torch.empty_like(means) if need[0] else None
        v_covars = torch.empty_like(covars) if (covars is not None and need[1]) else None
        v_quats = torch.empty_like(quats) if (quats is not None and need[2]) else None
        v_scales = torch.empty_like(scales) if (scales is not None and need[3]) else None
        v_viewmats = torch.empty_like(viewmats) if need[4] else None
        if compensations is None:
            v_compensations = None
        if v_means2d is None:
            v_means2d = torch.zeros((C, N, 2), dtype=torch.float32, device=means.device)
        if v_conics is None:
            v_conics = torch.zeros((C, N, 3), dtype=torch.float32, device=means.device)
        # converted cotangents stay bound until after the launch (a temporary freed before the launch could be
        # handed to the next conversion by the caching allocator and overwritten on the same stream)
        v_means2d, v_depths, v_conics, v_compensations = (
            _f32(t) for t in (v_means2d, v_depths, v_conics, v_compensations))
        lib.gsb_projection_bwd(
            C, N, ptr(means), ptr(covars), ptr(quats), ptr(scales), ptr(viewmats), ptr(Ks), ctx.width,
            ctx.height, ctx.eps2d, ctx.model, ptr(radii), ptr(conics), ptr(compensations),
            ptr(v_means2d), ptr(v_depths), ptr(v_conics), ptr(v_compensations),
            ptr(v_means), ptr(v_covars), ptr(v_quats), ptr(v_scales), ptr(v_viewmats), stream_of(means),
        )
        return (v_means, v_covars, v_quats, v_scales, v_viewmats) + (None,) * 9


_PACKED_COUNTS = [None]  # (camera_ids.data_ptr(), nnz, rows per camera) of the last packed projection


def packed_camera_counts(camera_ids: Tensor, n_cameras: int):
    """Rows per camera of a packed projection, as a host list without a device read-back when `camera_ids` is the
    tensor the last `fully_fused_projection(packed=True)` returned (its counts came back with nnz)."""
    last = _PACKED_COUNTS[0]
    if last is not None and last[0] == camera_ids.data_ptr() and last[1] == camera_ids.numel() \
            and len(last[2]) == n_cameras:
        return list(last[2])
    return torch.bincount(camera_ids, minlength=n_cameras).tolist()


class _FullyFusedProjectionPacked(torch.autograd.Function):
    """Packed (COO) projection. cf. cuda/_wrapper.py:998-1190."""

    @staticmethod
    def forward(ctx, means, covars, quats, scales, viewmats, Ks, width, height, eps2d, near_plane,
                far_plane, radius_clip, sparse_grad, calc_compensations, camera_model):
        C, N = viewmats.size(0), means.size(0)
        dev = means.device
        model = _camera_model_id(camera_model)
        st = stream_of(means)
        nblk = (N + 255) // 256
        block_counts = torch.empty((C * nblk,), dtype=torch.int32, device=dev)
        block_offsets = torch.empty((C * nblk + 1,), dtype=torch.int64, device=dev)
        args = (C, N, ptr(means), ptr(covars), ptr(quats), ptr(scales), ptr(viewmats), ptr(Ks), width, height,
                eps2d, near_plane, far_plane, radius_clip, model)
        lib.gsb_projection_packed_count(*args, ptr(block_counts), st)
        lib.gsb_scan_i32_to_i64(C * nblk, ptr(block_counts), ptr(block_offsets),
                                block_offsets.data_ptr() + 8 * C * nblk, st)
        # size read-back, as Projection.cpp:332-333; the same copy brings the per-camera row counts (the scan is
        # camera-major), which the distributed packed exchange needs on the host (rendering.py:416-418 of the reference
        # does a bincount + one .item() per rank for them)
        bounds = block_offsets[::nblk].tolist() if nblk > 0 else [0] * (C + 1)
        nnz = int(bounds[-1])
        camera_ids = torch.empty((nnz,), dtype=torch.int64, device=dev)
        gaussian_ids = torch.empty((nnz,), dtype=torch.int64, device=dev)
        radii = torch.empty((nnz,), dtype=torch.int32, device=dev)
        means2d = torch.empty((nnz, 2), dtype=torch.float32, device=dev)
        depths = torch.empty((nnz,), dtype=torch.float32, device=dev)
        conics = torch.empty((nnz, 3), dtype=torch.float32, device=dev)
        compensations = torch.empty((nnz,), dtype=torch.float32, device=dev) if calc_compensations else None
        if nnz > 0:
            lib.gsb_projection_packed_write(
                *args, ptr(block_offsets), ptr(camera_ids), ptr(gaussian_ids), ptr(radii), ptr(means2d),
                ptr(depths), ptr(conics), ptr(compensations), st,
            )
        ctx.save_for_backward(camera_ids, gaussian_ids, means, covars, quats, scales, viewmats, Ks, conics,
                              compensations)
        ctx.width, ctx.height, ctx.eps2d, ctx.model, ctx.sparse_grad = width, height, eps2d, model, sparse_grad
        ctx.mark_non_differentiable(camera_ids, gaussian_ids, radii)
        ctx.set_materialize_grads(False)  # see _FullyFusedProjection
        _PACKED_COUNTS[0] = (camera_ids.data_ptr(), nnz, [int(b - a) for a, b in zip(bounds[:-1], bounds[1:])])
        return camera_ids, gaussian_ids, radii, means2d, depths, conics, compensations

    @staticmethod
    def backward(ctx, v_camera_ids, v_gaussian_ids, v_radii, v_means2d, v_depths, v_conics, v_compensations):
        (camera_ids, gaussian_ids, means, covars, quats, scales, viewmats, Ks, conics,
         compensations) = ctx.saved_tensors
        C, N, nnz = viewmats.size(0), means.size(0), camera_ids.numel()
        need = ctx.needs_input_grad
        sparse = ctx.sparse_grad
        dev = means.device
        rows = nnz if sparse else N

        def out(t, width, needed):
            if t is None or not needed:
                return None
            return torch.empty((rows, width), dtype=torch.float32, device=dev)

        v_means = out(means, 3, need[0])
        v_covars = out(covars, 6, need[1])
        v_quats = out(quats, 4, need[2])
        v_scales = out(scales, 3, need[3])
        v_viewmats = torch.empty_like(viewmats) if need[4] else None
        acc = None if sparse else torch.empty((N, 9), dtype=torch.float32, device=dev)
        if compensations is None:
            v_compensations = None
        if v_means2d is None:
            v_means2d = torch.zeros((nnz, 2), dtype=torch.float32, device=dev)
        if v_conics is None:
            v_conics = torch.zeros((nnz, 3), dtype=torch.float32, device=dev)
        v_means2d, v_depths, v_conics, v_compensations = (
            _f32(t) for t in (v_means2d, v_depths, v_conics, v_compensations))  # bound past the launch
        lib.gsb_projection_packed_bwd(
            C, N, nnz, ptr(means), ptr(covars), ptr(quats), ptr(scales), ptr(viewmats), ptr(Ks), ctx.width,
            ctx.height, ctx.eps2d, ctx.model, ptr(camera_ids), ptr(gaussian_ids), ptr(conics),
            ptr(compensations), ptr(v_means2d), ptr(v_depths), ptr(v_conics),
            ptr(v_compensations), int(sparse), ptr(acc), ptr(v_means), ptr(v_covars), ptr(v_quats),
            ptr(v_scales), ptr(v_viewmats), stream_of(means),
        )
        if sparse:
            def coo(values, like):
                if values is None:
                    return None
                return torch.sparse_coo_tensor(indices=gaussian_ids[None], values=values, size=like.size(),
                                               is_coalesced=(C == 1))

            v_means, v_covars = coo(v_means, means), coo(v_covars, covars) if covars is not None else None
            v_quats = coo(v_quats, quats) if quats is not None else None
            v_scales = coo(v_scales, scales) if scales is not None else None
        return (v_means, v_covars, v_quats, v_scales, v_viewmats) + (None,) * 10


# --------------------------------------------------------------------------------------------
# spherical_harmonics
# --------------------------------------------------------------------------------------------
def spherical_harmonics(
    degrees_to_use: int,
    dirs: Tensor,  # [..., 3]
    coeffs: Tensor,  # [..., K, 3]
    masks: Optional[Tensor] = None,  # [...]
) -> Tensor:
    """Colours from SH coefficients (degree <= 4); `dirs` need not be normalised.

    Same contract as triton_impl/_wrapper.py:596-620 / cuda/_wrapper.py:82-118. Additionally,
    `coeffs` may omit the leading camera dim (`dirs[C,N,3]`, `coeffs[N,K,3]`): the kernel then
    broadcasts inside the thread and `[C,N,K,3]` is never materialised. A stride-0
    `coeffs.expand(C, ...)` as built by the reference's rendering.py:402 is detected and
    handled the same way.
    """
    assert (degrees_to_use + 1) ** 2 <= coeffs.shape[-2], coeffs.shape
    assert dirs.shape[-1] == 3, dirs.shape
    assert coeffs.shape[-1] == 3, coeffs.shape
    if coeffs.dim() == dirs.dim() + 1:
        assert dirs.shape[:-1] == coeffs.shape[:-2], (dirs.shape, coeffs.shape)
        # undo a stride-0 broadcast over the leading dim (camera axis)
        if dirs.dim() == 3 and coeffs.stride(0) == 0 and coeffs.shape[0] > 1:
            coeffs = coeffs[0]
    else:
        assert dirs.dim() == 3 and coeffs.dim() == 3 and dirs.shape[1] == coeffs.shape[0], (
            dirs.shape, coeffs.shape)
    if masks is not None:
        assert masks.shape == dirs.shape[:-1], masks.shape
        masks = masks.contiguous()
    require_cuda(dirs, coeffs, masks)
    return _SphericalHarmonics.apply(degrees_to_use, dirs.contiguous(), coeffs.contiguous(), masks)


class _SphericalHarmonics(torch.autograd.Function):
    """cf. triton_impl/_wrapper.py:552-594."""

    @staticmethod
    def _geometry(dirs, coeffs):
        K = coeffs.shape[-2]
        if coeffs.dim() == dirs.dim() + 1:
            return 1, dirs.numel() // 3, K
        return dirs.shape[0], dirs.shape[1], K

    @staticmethod
    def forward(ctx, sh_degree, dirs, coeffs, masks):
        n_batch, n_rows, K = _SphericalHarmonics._geometry(dirs, coeffs)
        colors = torch.empty_like(dirs)
        m8 = masks.to(torch.uint8) if masks is not None else None
        lib.gsb_sh_fwd(sh_degree, n_rows, n_batch, K, ptr(dirs), ptr(coeffs), ptr(m8), ptr(colors),
                       stream_of(dirs))
        ctx.save_for_backward(dirs, coeffs, m8)
        ctx.sh_degree = sh_degree
        return colors

    @staticmethod
    def backward(ctx, v_colors):
        dirs, coeffs, m8 = ctx.saved_tensors
        n_batch, n_rows, K = _SphericalHarmonics._geometry(dirs, coeffs)
        v_coeffs = torch.empty_like(coeffs)
        v_dirs = torch.empty_like(dirs) if ctx.needs_input_grad[1] else None
        if n_rows == 0 or n_batch == 0:
            v_coeffs.zero_()
        v_colors = _f32(v_colors)
        lib.gsb_sh_bwd(ctx.sh_degree, n_rows, n_batch, K, ptr(dirs), ptr(coeffs), ptr(m8),
                       ptr(v_colors), ptr(v_coeffs), ptr(v_dirs), stream_of(dirs))
        if not ctx.needs_input_grad[2]:
            v_coeffs = None
        return None, v_dirs, v_coeffs, None


def view_dependent_colors(sh_degree: int, means: Tensor, viewmats: Tensor, coeffs: Tensor, radii: Tensor) -> Tensor:
    """`clamp_min(spherical_harmonics(deg, means - campos, coeffs, masks=radii > 0) + 0.5, 0)` in one kernel.

    The fused form of the reference's rendering.py:384-408 for unpacked rows and per-Gaussian coefficients
    (`means [N,3]`, `viewmats [C,4,4]` without gradient, `coeffs [N,K,3]`, `radii [C,N]`) -> `[C,N,3]`.
    b200-only helper used by this package's `rasterization()`; the five reference-facing functions are unchanged.
    """
    C, N = radii.shape
    assert means.shape == (N, 3) and viewmats.shape == (C, 4, 4), (means.shape, viewmats.shape)
    assert coeffs.dim() == 3 and coeffs.shape[0] == N and coeffs.shape[2] == 3, coeffs.shape
    assert (sh_degree + 1) ** 2 <= coeffs.shape[1], coeffs.shape
    assert not viewmats.requires_grad, "use spherical_harmonics() when camera poses need gradients"
    require_cuda(means, viewmats, coeffs, radii)
    return _ViewColors.apply(sh_degree, means.contiguous(), viewmats.contiguous(), coeffs.contiguous(),
                             radii.contiguous())


# The staged colour kernels (csrc/sh.cu) run persistent CTAs. Alone on the GPU they take every SM slot that fits;
# beside the intersection pipeline / the projection backward on rasterization()'s second stream they must not: a
# persistent CTA never gives its slot back, so the other stream's (critical-path) kernels would queue behind the whole
# colour kernel. Measured at cfg2 (ms per step): forward kernel on 1 CTA per SM + backward on 3: 1.639; both on all
# four: 1.676 (no better than the plain kernels, 1.675); forward plain + backward on 4 / 3 / 2 / 1: 1.650 / 1.644 /
# 1.644 / 1.684. rasterization() announces the two-stream case here; GSPLAT_B200_SH_FWD_CTAS / GSPLAT_B200_SH_CTAS pin
# the caps for experiments. The forward cap is for ONE camera, where the colour kernel has 300 us of slack; with more
# cameras it is real work (cfg3, 8 cameras: 1.6 ms on 1 CTA per SM, 15.39 ms per step against 14.78 uncapped).
_COLOURS_BESIDE = [False]
_GSB_OPT_SH_CTAS, _GSB_OPT_SH_FWD_CTAS = 5, 7


def colour_kernels_share_the_gpu(flag: bool) -> None:
    _COLOURS_BESIDE[0] = bool(flag)


def _cap_colour_ctas(option: int, env: str, ctas_per_sm: int) -> None:
    if os.environ.get(env) is None:
        lib.gsb_set_option(option, ctas_per_sm)


class _ViewColors(torch.autograd.Function):
    @staticmethod
    def forward(ctx, sh_degree, means, viewmats, coeffs, radii):
        C, N = radii.shape
        K = coeffs.shape[1]
        st = stream_of(means)
        campos = torch.empty((C, 3), dtype=torch.float32, device=means.device)
        lib.gsb_camera_centers(C, ptr(viewmats), ptr(campos), st)
        colors = torch.empty((C, N, 3), dtype=torch.float32, device=means.device)
        ex = _PUSH_CONTEXT[0]
        if ex is not None:
            world, C_local, N_total, n_base, peers, region = ex.push_args()
            assert C == world * C_local and N == ex.N_local and ex.D == 3
            lib.gsb_sh_colors_fwd_push(sh_degree, N, C, K, ptr(means), ptr(campos), ptr(coeffs), ptr(radii),
                                       ptr(colors), world, C_local, N_total, n_base, peers, region, st)
            ex.pushed.add("colors")
        else:
            _cap_colour_ctas(_GSB_OPT_SH_FWD_CTAS, "GSPLAT_B200_SH_FWD_CTAS", 1 if (_COLOURS_BESIDE[0] and C == 1) else 0)
            lib.gsb_sh_colors_fwd(sh_degree, N, C, K, ptr(means), ptr(campos), ptr(coeffs), ptr(radii), ptr(colors), st)
        ctx.save_for_backward(means, campos, coeffs, radii)
        ctx.sh_degree = sh_degree
        ctx.beside = _COLOURS_BESIDE[0]
        return colors

    @staticmethod
    def backward(ctx, v_colors):
        means, campos, coeffs, radii = ctx.saved_tensors
        C, N = radii.shape
        v_coeffs = torch.empty_like(coeffs)
        v_means = torch.empty_like(means) if ctx.needs_input_grad[1] else None
        v_colors = _f32(v_colors)
        _cap_colour_ctas(_GSB_OPT_SH_CTAS, "GSPLAT_B200_SH_CTAS", 3 if ctx.beside else 0)
        lib.gsb_sh_colors_bwd(ctx.sh_degree, N, C, coeffs.shape[1], ptr(means), ptr(campos), ptr(coeffs), ptr(radii),
                              ptr(v_colors), ptr(v_coeffs), ptr(v_means), stream_of(means))
        return None, v_means, None, (v_coeffs if ctx.needs_input_grad[3] else None), None


class _GatherRows(torch.autograd.Function):
    """`table[ids]` whose backward is one `index_add_` (atomics) instead of torch's sort-based
    `index_put_(accumulate=True)`: the packed path gathers per-Gaussian rows for up to C*N projected rows."""

    @staticmethod
    def forward(ctx, table, ids):
        ctx.save_for_backward(ids)
        ctx.rows = table.shape[0]
        return table.index_select(0, ids)

    @staticmethod
    def backward(ctx, grad):
        (ids,) = ctx.saved_tensors
        out = torch.zeros((ctx.rows,) + tuple(grad.shape[1:]), dtype=grad.dtype, device=grad.device)
        out.index_add_(0, ids, grad)
        return out, None


def gather_rows(table: Tensor, ids: Tensor) -> Tensor:
    """Differentiable `table[ids]` for a 1-D int64 `ids` with many duplicates (packed rows -> their Gaussian)."""
    return _GatherRows.apply(table, ids)


def view_dependent_colors_packed(sh_degree: int, means: Tensor, viewmats: Tensor, coeffs: Tensor,
                                 camera_ids: Tensor, gaussian_ids: Tensor) -> Tensor:
    """Packed form of `view_dependent_colors`: colours `[nnz, 3]` of the projected rows (camera_ids, gaussian_ids).

    The reference gathers `coeffs[gaussian_ids]` into `[nnz, K, 3]` (rendering.py:390-396: 3.3 GB at 3 M Gaussians
    x 8 cameras, and a scatter-add of the same size backward). Here the unpacked kernel runs over the dense
    `[C, N]` grid with the visibility mask scattered from the packed ids (thread per Gaussian, cameras looped inside,
    coefficients read once, no atomics backward) and only the 3-float colours are gathered / scattered."""
    C, N = viewmats.shape[0], means.shape[0]
    assert coeffs.dim() == 3 and coeffs.shape[0] == N and coeffs.shape[2] == 3, coeffs.shape
    assert (sh_degree + 1) ** 2 <= coeffs.shape[1], coeffs.shape
    assert not viewmats.requires_grad, "use spherical_harmonics() when camera poses need gradients"
    require_cuda(means, viewmats, coeffs, camera_ids, gaussian_ids)
    return _ViewColorsPacked.apply(sh_degree, means.contiguous(), viewmats.contiguous(), coeffs.contiguous(),
                                   camera_ids, gaussian_ids)


class _ViewColorsPacked(torch.autograd.Function):
    @staticmethod
    def forward(ctx, sh_degree, means, viewmats, coeffs, camera_ids, gaussian_ids):
        C, N, K = viewmats.shape[0], means.shape[0], coeffs.shape[1]
        dev, st = means.device, stream_of(means)
        rows = camera_ids * N + gaussian_ids  # [nnz] flat (camera, Gaussian) index, unique
        visible = torch.zeros((C * N,), dtype=torch.int32, device=dev)
        visible.index_fill_(0, rows, 1)
        campos = torch.empty((C, 3), dtype=torch.float32, device=dev)
        lib.gsb_camera_centers(C, ptr(viewmats), ptr(campos), st)
        dense = torch.empty((C * N, 3), dtype=torch.float32, device=dev)
        _cap_colour_ctas(_GSB_OPT_SH_FWD_CTAS, "GSPLAT_B200_SH_FWD_CTAS", 1 if (_COLOURS_BESIDE[0] and C == 1) else 0)
        lib.gsb_sh_colors_fwd(sh_degree, N, C, K, ptr(means), ptr(campos), ptr(coeffs), ptr(visible), ptr(dense), st)
        ctx.save_for_backward(means, campos, coeffs, visible, rows)
        ctx.sh_degree = sh_degree
        ctx.beside = _COLOURS_BESIDE[0]
        return dense.index_select(0, rows)

    @staticmethod
    def backward(ctx, v_colors):
        means, campos, coeffs, visible, rows = ctx.saved_tensors
        C, N = campos.shape[0], means.shape[0]
        v_dense = torch.zeros((C * N, 3), dtype=torch.float32, device=means.device)
        v_dense.index_copy_(0, rows, _f32(v_colors))
        v_coeffs = torch.empty_like(coeffs)
        v_means = torch.empty_like(means) if ctx.needs_input_grad[1] else None
        _cap_colour_ctas(_GSB_OPT_SH_CTAS, "GSPLAT_B200_SH_CTAS", 3 if ctx.beside else 0)
        lib.gsb_sh_colors_bwd(ctx.sh_degree, N, C, coeffs.shape[1], ptr(means), ptr(campos), ptr(coeffs), ptr(visible),
                              ptr(v_dense), ptr(v_coeffs), ptr(v_means), stream_of(means))
        return None, v_means, None, (v_coeffs if ctx.needs_input_grad[3] else None), None, None


# --------------------------------------------------------------------------------------------
# isect_tiles / isect_offset_encode
# --------------------------------------------------------------------------------------------
@torch.no_grad()
def isect_tiles(
    means2d: Tensor,  # [C, N, 2] or [nnz, 2]
    radii: Tensor,  # [C, N] or [nnz]
    depths: Tensor,  # [C, N] or [nnz]
    tile_size: int,
    tile_width: int,
    tile_height: int,
    sort: bool = True,
    packed: bool = False,
    n_cameras: Optional[int] = None,
    camera_ids: Optional[Tensor] = None,
    gaussian_ids: Optional[Tensor] = None,
    _speculate: bool = False,
) -> Tuple[Tensor, Tensor, Tensor]:
    """Maps projected Gaussians to the tiles they overlap.

    Returns `(tiles_per_gauss int32 [C,N]|[nnz], isect_ids int64 [n_isects],
    flatten_ids int32 [n_isects])`, sorted by key when `sort`. Contract:
    triton_impl/isect_tiles.py:14-133, cuda/_wrapper.py:349-419.
    """
    if packed:
        nnz = means2d.size(0)
        assert means2d.shape == (nnz, 2), means2d.size()
        assert radii.shape == (nnz,), radii.size()
        assert depths.shape == (nnz,), depths.size()
        assert camera_ids is not None, "camera_ids is required if packed is True"
        assert gaussian_ids is not None, "gaussian_ids is required if packed is True"
        assert n_cameras is not None, "n_cameras is required if packed is True"
        camera_ids = camera_ids.contiguous()
        if camera_ids.dtype != torch.int64:
            camera_ids = camera_ids.long()
        C, N, P = n_cameras, 0, nnz
    else:
        C, N, _ = means2d.shape
        assert means2d.shape == (C, N, 2), means2d.size()
        assert radii.shape == (C, N), radii.size()
        assert depths.shape == (C, N), depths.size()
        P = C * N
    require_cuda(means2d, radii, depths, camera_ids)
    means2d = _f32(means2d)
    depths = _f32(depths)
    radii = radii.contiguous()
    if radii.dtype != torch.int32:
        radii = radii.int()
    dev = means2d.device
    st = stream_of(means2d)

    n_tiles = tile_width * tile_height
    tb = tile_n_bits(n_tiles)
    cb = max(C - 1, 0).bit_length()
    assert 32 + tb + cb <= 64, "camera/tile ids do not fit a 64-bit intersection key"

    if sort and tb + cb <= 32 and os.environ.get("GSPLAT_B200_SORT", "presort") != "global":
        tiles_per_gauss, isect_ids, flatten_ids, offsets, pending = _isect_presorted(
            means2d, radii, depths, camera_ids, C, N, P, packed, tile_size, tile_width, tile_height, tb, cb,
            _speculate)
        if _speculate:
            return tiles_per_gauss, isect_ids, flatten_ids, offsets, pending
        return tiles_per_gauss, isect_ids, flatten_ids

    tiles_per_gauss = torch.empty(radii.shape, dtype=torch.int32, device=dev)
    nblk = int(lib.gsb_isect_num_blocks(P))
    block_sums = torch.empty((max(nblk, 1),), dtype=torch.int32, device=dev)
    block_offsets = torch.empty((nblk + 1,), dtype=torch.int64, device=dev)
    lib.gsb_isect_count(P, ptr(means2d), ptr(radii), tile_size, tile_width, tile_height, ptr(tiles_per_gauss),
                        ptr(block_sums), ptr(block_offsets), block_offsets.data_ptr() + 8 * nblk, st)
    n_isects = int(block_offsets[-1].item())  # size read-back (isect_tiles.py:102, Intersect.cpp:77-78)
    isect_ids = torch.empty((n_isects,), dtype=torch.int64, device=dev)
    flatten_ids = torch.empty((n_isects,), dtype=torch.int32, device=dev)
    if n_isects == 0:
        return (tiles_per_gauss, isect_ids, flatten_ids) + ((None, None) if _speculate else ())
    lib.gsb_isect_emit(P, N, int(packed), ptr(means2d), ptr(radii), ptr(depths), ptr(camera_ids), tile_size,
                       tile_width, tile_height, tb, ptr(tiles_per_gauss), ptr(block_offsets), ptr(isect_ids),
                       ptr(flatten_ids), st)
    if sort:
        isect_ids, flatten_ids = _sort_pairs(isect_ids, flatten_ids, 32 + tb + cb)
    if _speculate:
        return tiles_per_gauss, isect_ids, flatten_ids, None, None
    return tiles_per_gauss, isect_ids, flatten_ids


# last n_isects seen per (device, cameras, tile grid, packed): sizes the next call's buffers without a read-back
_ISECT_ESTIMATE = {}


class _PendingCount:
    """n_isects of a speculatively sized isect_tiles call: its device-to-host copy is queued right after the
    count kernel and awaited only after the compositor has been launched, so the GPU never idles on it."""

    def __init__(self, n_dev: Tensor, capacity: int, key):
        self.n_dev, self.capacity, self.key = n_dev, capacity, key
        self.host = torch.empty((1,), dtype=torch.int64, pin_memory=True)
        self.host.copy_(n_dev, non_blocking=True)
        self.event = torch.cuda.Event()
        self.event.record()
        self.n = None

    def wait(self) -> int:
        if self.n is None:
            self.event.synchronize()
            self.n = int(self.host[0])
            _ISECT_ESTIMATE[self.key] = self.n
        return self.n


def _isect_presorted(means2d, radii, depths, camera_ids, C, N, P, packed, tile_size, tile_width, tile_height, tb, cb,
                     speculate):
    """Depth-presorted isect_tiles (include/gsplat_b200.h): sort the P rows by (camera, depth), emit in that
    order, then sort the intersections by their (camera | tile) bits only; offsets fall out of the last pass.

    `speculate`: size the intersection buffers from the previous call's count (+12.5 %) and keep the real count
    on the device; returns a `_PendingCount` the caller must `wait()` on (and redo the call exactly if the count
    exceeded the capacity). Without an estimate, or when not speculating, the count is read back here."""
    dev = means2d.device
    st = stream_of(means2d)
    tiles_per_gauss = torch.empty(radii.shape, dtype=torch.int32, device=dev)
    nblk = int(lib.gsb_isect_num_blocks(P))
    rows = torch.empty((P,), dtype=torch.int32, device=dev)
    # the row key is the depth bit pattern itself (positive floats order like their bits). Cameras need no bits
    # here: rows are emitted in global depth order and the (camera | tile) sort below is stable, so inside every
    # (camera, tile) bucket the intersections stay depth-ordered, ties by ascending row like the reference's sort.
    # tile counts are taken in natural order (coalesced); the depth-ordered passes only gather them
    records = torch.empty((P, 4), dtype=torch.int32, device=dev)
    # rows without tiles get an all-ones key: the emitting rows are one contiguous run of the sorted order
    row_keys = torch.empty((P,), dtype=torch.int32, device=dev)
    # the kernels that write sort keys also count their digits: the sorts then start with their first pass
    row_temp = _radix_temp(P, 32, dev, True) if (_own_sort() and P > 0) else None
    lib.gsb_isect_row_tiles_keys(P, N, int(packed), ptr(means2d), ptr(radii), ptr(depths), ptr(camera_ids), tile_size,
                                 tile_width, tile_height, tb, ptr(rows), ptr(tiles_per_gauss), ptr(records),
                                 ptr(row_keys), ptr(row_temp), st)
    sorted_keys, perm = _sort_pairs_u32(row_keys, rows, 32, None, row_temp)
    block_sums = torch.empty((max(nblk, 1),), dtype=torch.int32, device=dev)
    block_offsets = torch.empty((nblk + 1,), dtype=torch.int64, device=dev)
    # GSPLAT_B200_ISECT=bin: tile binning (csrc/isect_bin.cu) when the (camera, tile) bins fit its count matrix;
    # default: two radix passes over the intersections (measured faster so far: DESIGN.md 4)
    binned = (C * tile_width * tile_height <= int(lib.gsb_isect_bin_max_bins())
              and os.environ.get("GSPLAT_B200_ISECT", "radix") == "bin")
    # GSPLAT_B200_EMIT=chunks: emission over chunks of 8192 slots with the owner rows from a marker max-scan
    # (csrc/isect_bin.cu) instead of a binary search per slot over 256-row blocks; needs the records gathered into
    # depth order. Same output, measured equal (54 vs 56 us per launch, 1.674 vs 1.672 ms per step at cfg2): both
    # kernels are bound by their stores and histogram counts, not by the search. Default: the search.
    chunked = binned or (_own_sort() and os.environ.get("GSPLAT_B200_EMIT", "search") == "chunks")
    if chunked:
        sorted_records = torch.empty((P, 4), dtype=torch.int32, device=dev)
        lib.gsb_isect_count_perm_records(P, ptr(perm), ptr(records), ptr(sorted_records), ptr(block_sums),
                                         ptr(block_offsets), block_offsets.data_ptr() + 8 * nblk, st)
    else:
        lib.gsb_isect_count_perm(P, ptr(perm), None, None, tile_size, tile_width, tile_height,
                                 ptr(tiles_per_gauss), ptr(block_sums), ptr(block_offsets),
                                 block_offsets.data_ptr() + 8 * nblk, st)
    key = (dev.index, C, tile_width, tile_height, tile_size, bool(packed))
    estimate = _ISECT_ESTIMATE.get(key) if speculate else None
    pending = None
    if estimate is None:
        n_isects = int(block_offsets[-1].item())  # the one size read-back (isect_tiles.py:102, Intersect.cpp:77-78)
        _ISECT_ESTIMATE[key] = n_isects
        n_dev = None
    else:
        n_isects = min(estimate + estimate // 8 + 4096, (1 << 31) - 1)  # capacity, not the count
        pending = _PendingCount(block_offsets[nblk:], n_isects, key)
        n_dev = block_offsets.data_ptr() + 8 * nblk
    isect_ids = torch.empty((n_isects,), dtype=torch.int64, device=dev)
    offsets = torch.empty((C, tile_height, tile_width), dtype=torch.int32, device=dev)
    flatten_ids = torch.empty((n_isects,), dtype=torch.int32, device=dev)
    if binned:
        bins = C * tile_width * tile_height
        temp_bytes = int(lib.gsb_isect_bin_temp_bytes(n_isects, bins))
        temp = torch.empty((max(temp_bytes, 1),), dtype=torch.uint8, device=dev)
        lib.gsb_isect_bin(P, n_isects, n_dev if n_isects > 0 else None, ptr(sorted_records), ptr(perm),
                          ptr(sorted_keys), ptr(block_offsets), C, tile_width, tile_height, tb, ptr(temp), temp_bytes,
                          ptr(isect_ids), ptr(flatten_ids), ptr(offsets), st)
    else:
        tile_keys = torch.empty((n_isects,), dtype=torch.int32, device=dev)
        if n_isects > 0:
            # the hand-written sort takes the device-side count itself; the library sort needs the tail padded with
            # keys that sort last
            pad = n_dev if not _own_sort() else None
            tile_temp = _radix_temp(n_isects, tb + cb, dev, True) if (_own_sort() and tb + cb > 0) else None
            if chunked:
                lib.gsb_isect_emit_chunks(P, n_isects, n_dev, ptr(sorted_records), ptr(perm), ptr(block_offsets),
                                          tile_width, ptr(tile_keys), ptr(flatten_ids), ptr(tile_temp), tb + cb, st)
            else:
                lib.gsb_isect_emit_perm_hist(P, N, int(packed), ptr(perm), ptr(means2d), ptr(radii), ptr(camera_ids),
                                             tile_size, tile_width, tile_height, tb, ptr(tiles_per_gauss),
                                             ptr(block_offsets), ptr(tile_keys), ptr(flatten_ids), n_isects, pad,
                                             ptr(records), ptr(tile_temp), tb + cb, st)
            tile_keys, flatten_ids = _sort_pairs_u32(tile_keys, flatten_ids, tb + cb, n_dev, tile_temp)
        lib.gsb_isect_finalize_cap(n_isects, n_dev if n_isects > 0 else None, ptr(tile_keys), ptr(flatten_ids),
                                   ptr(depths), C, tile_width, tile_height, tb, ptr(isect_ids), ptr(offsets), st)
    if pending is None:
        # the finalize pass already produced isect_offset_encode's answer for these ids: hand it over with them
        isect_ids._b200_offsets = (C, tile_width, tile_height, tb, offsets)
    else:
        flatten_ids._b200_pending = pending
    return tiles_per_gauss, isect_ids, flatten_ids, offsets, pending


def _own_sort() -> bool:
    # GSPLAT_B200_SORT_IMPL=cub keeps the library sort for A/B measurements (tools/sort_bench.py)
    return os.environ.get("GSPLAT_B200_SORT_IMPL", "onesweep") != "cub"


def _radix_temp(n: int, end_bit: int, device, prepare: bool) -> Tensor:
    """Scratch of one onesweep sort; `prepare`: cleared now, so that the kernel that produces the keys can add its
    digit counts and the sort skips its histogram pass (gsb_radix_prepare)."""
    temp_bytes = int(lib.gsb_radix_temp_bytes(n, end_bit))
    temp = torch.empty((max(temp_bytes, 1),), dtype=torch.uint8, device=device)
    if prepare:
        lib.gsb_radix_prepare(n, end_bit, ptr(temp), temp.numel(), stream_of(temp))
    return temp


def _radix_sort(keys: Tensor, vals: Tensor, end_bit: int, n_dev: Optional[int] = None,
                prepared: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """Stable LSD radix sort of (uint32|int64 bit pattern, int32) pairs over bits [0, end_bit) with the hand-written
    onesweep kernels (csrc/radix_sort.cu). `n_dev`: device address of an int64 item count (<= keys.numel());
    `prepared`: a `_radix_temp(..., prepare=True)` buffer whose histograms the producer of the keys has filled."""
    import ctypes

    n = keys.numel()
    keys_alt = torch.empty_like(keys)
    vals_alt = torch.empty_like(vals)
    temp = prepared if prepared is not None else _radix_temp(n, end_bit, keys.device, False)
    in_alt = ctypes.c_int(0)
    if prepared is not None:
        fn = lib.gsb_radix_sort_pairs_u32_prepared
    else:
        fn = lib.gsb_radix_sort_pairs_u32 if keys.dtype == torch.int32 else lib.gsb_radix_sort_pairs_u64
    fn(n, n_dev, end_bit, ptr(keys), ptr(keys_alt), ptr(vals), ptr(vals_alt), ptr(temp), temp.numel(),
       ctypes.addressof(in_alt), stream_of(keys))
    return (keys_alt, vals_alt) if in_alt.value else (keys, vals)


def _sort_pairs(keys: Tensor, vals: Tensor, end_bit: int) -> Tuple[Tensor, Tensor]:
    """Stable LSD radix sort of (int64, int32) pairs over bits [0, end_bit) (radix_sort.cu:10-62)."""
    import ctypes

    if _own_sort():
        return _radix_sort(keys, vals, end_bit)
    n = keys.numel()
    keys_alt = torch.empty_like(keys)
    vals_alt = torch.empty_like(vals)
    temp_bytes = int(lib.gsb_sort_pairs_temp_bytes(n))
    temp = torch.empty((max(temp_bytes, 1),), dtype=torch.uint8, device=keys.device)
    in_alt = ctypes.c_int(0)
    lib.gsb_sort_pairs(n, end_bit, ptr(keys), ptr(keys_alt), ptr(vals), ptr(vals_alt), ptr(temp), temp_bytes,
                       ctypes.addressof(in_alt), stream_of(keys))
    return (keys_alt, vals_alt) if in_alt.value else (keys, vals)


def _sort_pairs_u32(keys: Tensor, vals: Tensor, end_bit: int, n_dev: Optional[int] = None,
                    prepared: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """Stable LSD radix sort of (uint32 bit pattern, int32) pairs over bits [0, end_bit)."""
    import ctypes

    if _own_sort():
        return _radix_sort(keys, vals, end_bit, n_dev, prepared)
    n = keys.numel()
    keys_alt = torch.empty_like(keys)
    vals_alt = torch.empty_like(vals)
    temp_bytes = int(lib.gsb_sort_pairs_u32_temp_bytes(n))
    temp = torch.empty((max(temp_bytes, 1),), dtype=torch.uint8, device=keys.device)
    in_alt = ctypes.c_int(0)
    lib.gsb_sort_pairs_u32(n, end_bit, ptr(keys), ptr(keys_alt), ptr(vals), ptr(vals_alt), ptr(temp), temp_bytes,
                           ctypes.addressof(in_alt), stream_of(keys))
    return (keys_alt, vals_alt) if in_alt.value else (keys, vals)


@torch.no_grad()
def isect_offset_encode(isect_ids: Tensor, n_cameras: int, tile_width: int, tile_height: int) -> Tensor:
    """First sorted-intersection index of every (camera, tile): int32 [C, tile_height, tile_width].

    Contract: triton_impl/isect_offset.py:9-35, cuda/_wrapper.py:422-438.
    """
    require_cuda(isect_ids)
    cached = getattr(isect_ids, "_b200_offsets", None)
    if cached is not None and cached[:4] == (n_cameras, tile_width, tile_height,
                                             tile_n_bits(tile_width * tile_height)):
        return cached[4]  # produced by isect_tiles' bucket scan for exactly these ids
    isect_ids = isect_ids.contiguous()
    offsets = torch.empty((n_cameras, tile_height, tile_width), dtype=torch.int32, device=isect_ids.device)
    lib.gsb_isect_offsets(isect_ids.numel(), ptr(isect_ids), n_cameras, tile_width, tile_height,
                          tile_n_bits(tile_width * tile_height), ptr(offsets), stream_of(isect_ids))
    return offsets


# --------------------------------------------------------------------------------------------
# rasterize_to_pixels
# --------------------------------------------------------------------------------------------
def rasterize_to_pixels(
    means2d: Tensor,  # [C, N, 2] or [nnz, 2]
    conics: Tensor,  # [C, N, 3] or [nnz, 3]
    colors: Tensor,  # [C, N, channels] or [nnz, channels]
    opacities: Tensor,  # [C, N] or [nnz]
    image_width: int,
    image_height: int,
    tile_size: int,
    isect_offsets: Tensor,  # [C, tile_height, tile_width]
    flatten_ids: Tensor,  # [n_isects]
    backgrounds: Optional[Tensor] = None,  # [C, channels]
    masks: Optional[Tensor] = None,  # [C, tile_height, tile_width]
    packed: bool = False,
    absgrad: bool = False,
) -> Tuple[Tensor, Tensor]:
    """Alpha-composites the sorted Gaussians of every tile.

    Returns `(render_colors [C,H,W,channels], render_alphas [C,H,W,1])`. Contract:
    triton_impl/_wrapper.py:185-296, cuda/_wrapper.py:441-575 (masks: True = render the tile,
    the upstream/CUDA polarity).
    """
    C = isect_offsets.size(0)
    if packed:
        nnz = means2d.size(0)
        assert means2d.shape == (nnz, 2), means2d.shape
        assert conics.shape == (nnz, 3), conics.shape
        assert colors.shape[0] == nnz, colors.shape
        assert opacities.shape == (nnz,), opacities.shape
    else:
        N = means2d.size(1)
        assert means2d.shape == (C, N, 2), means2d.shape
        assert conics.shape == (C, N, 3), conics.shape
        assert colors.shape[:2] == (C, N), colors.shape
        assert opacities.shape == (C, N), opacities.shape
    channels = colors.shape[-1]
    if backgrounds is not None:
        assert backgrounds.shape == (C, channels), backgrounds.shape
        backgrounds = backgrounds.contiguous()
    if masks is not None:
        assert masks.shape == isect_offsets.shape, masks.shape
        masks = masks.contiguous()
    if channels > 513 or channels == 0:
        raise ValueError(f"Unsupported number of color channels: {channels}")
    tile_height, tile_width = isect_offsets.shape[1:3]
    assert tile_height * tile_size >= image_height, f"Assert Failed: {tile_height} * {tile_size} >= {image_height}"
    assert tile_width * tile_size >= image_width, f"Assert Failed: {tile_width} * {tile_size} >= {image_width}"
    require_cuda(means2d, conics, colors, opacities, isect_offsets, flatten_ids, backgrounds, masks)

    pending = getattr(flatten_ids, "_b200_pending", None)  # speculatively sized ids (see _isect_presorted)

    def run(colors_c, backgrounds_c):
        ch = colors_c.shape[-1]
        target = next(c for c in _RASTER_CHANNELS if c >= ch)
        pad = target - ch
        if pad:
            colors_c = torch.nn.functional.pad(colors_c, (0, pad))
            if backgrounds_c is not None:
                backgrounds_c = torch.nn.functional.pad(backgrounds_c, (0, pad))
        rc, ra = _RasterizeToPixels.apply(
            means2d.contiguous(), conics.contiguous(), colors_c.contiguous(), opacities.contiguous(),
            backgrounds_c.contiguous() if backgrounds_c is not None else None, masks, image_width, image_height,
            tile_size, isect_offsets.contiguous(),
            flatten_ids.contiguous(), absgrad, pending,
        )
        return (rc[..., :ch] if pad else rc), ra

    top = _RASTER_CHANNELS[-1]
    if channels <= top:
        return run(colors, backgrounds)
    outs, alphas = [], None
    for lo in range(0, channels, top):
        bg = backgrounds[..., lo:lo + top] if backgrounds is not None else None
        rc, ra = run(colors[..., lo:lo + top], bg)
        outs.append(rc)
        alphas = ra if alphas is None else alphas
    return torch.cat(outs, dim=-1), alphas


# splat records packed ahead of the compositor call (rasterization() packs them on the caller's stream while the
# intersection pipeline runs on the side stream): ((means2d, conics, opacities data pointers, rows), records,
# zero-filled gradient buffer for the backward pass or None)
_PREPACKED = [None]


def drop_prepacked() -> None:
    """Forgets records packed ahead of a compositor call that never came (rasterization() calls this when it starts and
    when its compositor call returns or raises): a stale entry is matched by data pointers, and a later call's tensors
    may well live at the addresses a failed call's tensors were freed from."""
    _PREPACKED[0] = None


def _grad_widths(D: int, absgrad: bool):
    """floats per projected row the backward compositor accumulates into: v_means2d, v_conics, v_colors, v_opacities
    (, absgrad)"""
    return (2, 3, D, 1) + ((2,) if absgrad else ())


def prepack_splats(means2d: Tensor, conics: Tensor, opacities: Tensor, grad_channels: Optional[int] = None,
                   absgrad: bool = False) -> None:
    """Runs gsb_rasterize_pack now for the next rasterize_to_pixels() call on exactly these (contiguous) tensors.
    With `grad_channels` (the colour channels of that call, when a gradient will be asked for) the zero-filled buffer
    the backward compositor accumulates into is also made now: the caller's stream is idle here (the intersection
    pipeline runs on the side stream), in the backward pass the fill would sit on the critical path."""
    if not (means2d.is_contiguous() and conics.is_contiguous() and opacities.is_contiguous()):
        return
    if means2d.dtype != torch.float32 or conics.dtype != torch.float32 or opacities.dtype != torch.float32:
        return
    P = opacities.numel()
    rec = torch.empty((2, max(P, 1), 4), dtype=torch.float32, device=means2d.device)
    lib.gsb_rasterize_pack(P, ptr(means2d), ptr(conics), ptr(opacities), ptr(rec[0]), ptr(rec[1]), stream_of(means2d))
    zero = None
    if grad_channels is not None:
        zero = torch.zeros((P * sum(_grad_widths(grad_channels, absgrad)),), dtype=torch.float32, device=means2d.device)
    _PREPACKED[0] = ((means2d.data_ptr(), conics.data_ptr(), opacities.data_ptr(), P), rec, zero)


class _RasterizeToPixels(torch.autograd.Function):
    """cf. triton_impl/_wrapper.py:42-183, cuda/_wrapper.py:845-995."""

    @staticmethod
    def forward(ctx, means2d, conics, colors, opacities, backgrounds, masks, width, height, tile_size,
                isect_offsets, flatten_ids, absgrad, pending=None):
        C, tile_height, tile_width = isect_offsets.shape
        D = colors.shape[-1]
        dev = means2d.device
        st = stream_of(means2d)
        P = opacities.numel()
        n_isects = flatten_ids.numel()
        if flatten_ids.dtype != torch.int32:
            flatten_ids = flatten_ids.int()
        m8 = masks.to(torch.uint8) if masks is not None else None
        render_colors = torch.empty((C, height, width, D), dtype=torch.float32, device=dev)
        render_alphas = torch.empty((C, height, width, 1), dtype=torch.float32, device=dev)
        last_ids = torch.empty((C, height, width), dtype=torch.int32, device=dev)
        if tile_size != 16:
            # the tuned compositor is built around 16x16 tiles; every other tile size takes the plain kernels
            if pending is not None:
                n_isects = min(n_isects, pending.wait())
            lib.gsb_rasterize_fwd_generic(C, n_isects, D, ptr(means2d), ptr(conics), ptr(opacities), ptr(colors),
                                          ptr(backgrounds), ptr(m8), width, height, tile_size, tile_width, tile_height,
                                          ptr(isect_offsets), ptr(flatten_ids), ptr(render_colors), ptr(render_alphas),
                                          ptr(last_ids), st)
            ctx.pending = pending
            ctx.save_for_backward(means2d, conics, colors, backgrounds, m8, isect_offsets, flatten_ids, render_alphas,
                                  last_ids, opacities)
            ctx.width, ctx.height, ctx.tile_size, ctx.absgrad = width, height, tile_size, absgrad
            ctx.shapes = (conics.shape, opacities.shape)
            return render_colors, render_alphas
        hit, _PREPACKED[0] = _PREPACKED[0], None
        ctx.zero_buf = None
        if hit is not None and hit[0] == (means2d.data_ptr(), conics.data_ptr(), opacities.data_ptr(), P):
            rec = hit[1]
            if hit[2] is not None and hit[2].numel() == P * sum(_grad_widths(D, absgrad)):
                ctx.zero_buf = hit[2]  # filled while this stream waited for the intersections
        else:
            rec = torch.empty((2, max(P, 1), 4), dtype=torch.float32, device=dev)
            lib.gsb_rasterize_pack(P, ptr(means2d), ptr(conics), ptr(opacities), ptr(rec[0]), ptr(rec[1]), st)
        n_dev = ptr(pending.n_dev) if (pending is not None and pending.n is None) else None
        if pending is not None and pending.n is not None:
            n_isects = min(n_isects, pending.n)
        # per-intersection culling result (which 8x4 blocks of the tile the Gaussian reaches), kept for the backward
        # pass; only needed when a gradient will be asked for
        needs_grad = any(ctx.needs_input_grad[:5])
        block_masks = torch.empty((max(flatten_ids.numel(), 1),), dtype=torch.uint8, device=dev) if needs_grad else None
        lib.gsb_rasterize_fwd_cap(C, n_isects, n_dev, D, ptr(rec[0]), ptr(rec[1]), ptr(colors), ptr(backgrounds),
                                  ptr(m8), width, height, tile_size, tile_width, tile_height, ptr(isect_offsets),
                                  ptr(flatten_ids), ptr(render_colors), ptr(render_alphas), ptr(last_ids),
                                  ptr(block_masks), st)
        ctx.pending = pending
        ctx.save_for_backward(means2d, rec, colors, backgrounds, m8, isect_offsets, flatten_ids, render_alphas,
                              last_ids, block_masks)
        ctx.width, ctx.height, ctx.tile_size, ctx.absgrad = width, height, tile_size, absgrad
        ctx.shapes = (conics.shape, opacities.shape)
        return render_colors, render_alphas

    @staticmethod
    def backward(ctx, v_render_colors, v_render_alphas):
        generic = ctx.tile_size != 16
        if generic:
            (means2d, conics, colors, backgrounds, m8, isect_offsets, flatten_ids, render_alphas, last_ids,
             opacities) = ctx.saved_tensors
        else:
            (means2d, rec, colors, backgrounds, m8, isect_offsets, flatten_ids, render_alphas,
             last_ids, block_masks) = ctx.saved_tensors
        C, tile_height, tile_width = isect_offsets.shape
        D = colors.shape[-1]
        dev = means2d.device
        conics_shape, opac_shape = ctx.shapes
        v_render_colors = _f32(v_render_colors)
        v_render_alphas = _f32(v_render_alphas)
        # one zero-filled buffer, five views: the kernel accumulates into them with reductions
        P = means2d.numel() // 2
        widths = _grad_widths(D, ctx.absgrad)
        flat, ctx.zero_buf = getattr(ctx, "zero_buf", None), None  # good for one backward pass only
        if flat is None:
            flat = torch.zeros((P * sum(widths),), dtype=torch.float32, device=dev)
        parts = flat.split([P * w for w in widths])
        v_means2d = parts[0].view(means2d.shape)
        v_conics = parts[1].view(conics_shape)
        v_colors = parts[2].view(colors.shape)
        v_opacities = parts[3].view(opac_shape)
        v_abs = parts[4].view(means2d.shape) if ctx.absgrad else None
        n_isects = flatten_ids.numel() if ctx.pending is None else min(flatten_ids.numel(), ctx.pending.wait())
        if generic:
            lib.gsb_rasterize_bwd_generic(C, n_isects, D, ptr(means2d), ptr(conics), ptr(opacities), ptr(colors),
                                          ptr(backgrounds), ptr(m8), ctx.width, ctx.height, ctx.tile_size, tile_width,
                                          tile_height, ptr(isect_offsets), ptr(flatten_ids), ptr(render_alphas),
                                          ptr(last_ids), ptr(v_render_colors), ptr(v_render_alphas), ptr(v_abs),
                                          ptr(v_means2d), ptr(v_conics), ptr(v_colors), ptr(v_opacities),
                                          stream_of(means2d))
        else:
            lib.gsb_rasterize_bwd(C, n_isects, D, ptr(rec[0]), ptr(rec[1]), ptr(colors), ptr(backgrounds),
                                  ptr(m8), ctx.width, ctx.height, ctx.tile_size, tile_width, tile_height,
                                  ptr(isect_offsets), ptr(flatten_ids), ptr(render_alphas), ptr(last_ids),
                                  ptr(block_masks), ptr(v_render_colors), ptr(v_render_alphas), ptr(v_abs),
                                  ptr(v_means2d), ptr(v_conics), ptr(v_colors), ptr(v_opacities), stream_of(means2d))
        if ctx.absgrad:
            means2d.absgrad = v_abs
        v_backgrounds = None
        if ctx.needs_input_grad[4]:
            v_backgrounds = (v_render_colors * (1.0 - render_alphas).float()).sum(dim=(1, 2))
        return (v_means2d, v_conics, v_colors, v_opacities, v_backgrounds) + (None,) * 8


# --------------------------------------------------------------------------------------------
# rasterize_to_indices_in_range
# --------------------------------------------------------------------------------------------
@torch.no_grad()
def rasterize_to_indices_in_range(
    range_start: int,
    range_end: int,
    transmittances: Tensor,  # [C, image_height, image_width]
    means2d: Tensor,  # [C, N, 2]
    conics: Tensor,  # [C, N, 3]
    opacities: Tensor,  # [C, N]
    image_width: int,
    image_height: int,
    tile_size: int,
    isect_offsets: Tensor,  # [C, tile_height, tile_width]
    flatten_ids: Tensor,  # [n_isects]
) -> Tuple[Tensor, Tensor, Tensor]:
    """The (Gaussian, pixel, camera) index lists of the pairs the compositor blends for the sorted-intersection
    batches `[range_start, range_end)` of every tile. Contract: cuda/_wrapper.py:577-648 (the op the reference's
    pure-torch compositor `_torch_impl._rasterize_to_pixels` iterates on). Returns int64 `[M]` tensors."""
    C, N, _ = means2d.shape
    assert conics.shape == (C, N, 3), conics.shape
    assert opacities.shape == (C, N), opacities.shape
    assert isect_offsets.shape[0] == C, isect_offsets.shape
    assert transmittances.shape == (C, image_height, image_width), transmittances.shape
    tile_height, tile_width = isect_offsets.shape[1:3]
    assert tile_height * tile_size >= image_height, f"Assert Failed: {tile_height} * {tile_size} >= {image_height}"
    assert tile_width * tile_size >= image_width, f"Assert Failed: {tile_width} * {tile_size} >= {image_width}"
    require_cuda(transmittances, means2d, conics, opacities, isect_offsets, flatten_ids)
    dev, st = means2d.device, stream_of(means2d)
    range_end = int(min(range_end, 2**31 - 1))
    # every converted copy stays bound to a local until both launches are queued
    transmittances, means2d, conics, opacities = (_f32(t) for t in (transmittances, means2d, conics, opacities))
    isect_offsets, flatten_ids = isect_offsets.contiguous().int(), flatten_ids.contiguous().int()
    args = (C, N, flatten_ids.numel(), int(range_start), range_end, ptr(transmittances), ptr(means2d),
            ptr(conics), ptr(opacities), image_width, image_height, tile_size, tile_width, tile_height,
            ptr(isect_offsets), ptr(flatten_ids))
    counts = torch.empty((C * image_height * image_width,), dtype=torch.int32, device=dev)
    lib.gsb_rasterize_indices_count(*args, ptr(counts), st)
    ends = torch.cumsum(counts, dim=0, dtype=torch.int64)
    total = int(ends[-1].item()) if ends.numel() else 0
    starts = (ends - counts).contiguous()
    gaussian_ids = torch.empty((total,), dtype=torch.int64, device=dev)
    indices = torch.empty((total,), dtype=torch.int64, device=dev)
    if total > 0:
        lib.gsb_rasterize_indices_write(*args, ptr(starts), ptr(gaussian_ids), ptr(indices), st)
    per_camera = image_width * image_height
    return gaussian_ids, indices % per_camera, indices // per_camera
