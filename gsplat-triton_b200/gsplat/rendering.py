"""`rasterization()` — the public entry of the hot path, with the `GSPLAT_BACKEND` switch.

Same signature, defaults, outputs and `meta` keys as the reference's gsplat/rendering.py:44-598.
`GSPLAT_BACKEND=b200` (default here) selects the sm_100a CUDA kernels of `gsplat.b200_impl`;
`triton` / `cuda` import the reference's own backends when they are importable (parity runs
with GSPLAT_REFERENCE_PATH set, see gsplat/__init__.py). Unknown names raise ValueError like
rendering.py:28-29.
"""
import math
import os
from typing import Dict, Optional, Tuple

import torch
from torch import Tensor
from typing_extensions import Literal

BACKEND = os.environ.get("GSPLAT_BACKEND", "b200")
if BACKEND == "b200":
    from .b200_impl._wrapper import (
        fully_fused_projection,
        isect_offset_encode,
        isect_tiles,
        rasterize_to_pixels,
        spherical_harmonics,
    )
elif BACKEND == "cuda":
    from .cuda._wrapper import (  # type: ignore  # reference backend via GSPLAT_REFERENCE_PATH
        fully_fused_projection,
        isect_offset_encode,
        isect_tiles,
        rasterize_to_pixels,
        spherical_harmonics,
    )
elif BACKEND == "triton":
    from .triton_impl._wrapper import (  # type: ignore  # reference backend via GSPLAT_REFERENCE_PATH
        fully_fused_projection,
        isect_offset_encode,
        isect_tiles,
        rasterize_to_pixels,
        spherical_harmonics,
    )
else:
    raise ValueError(f"Unknown GSPLAT_BACKEND: {BACKEND}")

from .distributed import (
    all_gather_int32,
    all_gather_int_host,
    all_gather_ints_host,
    all_gather_tensor_list,
    all_to_all_int32,
    all_to_all_rows,
    all_to_all_tensor_list,
    interleave_camera_major,
    host_trace,
    pack_rows,
    peer_exchange_available,
    PushExchange,
    SplitExchange,
    check_mailboxes,
    peer_exchange_packed,
    peer_exchange_rows,
    unpack_rows,
)

_RENDER_MODES = ("RGB", "D", "ED", "RGB+D", "RGB+ED")


def _rows_of(table: Tensor, ids: Tensor) -> Tensor:
    """`table[ids]` (packed rows -> their Gaussian's attribute); the b200 build uses an index_add backward."""
    if BACKEND == "b200":
        from .b200_impl._wrapper import gather_rows

        return gather_rows(table, ids)
    return table[ids]


def rasterization(
    means: Tensor,  # [N, 3]
    quats: Tensor,  # [N, 4]
    scales: Tensor,  # [N, 3]
    opacities: Tensor,  # [N]
    colors: Tensor,  # [(C,) N, D] or [(C,) N, K, 3]
    viewmats: Tensor,  # [C, 4, 4]
    Ks: Tensor,  # [C, 3, 3]
    width: int,
    height: int,
    near_plane: float = 0.01,
    far_plane: float = 1e10,
    radius_clip: float = 0.0,
    eps2d: float = 0.3,
    sh_degree: Optional[int] = None,
    packed: bool = True,
    tile_size: int = 16,
    backgrounds: Optional[Tensor] = None,
    render_mode: Literal["RGB", "D", "ED", "RGB+D", "RGB+ED"] = "RGB",
    sparse_grad: bool = False,
    absgrad: bool = False,
    rasterize_mode: Literal["classic", "antialiased"] = "classic",
    channel_chunk: int = 32,
    distributed: bool = False,
    camera_model: Literal["pinhole", "ortho", "fisheye"] = "pinhole",
    covars: Optional[Tensor] = None,
) -> Tuple[Tensor, Tensor, Dict]:
    """Rasterizes N 3D Gaussians to C image planes; differentiable w.r.t. the Gaussian parameters.

    Returns `(render_colors [C,H,W,X], render_alphas [C,H,W,1], meta)`; X is D, 1 (depth modes)
    or D+1 ("RGB+D"/"RGB+ED"). `meta` carries the intermediate tensors the training strategies
    read (`means2d` is a live autograd intermediate: `.retain_grad()`, `.absgrad`).
    See the reference docstring (rendering.py:70-239) for the argument semantics; they are kept.
    """
    meta: Dict = {}
    N = means.shape[0]
    C = viewmats.shape[0]
    device = means.device
    assert means.shape == (N, 3), means.shape
    if covars is None:
        assert quats.shape == (N, 4), quats.shape
        assert scales.shape == (N, 3), scales.shape
    else:
        assert covars.shape == (N, 3, 3), covars.shape
        quats, scales = None, None
        rows, cols = (0, 0, 0, 1, 1, 2), (0, 1, 2, 1, 2, 2)
        covars = covars[..., rows, cols]  # upper triangle, [N, 6]
    assert opacities.shape == (N,), opacities.shape
    assert viewmats.shape == (C, 4, 4), viewmats.shape
    assert Ks.shape == (C, 3, 3), Ks.shape
    assert render_mode in _RENDER_MODES, render_mode

    if sh_degree is None:
        # post-activation colours [N, D] or [C, N, D]
        assert (colors.dim() == 2 and colors.shape[0] == N) or (
            colors.dim() == 3 and colors.shape[:2] == (C, N)
        ), colors.shape
        if distributed:
            assert colors.dim() == 2, "Distributed mode only supports per-Gaussian colors."
    else:
        # SH coefficients [N, K, 3] or [C, N, K, 3]
        assert (colors.dim() == 3 and colors.shape[0] == N and colors.shape[2] == 3) or (
            colors.dim() == 4 and colors.shape[:2] == (C, N) and colors.shape[3] == 3
        ), colors.shape
        assert (sh_degree + 1) ** 2 <= colors.shape[-2], colors.shape
        if distributed:
            assert colors.dim() == 3, "Distributed mode only supports per-Gaussian colors."
    if absgrad:
        assert not distributed, "AbsGrad is not supported in distributed mode."

    # Distributed: Gaussians are sharded over ranks for projection, cameras for compositing.
    host_trace.mark()
    if distributed:
        world_rank = torch.distributed.get_rank()
        world_size = torch.distributed.get_world_size()
        if BACKEND == "b200" and peer_exchange_available() and os.environ.get("GSPLAT_B200_HOST_COUNTS", "1") == "1":
            N_world = all_gather_int_host(world_size, N)  # host-side channel: no GPU read-back per step
        else:
            N_world = all_gather_int32(world_size, N, device=device)
        host_trace.mark("us_counts")
        C_world = [C] * world_size  # every rank must bring the same number of cameras
        viewmats, Ks = all_gather_tensor_list(world_size, [viewmats, Ks])
        C = len(viewmats)  # global camera count from here on
        host_trace.mark("us_cameras")

    # b200, one node, unpacked rows, fused SH colours, GSPLAT_B200_EXCHANGE_PUSH=1: no put kernel, the projection and
    # the colour kernel store their rows straight into the camera owners' mailboxes. Off by default: measured no
    # faster than the put kernel (the producers then wait for the same NVLink stores).
    push = None
    if (distributed and not packed and BACKEND == "b200" and sh_degree is not None and colors.dim() == 3
            and not viewmats.requires_grad and peer_exchange_available()
            and os.environ.get("GSPLAT_B200_EXCHANGE_PUSH", "0") == "1"):
        from .b200_impl._wrapper import set_push_context

        push = PushExchange(C_world[world_rank], N_world, 3, opacities)
        set_push_context(push)
    # b200, one GPU, SH colours: the projection and the intersection pipeline (a dozen latency-bound kernels that leave
    # the memory system idle) run on a high-priority second stream, the HBM-bound colour kernels on the caller's; both
    # are joined before the compositor. Autograd replays the split backward: the projection's backward kernel then
    # shares the memory system with the colour backward instead of queueing behind it. Not with distributed=True: the
    # exchange sits between the colours and the intersections, nothing is left to run beside them (measured at 2
    # ranks: 1.86-1.90 ms with the second stream against 1.85 without).
    overlap, use_side = None, False
    dist_overlap = int(os.environ.get("GSPLAT_B200_OVERLAP_DIST", "0")) if distributed else 0
    if BACKEND == "b200" and (not distributed or dist_overlap) and sh_degree is not None:
        from .b200_impl._wrapper import overlap_enabled, side_stream

        use_side = overlap_enabled()
    if use_side:
        main, side = torch.cuda.current_stream(device), side_stream(device)
        start = torch.cuda.Event()
        start.record(main)

        # leaves enter the graph on the caller's stream (a free view), not on the side stream: autograd ties a leaf's
        # gradient accumulation to the stream current at its first use
        def enter(t):
            return t.view_as(t) if (t is not None and t.requires_grad and t.is_leaf) else t

        means, covars, quats, scales, viewmats, Ks = (enter(t) for t in (means, covars, quats, scales, viewmats, Ks))

    def project():
        return fully_fused_projection(
            means, covars, quats, scales, viewmats, Ks, width, height,
            eps2d=eps2d, packed=packed, near_plane=near_plane, far_plane=far_plane, radius_clip=radius_clip,
            sparse_grad=sparse_grad, calc_compensations=(rasterize_mode == "antialiased"),
            camera_model=camera_model,
        )

    try:
        if use_side:
            with torch.cuda.stream(side):
                side.wait_event(start)
                proj_results = project()
                overlap = torch.cuda.Event()
                overlap.record(side)
            main.wait_event(overlap)
            for t in proj_results:
                if isinstance(t, Tensor):
                    t.record_stream(main)  # allocated under the side stream, read on both
            if dist_overlap == 1:
                overlap = None  # experiment: only the projection (and so its backward) on the second stream
        else:
            proj_results = project()
    except Exception:
        if push is not None:
            set_push_context(None)
        raise
    if packed:
        camera_ids, gaussian_ids, radii, means2d, depths, conics, compensations = proj_results
        opacities = _rows_of(opacities, gaussian_ids)  # [nnz]
    else:
        radii, means2d, depths, conics, compensations = proj_results
        opacities = opacities.unsqueeze(0) if C == 1 else opacities.repeat(C, 1)  # [C, N]
        camera_ids, gaussian_ids = None, None
    if compensations is not None:
        opacities = opacities * compensations

    meta.update({
        "camera_ids": camera_ids,  # global camera ids
        "gaussian_ids": gaussian_ids,  # local Gaussian ids
        "radii": radii,
        "means2d": means2d,
        "depths": depths,
        "conics": conics,
        "opacities": opacities,
    })

    # b200, unpacked, one node, GSPLAT_B200_EXCHANGE_SPLIT=1: the forward exchange is split. The projected geometry
    # leaves for the camera owners NOW, on a side stream, while the colours are still being computed; the colours
    # follow under the receivers' tile intersection (which needs the geometry only). Off by default: on one 8xB200 box
    # the extra launches and stream hand-overs cost more than the overlap returns (cfg2 sharded over 8 ranks: 1.958 ms
    # against 1.921 ms per step).
    split = None
    if (push is None and distributed and not packed and BACKEND == "b200" and peer_exchange_available()
            and os.environ.get("GSPLAT_B200_EXCHANGE_SPLIT", "0") == "1"):
        D_out = 3 if sh_degree is not None else colors.shape[-1]
        split = SplitExchange(C_world[world_rank], N_world, D_out)
        split.start_geometry(means2d, depths, conics, opacities, radii)

    # Colours -> [C, N, D] or [nnz, D]
    if sh_degree is None:
        if packed:
            colors = _rows_of(colors, gaussian_ids) if colors.dim() == 2 else colors[camera_ids, gaussian_ids]
        elif colors.dim() == 2:
            colors = colors.expand(C, -1, -1)
    else:
        if BACKEND == "b200" and not packed and colors.dim() == 3 and not viewmats.requires_grad:
            # dirs, the radii > 0 mask, + 0.5 and the clamp folded into the SH kernels (one pass over [C, N])
            from .b200_impl._wrapper import colour_kernels_share_the_gpu, view_dependent_colors

            try:
                colour_kernels_share_the_gpu(use_side)  # the second stream's kernels need SM slots beside them
                colors = view_dependent_colors(sh_degree, means, viewmats, colors, radii)  # [C, N, 3]
            finally:
                colour_kernels_share_the_gpu(False)
                if push is not None:
                    set_push_context(None)
        elif BACKEND == "b200" and packed and colors.dim() == 3 and not viewmats.requires_grad:
            # same kernels over the dense [C, N] grid (mask scattered from the packed ids); only colours are gathered
            from .b200_impl._wrapper import colour_kernels_share_the_gpu, view_dependent_colors_packed

            try:
                colour_kernels_share_the_gpu(use_side)
                colors = view_dependent_colors_packed(sh_degree, means, viewmats, colors, camera_ids, gaussian_ids)
            finally:
                colour_kernels_share_the_gpu(False)
        else:
            camtoworlds = torch.inverse(viewmats)
            if packed:
                dirs = means[gaussian_ids, :] - camtoworlds[camera_ids, :3, 3]  # [nnz, 3]
                shs = colors[gaussian_ids, :, :] if colors.dim() == 3 else colors[camera_ids, gaussian_ids, :, :]
                colors = spherical_harmonics(sh_degree, dirs, shs, masks=radii > 0)  # [nnz, 3]
            else:
                dirs = means[None, :, :] - camtoworlds[:, None, :3, 3]  # [C, N, 3]
                if BACKEND == "b200" or colors.dim() == 4:
                    shs = colors  # b200 broadcasts [N,K,3] over cameras inside the kernel
                else:
                    shs = colors.expand(C, -1, -1, -1)
                colors = spherical_harmonics(sh_degree, dirs, shs, masks=radii > 0)  # [C, N, 3]
            colors = torch.clamp_min(colors + 0.5, 0.0)  # as rendering.py:408

    host_trace.mark("us_project_colors")
    # Distributed: hand every projected Gaussian to the rank that owns its camera.
    if distributed:
        D = colors.shape[-1]
        if packed and BACKEND == "b200" and peer_exchange_available() and world_size <= 16:
            # rows per destination rank: host data already (they came back with nnz), exchanged on the host board;
            # then ONE fused pack + NVLink put per direction, ids rebased inside the kernel
            from .b200_impl._wrapper import packed_camera_counts

            per_cam = packed_camera_counts(camera_ids, C)
            C_l = C_world[world_rank]
            cnts = [sum(per_cam[i * C_l:(i + 1) * C_l]) for i in range(world_size)]
            matrix = all_gather_ints_host(world_size, cnts)
            means2d, depths, conics, opacities, colors, radii, camera_ids, gaussian_ids = peer_exchange_packed(
                means2d, depths, conics, opacities, colors, radii, camera_ids, gaussian_ids, matrix, C_l,
                sum(N_world[:world_rank]))
            C = C_l
        elif packed:
            cnts = torch.bincount(camera_ids, minlength=C).view(world_size, -1).sum(dim=1).tolist()  # one read-back
            recv = all_to_all_int32(world_size, cnts, device=device)
            rows = pack_rows(means2d, depths, conics, opacities, colors, radii)
            rows = all_to_all_rows(rows, cnts, recv)
            means2d, depths, conics, opacities, colors, radii = unpack_rows(rows, D)
            # camera ids global -> local to the destination; Gaussian ids local -> global
            cnts_t = torch.tensor(cnts, device=device)
            cam_base = torch.tensor([sum(C_world[:i]) for i in range(world_size)], device=device)
            gau_base = sum(N_world[:world_rank])
            ids = torch.stack([camera_ids - cam_base.repeat_interleave(cnts_t), gaussian_ids + gau_base], dim=-1)
            ids = all_to_all_rows(ids, cnts, recv)
            camera_ids, gaussian_ids = ids[:, 0].contiguous(), ids[:, 1].contiguous()
            C = C_world[world_rank]
        elif push is not None:
            C = C_world[world_rank]
            means2d, depths, conics, opacities, colors, radii = push.finish(means2d, depths, conics, opacities, colors,
                                                                            radii)
        elif split is not None:
            C = C_world[world_rank]
            split.start_colors(colors)
            means2d, depths, conics, opacities, colors, radii = split.finish(means2d, depths, conics, opacities,
                                                                             colors, radii)
            if render_mode != "RGB":  # the depth channel is concatenated to the colours below: they are needed now
                split.complete_colors()
        elif BACKEND == "b200" and peer_exchange_available():
            # fused pack + NVLink peer stores straight into the camera owner's [C_local, N_total] table
            C = C_world[world_rank]
            means2d, depths, conics, opacities, colors, radii = peer_exchange_rows(
                means2d, depths, conics, opacities, colors, radii, C, N_world)
        else:
            C = C_world[world_rank]
            rows = pack_rows(means2d.flatten(0, 1), depths.flatten(0, 1), conics.flatten(0, 1),
                             opacities.flatten(0, 1), colors.flatten(0, 1), radii.flatten(0, 1))
            rows = all_to_all_rows(rows, [C_i * N for C_i in C_world], [C * N_i for N_i in N_world])
            rows = interleave_camera_major(rows, C, N_world)  # [C, N_total, 8+D]
            N_total = rows.shape[1]
            means2d, depths, conics, opacities, colors, radii = unpack_rows(rows.reshape(C * N_total, -1), D)
            means2d, conics, colors = (t.reshape(C, N_total, -1) for t in (means2d, conics, colors))
            depths, opacities, radii = (t.reshape(C, N_total) for t in (depths, opacities, radii))

    host_trace.mark("us_exchange")
    if render_mode in ("RGB+D", "RGB+ED"):
        colors = torch.cat((colors, depths[..., None]), dim=-1)
        if backgrounds is not None:
            backgrounds = torch.cat([backgrounds, torch.zeros(C, 1, device=backgrounds.device)], dim=-1)
    elif render_mode in ("D", "ED"):
        colors = depths[..., None]
        if backgrounds is not None:
            backgrounds = torch.zeros(C, 1, device=backgrounds.device)

    # Tile intersection, sort, per-tile ranges
    tile_width = math.ceil(width / float(tile_size))
    tile_height = math.ceil(height / float(tile_size))
    def intersect(speculate):
        out = isect_tiles(
            means2d, radii, depths, tile_size, tile_width, tile_height,
            packed=packed, n_cameras=C, camera_ids=camera_ids, gaussian_ids=gaussian_ids,
            **({"_speculate": True} if speculate else {}),
        )
        if speculate and out[4] is not None:
            return out  # (tiles_per_gauss, isect_ids, flatten_ids, isect_offsets, pending count)
        return out[0], out[1], out[2], isect_offset_encode(out[1], C, tile_width, tile_height), None

    # b200: intersection buffers sized from the previous call's count, the real count is read back only after
    # the compositor has been queued (no GPU bubble at the reference's `.item()`, isect_tiles.py:102)
    speculate = BACKEND == "b200" and os.environ.get("GSPLAT_B200_SPECULATE", "1") == "1"
    if overlap is not None:
        from .b200_impl._wrapper import drop_prepacked, prepack_splats

        main, side = torch.cuda.current_stream(device), side_stream(device)
        drop_prepacked()  # leftovers of a call that failed between packing and compositing
        if tile_size == 16 and colors.shape[-1] <= channel_chunk:
            # the compositor's splat records (and, when a gradient will be asked for, its zero-filled gradient
            # buffer), also beside the side stream
            wants_grad = (torch.is_grad_enabled() and os.environ.get("GSPLAT_B200_PREZERO", "1") == "1"
                          and any(t.requires_grad for t in (means2d, conics, colors, opacities)))
            prepack_splats(means2d, conics, opacities, grad_channels=colors.shape[-1] if wants_grad else None,
                           absgrad=absgrad)
        with torch.cuda.stream(side):
            side.wait_event(overlap)
            tiles_per_gauss, isect_ids, flatten_ids, isect_offsets, pending = intersect(speculate)
            isect_done = torch.cuda.Event()
            isect_done.record(side)
        main.wait_event(isect_done)
        for t in (tiles_per_gauss, isect_ids, flatten_ids, isect_offsets):
            t.record_stream(main)  # allocated under the side stream, read by the compositor and its backward
    else:
        tiles_per_gauss, isect_ids, flatten_ids, isect_offsets, pending = intersect(speculate)
    host_trace.mark("us_isect")
    if split is not None:
        split.complete_colors()  # the colours arrived while the intersections were being sorted

    # Composite
    def composite(colors_c, backgrounds_c):
        return rasterize_to_pixels(
            means2d, conics, colors_c, opacities, width, height, tile_size, isect_offsets, flatten_ids,
            backgrounds=backgrounds_c, packed=packed, absgrad=absgrad,
        )

    def composite_all():
        if colors.shape[-1] > channel_chunk:
            chunks, alphas = [], None
            for lo in range(0, colors.shape[-1], channel_chunk):
                bg = backgrounds[..., lo:lo + channel_chunk] if backgrounds is not None else None
                rc, ra = composite(colors[..., lo:lo + channel_chunk], bg)
                chunks.append(rc)
                alphas = ra if alphas is None else alphas
            return torch.cat(chunks, dim=-1), alphas
        return composite(colors, backgrounds)

    try:
        render_colors, render_alphas = composite_all()
    finally:
        if overlap is not None:
            drop_prepacked()  # consumed by the compositor call above, or never to be used
    if pending is not None:
        n_isects = pending.wait()
        if distributed and BACKEND == "b200":
            check_mailboxes()  # the wait kernel of this step's exchange ran before the count that was just read back
        if n_isects > pending.capacity:  # the estimate was too small: redo this part with exact sizes
            tiles_per_gauss, isect_ids, flatten_ids, isect_offsets, pending = intersect(False)
            render_colors, render_alphas = composite_all()
        else:
            isect_ids, flatten_ids = isect_ids[:n_isects], flatten_ids[:n_isects]
    meta.update({
        "tile_width": tile_width,
        "tile_height": tile_height,
        "tiles_per_gauss": tiles_per_gauss,
        "isect_ids": isect_ids,
        "flatten_ids": flatten_ids,
        "isect_offsets": isect_offsets,
        "width": width,
        "height": height,
        "tile_size": tile_size,
        "n_cameras": C,
    })
    host_trace.mark("us_composite")

    if render_mode in ("ED", "RGB+ED"):
        render_colors = torch.cat(
            [render_colors[..., :-1], render_colors[..., -1:] / render_alphas.clamp(min=1e-10)], dim=-1
        )
    return render_colors, render_alphas, meta


def __getattr__(name):
    """`gsplat.rendering._rasterization`: the reference's pure-PyTorch twin of rasterization() (its
    `gsplat/rendering.py:601-760`, what its own `tests/test_rasterization.py` compares a backend against). It is the
    REFERENCE's function, taken at first use from the reference checkout named by GSPLAT_REFERENCE_PATH and evaluated
    in this module's namespace — exactly what happens when the reference's own rendering.py grows a `b200` branch
    (INTEGRATION.md option A): its backend-dispatched calls (`isect_tiles`, `isect_offset_encode`,
    `spherical_harmonics`) resolve to this backend, projection and compositing stay the reference's `_torch_impl`."""
    if name != "_rasterization":
        raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
    import ast

    ref = os.environ.get("GSPLAT_REFERENCE_PATH")
    path = os.path.join(ref, "gsplat", "rendering.py") if ref else None
    if not path or not os.path.exists(path):
        raise AttributeError("gsplat.rendering._rasterization is the reference's own function: set "
                             "GSPLAT_REFERENCE_PATH to a checkout of gsplat-triton (there is no copy in this package)")
    source = open(path).read()
    node = next(n for n in ast.parse(source).body if isinstance(n, ast.FunctionDef) and n.name == "_rasterization")
    code = compile(ast.Module(body=[node], type_ignores=[]), path, "exec")
    exec(code, globals())
    return globals()["_rasterization"]
