"""`accumulate()` and the batched pure-PyTorch compositor of the reference, without `nerfacc`.

The reference's `gsplat.accumulate` (gsplat/cuda/_torch_impl.py:432-519) is plain PyTorch glue around two `nerfacc`
primitives (a third-party package, not a dependency of either repo): it is the "playground" compositor that relies on
autograd, fed by the index lists of `rasterize_to_indices_in_range`, and `_rasterize_to_pixels` (:522-617) loops it over
batches. Both are restated here on torch ops only (segmented transmittance = exclusive cumulative sum of log(1 - alpha)
in float64, segmented sums = `index_add_`), so the reference's slow-but-simple compositor runs on top of OUR index kernel
(`csrc/indices.cu`). Neither function is on the hot path; nothing in `rasterization()` calls them.
"""
from typing import Optional, Tuple

import torch
from torch import Tensor


def _weights_along_rays(alphas: Tensor, ray_ids: Tensor) -> Tensor:
    """w_i = alpha_i * prod_{j < i, same ray} (1 - alpha_j) for samples stored ray by ray (what
    nerfacc.render_weight_from_alpha computes). `ray_ids` [M] must be grouped (equal ids contiguous), as the lists of
    rasterize_to_indices_in_range are."""
    if alphas.numel() == 0:
        return alphas
    log_keep = torch.log1p(-alphas.double())  # alpha <= 0.999: finite
    before = torch.cumsum(log_keep, dim=0) - log_keep  # exclusive, over the whole list
    first = torch.ones_like(ray_ids, dtype=torch.bool)
    first[1:] = ray_ids[1:] != ray_ids[:-1]
    segment = torch.cumsum(first.long(), dim=0) - 1
    trans = torch.exp(before - before[first][segment])  # restart at every ray's first sample
    return alphas * trans.to(alphas.dtype)


def accumulate(
    means2d: Tensor,  # [C, N, 2]
    conics: Tensor,  # [C, N, 3]
    opacities: Tensor,  # [C, N]
    colors: Tensor,  # [C, N, channels]
    gaussian_ids: Tensor,  # [M]
    pixel_ids: Tensor,  # [M]
    camera_ids: Tensor,  # [M]
    image_width: int,
    image_height: int,
) -> Tuple[Tensor, Tensor]:
    """Alpha compositing of the (Gaussian, pixel, camera) pairs in pure PyTorch; same arguments, outputs and arithmetic
    as the reference's `accumulate` (_torch_impl.py:432-519): renders [C, H, W, channels], alphas [C, H, W, 1]."""
    C = means2d.shape[0]
    channels = colors.shape[-1]
    pixel_xy = torch.stack([pixel_ids % image_width, pixel_ids // image_width], dim=-1) + 0.5  # [M, 2]
    deltas = pixel_xy - means2d[camera_ids, gaussian_ids]
    c = conics[camera_ids, gaussian_ids]
    sigmas = 0.5 * (c[:, 0] * deltas[:, 0] ** 2 + c[:, 2] * deltas[:, 1] ** 2) + c[:, 1] * deltas[:, 0] * deltas[:, 1]
    alphas = torch.clamp_max(opacities[camera_ids, gaussian_ids] * torch.exp(-sigmas), 0.999)

    rays = camera_ids * image_height * image_width + pixel_ids
    n_rays = C * image_height * image_width
    weights = _weights_along_rays(alphas, rays)
    renders = torch.zeros((n_rays, channels), dtype=colors.dtype, device=colors.device)
    renders = renders.index_add(0, rays, weights[:, None] * colors[camera_ids, gaussian_ids])
    acc = torch.zeros((n_rays,), dtype=weights.dtype, device=weights.device).index_add(0, rays, weights)
    return renders.reshape(C, image_height, image_width, channels), acc.reshape(C, image_height, image_width, 1)


def rasterize_to_pixels_torch(
    means2d: Tensor,  # [C, N, 2]
    conics: Tensor,  # [C, N, 3]
    colors: Tensor,  # [C, N, channels]
    opacities: Tensor,  # [C, N]
    image_width: int,
    image_height: int,
    tile_size: int,
    isect_offsets: Tensor,  # [C, tile_height, tile_width]
    flatten_ids: Tensor,  # [n_isects]
    backgrounds: Optional[Tensor] = None,  # [C, channels]
    batch_per_iter: int = 100,
) -> Tuple[Tensor, Tensor]:
    """The reference's `_rasterize_to_pixels` (_torch_impl.py:522-617): batches of `tile_size^2` sorted intersections,
    the pairs of each batch from `rasterize_to_indices_in_range` (our CUDA kernel), composited by `accumulate` with
    the transmittance carried from batch to batch; gradients by autograd."""
    from ._wrapper import rasterize_to_indices_in_range

    C = means2d.shape[0]
    n_isects = len(flatten_ids)
    device = means2d.device
    render_colors = torch.zeros((C, image_height, image_width, colors.shape[-1]), device=device)
    render_alphas = torch.zeros((C, image_height, image_width, 1), device=device)
    block_size = tile_size * tile_size
    ends = torch.cat([isect_offsets.flatten()[1:], torch.tensor([n_isects], device=device, dtype=isect_offsets.dtype)])
    max_range = int((ends - isect_offsets.flatten()).max().item()) if isect_offsets.numel() else 0
    num_batches = (max_range + block_size - 1) // block_size
    for step in range(0, num_batches, batch_per_iter):
        transmittances = 1.0 - render_alphas[..., 0]
        gs_ids, pixel_ids, camera_ids = rasterize_to_indices_in_range(
            step, step + batch_per_iter, transmittances, means2d, conics, opacities, image_width, image_height,
            tile_size, isect_offsets, flatten_ids)
        if len(gs_ids) == 0:
            break
        renders_step, accs_step = accumulate(means2d, conics, opacities, colors, gs_ids, pixel_ids, camera_ids,
                                             image_width, image_height)
        render_colors = render_colors + renders_step * transmittances[..., None]
        render_alphas = render_alphas + accs_step * transmittances[..., None]
    if backgrounds is not None:
        render_colors = render_colors + backgrounds[:, None, None, :] * (1.0 - render_alphas)
    return render_colors, render_alphas
