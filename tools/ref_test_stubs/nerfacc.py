"""Stand-in for the two `nerfacc` functions the reference's pure-PyTorch compositor calls
(`gsplat/cuda/_torch_impl.py:483`: `accumulate` -> render_weight_from_alpha, accumulate_along_rays). The package is
not in this image and cannot be installed (no network). Plain differentiable torch ops on whatever device the inputs
live on; `ray_indices` must be grouped in ascending order, as `rasterize_to_indices_in_range` produces them.
Used only by tools/run_reference_tests.py (the reference's own test file on the b200 backend); never by the product."""
import torch


def render_weight_from_alpha(alphas, ray_indices=None, n_rays=None, **_):
    """w_i = alpha_i * T_i, T_i = prod over the earlier samples j of the same ray of (1 - alpha_j). -> (w, T).
    The segmented exclusive product is taken as exp of a segmented exclusive sum of log(1 - alpha) in float64
    (alpha <= 0.999 in the caller, so the logarithm is finite); float32 results."""
    logs = torch.log1p(-alphas.double())
    csum = torch.cumsum(logs, 0)
    excl = csum - logs
    first = torch.ones_like(ray_indices, dtype=torch.bool)
    first[1:] = ray_indices[1:] != ray_indices[:-1]
    pos = torch.arange(ray_indices.numel(), device=ray_indices.device)
    seg_start = torch.cummax(torch.where(first, pos, torch.zeros_like(pos)), 0).values
    trans = torch.exp(excl - excl[seg_start]).to(alphas.dtype)
    return alphas * trans, trans


def accumulate_along_rays(weights, values=None, ray_indices=None, n_rays=None):
    """Segmented sum of weights (* values) per ray -> [n_rays, D] (D = 1 without values)."""
    src = weights[:, None] if values is None else weights[:, None] * values
    out = torch.zeros((n_rays, src.shape[-1]), dtype=src.dtype, device=src.device)
    return out.index_add(0, ray_indices, src)
