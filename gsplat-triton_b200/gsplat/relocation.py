"""compute_relocation on the b200 backend; signature and semantics of the reference's gsplat/relocation.py:10-50
(MCMC strategy: opacity and scale of a Gaussian split into `ratios` copies, arXiv 2404.09591 eq. 9)."""
from typing import Tuple

import torch
from torch import Tensor

from .b200_impl._lib import lib, ptr, require_cuda, stream_of


def compute_relocation(opacities: Tensor, scales: Tensor, ratios: Tensor, binoms: Tensor) -> Tuple[Tensor, Tensor]:
    """opacities [N], scales [N,3], ratios [N] (clamped in place to 1..n_max, as the reference does),
    binoms [n_max,n_max] -> (new_opacities [N], new_scales [N,3])."""
    N = opacities.shape[0]
    n_max, _ = binoms.shape
    assert scales.shape == (N, 3), scales.shape
    assert ratios.shape == (N,), ratios.shape
    require_cuda(opacities, scales, ratios, binoms)
    opacities = opacities.detach().float().contiguous()
    scales = scales.detach().float().contiguous()
    ratios.clamp_(min=1, max=n_max)
    counts = ratios.int().contiguous()
    table = binoms.float().contiguous()
    new_opacities, new_scales = torch.empty_like(opacities), torch.empty_like(scales)
    lib.gsb_relocation(N, ptr(opacities), ptr(scales), ptr(counts), ptr(table), int(n_max), ptr(new_opacities),
                       ptr(new_scales), stream_of(opacities))
    return new_opacities, new_scales
