"""`quat_scale_to_covar_preci` — off the hot path (the fused projection builds its covariances in-kernel); kept as a
plain differentiable torch function because the reference's densification code imports it from the package root
(`gsplat/strategy/ops.py:8`, used by `inject_noise_to_position` :352). Contract: reference gsplat/cuda/_wrapper.py:121-155,
math: gsplat/cuda/_torch_impl.py:10-75."""
from typing import Optional, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor


def quat_scale_to_covar_preci(
    quats: Tensor,  # [N, 4] (w, x, y, z), need not be normalised
    scales: Tensor,  # [N, 3]
    compute_covar: bool = True,
    compute_preci: bool = True,
    triu: bool = False,
) -> Tuple[Optional[Tensor], Optional[Tensor]]:
    """Covariance Sigma = (R S)(R S)^T and precision (R S^-1)(R S^-1)^T; [N,3,3], or the upper triangle [N,6]."""
    assert quats.dim() == 2 and quats.size(1) == 4, quats.size()
    assert scales.dim() == 2 and scales.size(1) == 3, scales.size()
    w, x, y, z = torch.unbind(F.normalize(quats, p=2, dim=-1), dim=-1)
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y),
    ], dim=-1).reshape(-1, 3, 3)

    def outer(M):
        out = M @ M.transpose(-1, -2)
        if triu:
            rows, cols = (0, 0, 0, 1, 1, 2), (0, 1, 2, 1, 2, 2)
            out = out[..., rows, cols]
        return out

    covars = outer(R * scales[:, None, :]) if compute_covar else None
    precis = outer(R * (1.0 / scales[:, None, :])) if compute_preci else None
    return covars, precis
