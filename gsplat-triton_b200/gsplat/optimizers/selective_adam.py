"""SelectiveAdam on the b200 backend: same class, arguments and state layout as the reference's
gsplat/optimizers/selective_adam.py:6-94 (one tensor per parameter group, `step(visibility)`), with the fused update
in csrc/adam.cu through the C ABI."""
import torch

from ..b200_impl._lib import lib, ptr, require_cuda, stream_of


class SelectiveAdam(torch.optim.Adam):
    """Adam that updates only the Gaussians marked visible (parameters and moments of the others stay untouched),
    fused into one kernel per parameter tensor. No bias correction, as in the reference kernel."""

    def __init__(self, params, eps, betas):
        super().__init__(params=params, eps=eps, betas=betas)

    @torch.no_grad()
    def step(self, visibility):
        N = visibility.numel()
        mask = visibility.reshape(-1)
        mask = (mask if mask.dtype == torch.bool else mask != 0).to(torch.uint8).contiguous()
        for group in self.param_groups:
            lr, eps = group["lr"], group["eps"]
            beta1, beta2 = group["betas"]
            assert len(group["params"]) == 1, "more than one tensor in group"
            param = group["params"][0]
            if param.grad is None:
                continue
            state = self.state[param]
            if len(state) == 0:
                state["step"] = torch.tensor(0.0, dtype=torch.float32)
                state["exp_avg"] = torch.zeros_like(param, memory_format=torch.preserve_format)
                state["exp_avg_sq"] = torch.zeros_like(param, memory_format=torch.preserve_format)
            exp_avg, exp_avg_sq = state["exp_avg"], state["exp_avg_sq"]
            assert param.numel() % max(N, 1) == 0, "visibility must have one entry per Gaussian"
            assert param.dtype == torch.float32 and param.is_contiguous(), "float32 contiguous parameters"
            require_cuda(param, mask)
            grad = param.grad if param.grad.is_contiguous() else param.grad.contiguous()
            M = param.numel() // max(N, 1)
            lib.gsb_selective_adam(N, max(M, 1), ptr(param), ptr(grad.float()), ptr(exp_avg), ptr(exp_avg_sq),
                                   ptr(mask), float(lr), float(beta1), float(beta2), float(eps), stream_of(param))
