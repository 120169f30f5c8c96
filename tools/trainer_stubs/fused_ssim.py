"""fused_ssim stand-in: the same quantity (mean SSIM, 11x11 Gaussian window, sigma 1.5) in plain differentiable torch."""
import torch
import torch.nn.functional as F


def _window(channels, device, dtype):
    x = torch.arange(11, device=device, dtype=dtype) - 5
    g = torch.exp(-(x ** 2) / (2 * 1.5 ** 2))
    g = g / g.sum()
    return (g[:, None] * g[None, :]).expand(channels, 1, 11, 11).contiguous()


def fused_ssim(img1, img2, padding="same", train=True):
    """img: [B, C, H, W] in [0, 1] -> scalar mean SSIM."""
    C = img1.shape[1]
    w = _window(C, img1.device, img1.dtype)
    pad = 5 if padding == "same" else 0
    mu1, mu2 = F.conv2d(img1, w, padding=pad, groups=C), F.conv2d(img2, w, padding=pad, groups=C)
    s11 = F.conv2d(img1 * img1, w, padding=pad, groups=C) - mu1 * mu1
    s22 = F.conv2d(img2 * img2, w, padding=pad, groups=C) - mu2 * mu2
    s12 = F.conv2d(img1 * img2, w, padding=pad, groups=C) - mu1 * mu2
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu1 * mu2 + c1) * (2 * s12 + c2)) / ((mu1 * mu1 + mu2 * mu2 + c1) * (s11 + s22 + c2))).mean()
