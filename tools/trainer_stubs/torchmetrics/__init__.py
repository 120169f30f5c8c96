"""torchmetrics stand-in: PSNR and SSIM computed directly; LPIPS needs pretrained weights (no network) -> NaN."""
