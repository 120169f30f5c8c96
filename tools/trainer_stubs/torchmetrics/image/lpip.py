import torch


class LearnedPerceptualImagePatchSimilarity(torch.nn.Module):
    """LPIPS needs pretrained AlexNet / VGG weights, which cannot be downloaded here: reports NaN."""

    def __init__(self, *a, **k):
        super().__init__()

    def forward(self, preds, target):
        return torch.full((), float("nan"), device=preds.device)
