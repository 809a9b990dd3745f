"""Seeded synthetic LR/HR slice pairs used by bench.py and the tests.

HR: smoothed uniform noise, per-slice min-max to [0, 1]; LR: 2x2 area down-sample + noise, clamped --
the reference's data preparation makes LR the half-size image of HR (utils/extraction_utils.py:150-157).
"""
import torch
import torch.nn.functional as F


def make_pairs(n: int, h: int, w: int, seed: int = 1234, rank: int = 0, device="cpu"):
    """(lr [n,1,h,w], hr [n,1,2h,2w]) fp32; generated on CPU for reproducibility, then moved."""
    g = torch.Generator("cpu").manual_seed(seed + rank)
    hr = torch.rand(n, 1, 2 * h, 2 * w, generator=g)
    hr = F.avg_pool2d(hr, 9, stride=1, padding=4)
    lo, hi = hr.amin((2, 3), True), hr.amax((2, 3), True)
    hr = (hr - lo) / (hi - lo)
    lr = (F.avg_pool2d(hr, 2) + 0.02 * torch.randn(n, 1, h, w, generator=g)).clamp(0, 1)
    return lr.contiguous().to(device), hr.contiguous().to(device)
