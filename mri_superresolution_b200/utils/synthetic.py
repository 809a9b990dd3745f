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


VGG19_CONV_CHANNELS = [64, 64, 128, 128, 256, 256, 256, 256, 512, 512, 512, 512, 512, 512, 512, 512]


def make_vgg19_weights(seed: int = 1235):
    """Random stand-in for torchvision's ImageNet VGG19 convolution weights (the .pth is not on disk and there is no
    network): torchvision's own ``vgg19(weights=None)`` initialisation (kaiming_normal_(fan_out, relu)), seeded.
    Used by bench.py for the perceptual-loss workload (BASELINE.json configs[2]); same shapes, same FLOPs."""
    g = torch.Generator("cpu").manual_seed(seed)
    out, cin = [], 3
    for cout in VGG19_CONV_CHANNELS:
        std = (2.0 / (cout * 9)) ** 0.5
        out.append((torch.randn((cout, cin, 3, 3), generator=g) * std, 0.05 * torch.randn((cout,), generator=g)))
        cin = cout
    return out
