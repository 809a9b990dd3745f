"""Functional fp32 restatement of the reference losses (TEST INFRASTRUCTURE, see oracle/__init__.py).

Follows /root/reference/utils/losses.py.  Pinned against the reference by
``oracle/make_golden.py`` -> ``tests/golden/*.npz``.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

VGG_MEAN = (0.485, 0.456, 0.406)     # utils/losses.py:7
VGG_STD = (0.229, 0.224, 0.225)      # utils/losses.py:8

# torchvision vgg19 "E" configuration restricted to features[:36] (utils/losses.py:90-95):
# 16 x (conv3x3 + ReLU), max-pool after convs 2, 4, 8 and 12; the pool after conv 16 is index 36 -> excluded.
VGG19_CFG = (64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512)


def gaussian_taps(window_size: int = 11, sigma: float = 1.5) -> torch.Tensor:
    """1-D Gaussian normalised to sum 1, utils/losses.py:10-18."""
    c = torch.arange(window_size, dtype=torch.float32) - (window_size // 2)
    g = torch.exp(-(c ** 2) / (2 * sigma ** 2))
    return g / g.sum()


def gaussian_window2d(window_size: int = 11, sigma: float = 1.5) -> torch.Tensor:
    """Outer product window, utils/losses.py:20-25 (fp32 ``mm`` of the 1-D taps)."""
    t = gaussian_taps(window_size, sigma).unsqueeze(1)
    return t.mm(t.t())


def ssim_map(x: torch.Tensor, y: torch.Tensor, window_size: int = 11, sigma: float = 1.5,
             val_range: float = 1.0) -> torch.Tensor:
    """Per-pixel SSIM map, utils/losses.py:41-70 (zero-padded 'same' depthwise 11x11 filters)."""
    x = x.to(torch.float32)
    y = y.to(torch.float32)
    ch = x.size(1)
    w = gaussian_window2d(window_size, sigma).to(x.device).expand(ch, 1, window_size, window_size).contiguous()
    pad = window_size // 2
    blur = lambda t: F.conv2d(t, w, padding=pad, groups=ch)
    mu1, mu2 = blur(x), blur(y)
    s11 = blur(x * x) - mu1 * mu1
    s22 = blur(y * y) - mu2 * mu2
    s12 = blur(x * y) - mu1 * mu2
    c1 = (0.01 * val_range) ** 2
    c2 = (0.03 * val_range) ** 2
    return ((2 * mu1 * mu2 + c1) * (2 * s12 + c2)) / ((mu1 * mu1 + mu2 * mu2 + c1) * (s11 + s22 + c2))


def ssim(x: torch.Tensor, y: torch.Tensor, window_size: int = 11, sigma: float = 1.5, val_range: float = 1.0,
         size_average: bool = True) -> torch.Tensor:
    """utils/losses.py:27-81 (mean over everything, or per-sample mean when size_average=False)."""
    m = ssim_map(x, y, window_size, sigma, val_range)
    res = m.mean() if size_average else m.mean(1).mean(1).mean(1)
    if x.dtype != torch.float32 and x.dtype == y.dtype:      # :78-79
        res = res.to(x.dtype)
    return res


def vgg_features(vgg_params: Sequence[Tuple[torch.Tensor, torch.Tensor]], x: torch.Tensor,
                 cfg: Sequence = VGG19_CFG) -> torch.Tensor:
    """VGGFeatureExtractor.forward, utils/losses.py:105-118, over explicit (weight, bias) pairs."""
    if x.shape[1] == 1:
        x = x.repeat(1, 3, 1, 1)                                                    # :107-108
    mean = torch.tensor(VGG_MEAN, dtype=x.dtype, device=x.device).view(1, 3, 1, 1)
    std = torch.tensor(VGG_STD, dtype=x.dtype, device=x.device).view(1, 3, 1, 1)
    x = (x - mean) / std                                                            # :112-114
    i = 0
    for item in cfg:
        if item == "M":
            x = F.max_pool2d(x, 2)
        else:
            w, b = vgg_params[i]
            i += 1
            x = F.relu(F.conv2d(x, w, b, padding=1))
    return x


def vgg_features_bf16_storage(vgg_params: Sequence[Tuple[torch.Tensor, torch.Tensor]], x: torch.Tensor,
                              cfg: Sequence = VGG19_CFG) -> torch.Tensor:
    """vgg_features with every stored tensor (normalised input, weights, each conv + ReLU output) rounded to bf16 and fp32
    arithmetic in between: the operand rounding of the tensor-core path, used by the GPU tests to measure the bf16 noise
    floor of the perceptual gradient (test infrastructure only)."""
    from oracle.unet_oracle import _q
    if x.shape[1] == 1:
        x = x.repeat(1, 3, 1, 1)
    mean = torch.tensor(VGG_MEAN, dtype=x.dtype, device=x.device).view(1, 3, 1, 1)
    std = torch.tensor(VGG_STD, dtype=x.dtype, device=x.device).view(1, 3, 1, 1)
    x = _q((x - mean) / std)
    i = 0
    for item in cfg:
        if item == "M":
            x = F.max_pool2d(x, 2)
        else:
            w, b = vgg_params[i]
            i += 1
            x = _q(F.relu(F.conv2d(x, _q(w), b, padding=1)))
    return x


def make_vgg_params(seed: int = 1235, cfg: Sequence = VGG19_CFG) -> List[Tuple[torch.Tensor, torch.Tensor]]:
    """Random stand-in for the ImageNet VGG19 weights (the .pth is not on disk and there is no
    network): torchvision's own init for ``vgg19(weights=None)`` --
    kaiming_normal_(fan_out, relu) weights, zero biases -- restated with an explicit generator."""
    g = torch.Generator("cpu").manual_seed(seed)
    out = []
    cin = 3
    for item in cfg:
        if item == "M":
            continue
        std = (2.0 / (item * 9)) ** 0.5
        w = torch.randn((item, cin, 3, 3), generator=g, dtype=torch.float32) * std
        b = 0.05 * torch.randn((item,), generator=g, dtype=torch.float32)   # non-zero so the bias path is exercised
        out.append((w, b))
        cin = item
    return out


def combined_loss(output: torch.Tensor, target: torch.Tensor, ssim_weight: float = 0.5,
                  perceptual_weight: float = 0.0, vgg_params=None, perceptual_loss_type: str = "l1",
                  window_size: int = 11, sigma: float = 1.5, val_range: float = 1.0,
                  components: Optional[Dict[str, torch.Tensor]] = None) -> torch.Tensor:
    """CombinedLoss.forward, utils/losses.py:200-240 (validation of the weights: :166-171)."""
    if not (0 <= ssim_weight <= 1):
        raise ValueError("ssim_weight must be between 0 and 1")
    if not (0 <= perceptual_weight <= 1):
        raise ValueError("perceptual_weight must be between 0 and 1")
    if ssim_weight + perceptual_weight > 1:
        raise ValueError("Sum of ssim_weight and perceptual_weight cannot exceed 1")
    l1_weight = 1.0 - ssim_weight - perceptual_weight                              # :175
    total = 0.0
    if l1_weight > 0:                                                              # :205-208
        l1 = (output - target).abs().mean()
        total = total + l1_weight * l1
        if components is not None:
            components["l1"] = l1.detach()
    if ssim_weight > 0:                                                            # :211-226
        s = torch.clamp(ssim(output, target, window_size, sigma, val_range), min=0.0, max=1.0)
        total = total + ssim_weight * (1 - s)
        if components is not None:
            components["ssim"] = s.detach()
    if perceptual_weight > 0:                                                      # :229-236, :138-151
        fg = vgg_features(vgg_params, output)
        with torch.no_grad():
            ft = vgg_features(vgg_params, target)
        if perceptual_loss_type == "l1":
            perc = (fg - ft).abs().mean()
        elif perceptual_loss_type in ("l2", "mse"):
            perc = ((fg - ft) ** 2).mean()
        else:
            raise ValueError(f"Unsupported loss type for PerceptualLoss: {perceptual_loss_type}")
        total = total + perceptual_weight * perc
        if components is not None:
            components["perceptual"] = perc.detach()
    return total


def psnr(a: torch.Tensor, b: torch.Tensor) -> float:
    """scripts/test_comparison.py:191-194 (skimage PSNR with data_range 1.0; 100.0 when mse < 1e-10)."""
    mse = float(((a.double() - b.double()) ** 2).mean())
    if mse < 1e-10:
        return 100.0
    import math
    return 10.0 * math.log10(1.0 / mse)
