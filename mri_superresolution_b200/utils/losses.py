"""Drop-in ``ssim`` / ``SSIM`` / ``CombinedLoss`` backed by the fused sm_100a loss kernel.

Mirrors the public surface of the reference's utils/losses.py (function ``ssim`` :27, ``CombinedLoss``
:153-240, ``SSIM`` :242-266): same constructor keywords, attributes, argument validation
(ValueError, :166-171) and return types (0-dim tensors supporting ``.backward()`` / ``.item()``).
One kernel (``mrisr_l1_ssim_fwd_bwd``) produces the L1 term, the SSIM term and d(loss)/d(output) in a
single pass; the per-term values stay on the device (``last_components``) instead of the reference's
four ``.item()`` host syncs per call (:208,:225,:226,:236).  No CPU fallback.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import _native

__all__ = ["gaussian_window", "create_window", "ssim", "SSIM", "CombinedLoss"]


def gaussian_window(window_size: int, sigma: float) -> torch.Tensor:
    """1-D normalised Gaussian (utils/losses.py:10-18)."""
    x = torch.arange(window_size, dtype=torch.float32) - (window_size // 2)
    g = torch.exp(-(x * x) / (2 * sigma ** 2))
    return g / g.sum()


def create_window(window_size: int, channel: int, sigma: float, device) -> torch.Tensor:
    """(channel, 1, k, k) outer-product window (utils/losses.py:20-25); kept as a buffer for state parity."""
    g = gaussian_window(window_size, sigma).to(device)
    return torch.outer(g, g).expand(channel, 1, window_size, window_size).contiguous()


def _as_images(t: torch.Tensor, name: str) -> torch.Tensor:
    if t.dim() != 4:
        raise ValueError(f"{name}: expected N x C x H x W, got {tuple(t.shape)}")
    if not t.is_cuda:
        raise RuntimeError(f"{name}: the B200 loss kernels need CUDA tensors (no CPU fallback)")
    t = t.detach().to(torch.float32).contiguous()
    n, c, h, w = t.shape
    return t.view(n * c, 1, h, w)      # depthwise filters: channels behave like extra samples


def _ssim_per_image(img1: torch.Tensor, img2: torch.Tensor, window_size: int, sigma: float, val_range: float):
    a, b = _as_images(img1, "img1"), _as_images(img2, "img2")
    if a.shape != b.shape:
        raise ValueError(f"ssim: shape mismatch {tuple(img1.shape)} vs {tuple(img2.shape)}")
    n, _, h, w = a.shape
    L = _native.lib()
    scratch = torch.empty(L.mrisr_l1_ssim_scratch_bytes(n, h, w), dtype=torch.uint8, device=a.device)
    out = torch.empty(n, dtype=torch.float32, device=a.device)
    rc = L.mrisr_ssim_forward(_native.ptr(a), _native.ptr(b), _native.ptr(out), _native.ptr(scratch), n, h, w,
                              int(window_size), float(sigma), float(val_range), _native.stream_ptr())
    _native.check(rc, "mrisr_ssim_forward")
    return out


def ssim(img1, img2, window_size=11, sigma=1.5, val_range=1.0, device=None, window=None, size_average=True):
    """SSIM index, forward only (metric use; utils/losses.py:27-81).  ``window`` is accepted for signature
    parity; the kernel derives its taps from (window_size, sigma) exactly as ``create_window`` does."""
    n, c = img1.shape[0], img1.shape[1]
    per = _ssim_per_image(img1, img2, window_size, sigma, val_range).view(n, c)
    res = per.mean() if size_average else per.mean(1)
    if img1.dtype != torch.float32 and img1.dtype == img2.dtype:      # :78-79
        res = res.to(img1.dtype)
    return res


class SSIM(nn.Module):
    """Metric wrapper (utils/losses.py:242-266)."""

    def __init__(self, window_size=11, sigma=1.5, val_range=1.0, device=None):
        super().__init__()
        self.window_size = window_size
        self.sigma = sigma
        self.val_range = val_range
        self.device = device if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.register_buffer("window", create_window(window_size, 1, sigma, self.device))

    def forward(self, img1, img2):
        return ssim(img1, img2, self.window_size, self.sigma, self.val_range, img1.device, None)


class _L1SSIMFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, output, target, l1_w, ssim_w, window_size, sigma, val_range, holder):
        pred = _as_images(output, "output")
        tgt = _as_images(target, "target")
        if pred.shape != tgt.shape:
            raise ValueError(f"CombinedLoss: shape mismatch {tuple(output.shape)} vs {tuple(target.shape)}")
        n, _, h, w = pred.shape
        L = _native.lib()
        need_grad = output.requires_grad
        dpred = torch.empty_like(pred) if need_grad else None
        out5 = torch.empty(5, dtype=torch.float32, device=pred.device)
        scratch = torch.empty(L.mrisr_l1_ssim_scratch_bytes(n, h, w), dtype=torch.uint8, device=pred.device)
        rc = L.mrisr_l1_ssim_fwd_bwd(_native.ptr(pred), _native.ptr(tgt), _native.ptr(dpred), _native.ptr(out5),
                                     _native.ptr(scratch), n, h, w, float(l1_w), float(ssim_w), int(window_size),
                                     float(sigma), float(val_range), _native.stream_ptr())
        _native.check(rc, "mrisr_l1_ssim_fwd_bwd")
        holder["components"] = out5
        ctx.shape = tuple(output.shape)
        ctx.dtype = output.dtype
        if need_grad:
            ctx.save_for_backward(dpred)
        return out5[2].clone()

    @staticmethod
    def backward(ctx, grad_out):
        (dpred,) = ctx.saved_tensors
        g = (dpred * grad_out).view(ctx.shape).to(ctx.dtype)
        return g, None, None, None, None, None, None, None


class CombinedLoss(nn.Module):
    """total = l1_w * L1 + ssim_w * (1 - clamp(SSIM, 0, 1)) [+ perceptual_w * VGG19 feature loss]
    with l1_w = 1 - ssim_w - perceptual_w (utils/losses.py:153-240)."""

    def __init__(self, ssim_weight=0.5, perceptual_weight=0.0, vgg_layer_idx=35, perceptual_loss_type="l1",
                 window_size=11, sigma=1.5, val_range=1.0, device=torch.device("cpu"), vgg_weights=None, precision=None):
        """``vgg_weights`` (extra, optional): (weight, bias) pairs for the VGG19 convolutions instead of torchvision's
        ImageNet download (utils/losses.py:90) -- for offline machines and tests.  ``precision`` (extra): 'fp32' runs the
        perceptual term in its CUDA-core parity mode (the L1 + SSIM kernel is fp32 in either mode)."""
        super().__init__()
        if not (0 <= ssim_weight <= 1):
            raise ValueError("ssim_weight must be between 0 and 1")
        if not (0 <= perceptual_weight <= 1):
            raise ValueError("perceptual_weight must be between 0 and 1")
        if ssim_weight + perceptual_weight > 1:
            raise ValueError("Sum of ssim_weight and perceptual_weight cannot exceed 1")
        self.ssim_weight = ssim_weight
        self.perceptual_weight = perceptual_weight
        self.l1_weight = 1.0 - ssim_weight - perceptual_weight
        self.l1_loss = nn.L1Loss()              # attribute kept for interface parity; not on the compute path
        self.window_size = window_size
        self.sigma = sigma
        self.val_range = val_range
        self.device = device
        self.register_buffer("window", create_window(window_size, 1, sigma, device))
        self.use_perceptual = perceptual_weight > 0
        if self.use_perceptual:
            from .perceptual import PerceptualLoss
            self.perceptual_loss = PerceptualLoss(feature_layer_idx=vgg_layer_idx, loss_type=perceptual_loss_type,
                                                  vgg_weights=vgg_weights, precision=precision)
        else:
            self.perceptual_loss = None
        self._holder = {}

    @property
    def last_components(self):
        """Device tensor [l1, ssim (unclamped), l1+ssim total, perceptual, total] of the last call (no sync)."""
        return self._holder.get("components")

    def forward(self, output, target):
        total = _L1SSIMFunction.apply(output, target, max(self.l1_weight, 0.0), self.ssim_weight, self.window_size,
                                      self.sigma, self.val_range, self._holder)
        if self.use_perceptual:
            total = total + self.perceptual_weight * self.perceptual_loss(output, target)
        return total
