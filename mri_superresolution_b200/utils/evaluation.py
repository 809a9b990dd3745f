"""Batched evaluation on the device: the per-image numpy / cv2 / skimage code of the reference's
comparison scripts as CUDA kernels over whole batches.

  calculate_metrics            scripts/test_comparison.py:164-202   SSIM, MSE, RMSE, MAE, PSNR per image
  upscale_with_interpolation   scripts/test_comparison.py:92-134    cv2 'bilinear' / 'sharp_bilinear' / 'bicubic' x2 on uint8
  match_histograms             scripts/infer.py:285-314             skimage.exposure.match_histograms + clip to [0, 1]
No CPU fallback: tensors must be CUDA tensors.
"""
from __future__ import annotations

from typing import Dict

import torch

from .. import _native
from .losses import ssim

__all__ = ["calculate_metrics", "upscale_with_interpolation", "match_histograms"]

_METHODS = {"bilinear": 0, "sharp_bilinear": 1, "bicubic": 2}


def _images(x: torch.Tensor, name: str, dtype) -> torch.Tensor:
    if not x.is_cuda:
        raise RuntimeError(f"{name}: CUDA tensors only (no CPU fallback)")
    if x.dim() == 4 and x.shape[1] == 1:
        x = x[:, 0]
    if x.dim() == 2:
        x = x[None]
    if x.dim() != 3:
        raise ValueError(f"{name}: expected N x H x W (or N x 1 x H x W, or H x W), got {tuple(x.shape)}")
    if x.dtype != dtype:
        if dtype == torch.uint8:
            raise ValueError(f"{name}: expected uint8, got {x.dtype}")
        x = x.to(dtype)
    return x.contiguous()


def calculate_metrics(hr_image: torch.Tensor, upscaled_image: torch.Tensor) -> Dict[str, torch.Tensor]:
    """Per-image ``{'ssim', 'mse', 'rmse', 'mae', 'psnr'}`` (each a length-N float32 device tensor) of images in [0, 1]."""
    hr, up = _images(hr_image, "hr_image", torch.float32), _images(upscaled_image, "upscaled_image", torch.float32)
    if hr.shape != up.shape:
        raise ValueError(f"shape mismatch {tuple(hr.shape)} vs {tuple(up.shape)}")
    n, h, w = hr.shape
    L = _native.lib()
    out = torch.empty((n, 4), dtype=torch.float32, device=hr.device)
    scratch = torch.empty(L.mrisr_image_metrics_scratch_bytes(n, h, w), dtype=torch.uint8, device=hr.device)
    _native.check(L.mrisr_image_metrics(_native.ptr(hr), _native.ptr(up), _native.ptr(out), _native.ptr(scratch), n, h, w,
                                        _native.stream_ptr()), "mrisr_image_metrics")
    s = ssim(up[:, None], hr[:, None], size_average=False)           # SSIM(window 11, sigma 1.5)(up, hr), :179-180
    return {"ssim": s, "mse": out[:, 0], "rmse": out[:, 1], "mae": out[:, 2], "psnr": out[:, 3]}


def upscale_with_interpolation(lr_images: torch.Tensor, method: str, scale_factor: int = 2, as_uint8: bool = False) -> torch.Tensor:
    """cv2.resize x2 of uint8 slices (N x H x W) -> N x 2H x 2W float32 in [0, 1] (``as_uint8``: the uint8 image instead)."""
    if method not in _METHODS:
        raise ValueError(f"Unknown interpolation method: {method}")                     # :118
    if scale_factor != 2:
        raise ValueError("the B200 path implements the reference's scale_factor = 2")
    src = _images(lr_images, "lr_images", torch.uint8)
    n, h, w = src.shape
    tmp = torch.empty((n, 2 * h, 2 * w), dtype=torch.uint8, device=src.device)
    dst = torch.empty((n, 2 * h, 2 * w), dtype=torch.uint8 if as_uint8 else torch.float32, device=src.device)
    rc = _native.lib().mrisr_upscale2x_u8(_native.ptr(src), None if as_uint8 else _native.ptr(dst), _native.ptr(dst) if as_uint8 else None,
                                          _native.ptr(tmp), n, h, w, _METHODS[method], _native.stream_ptr())
    _native.check(rc, "mrisr_upscale2x_u8")
    return dst


def match_histograms(output: torch.Tensor, reference: torch.Tensor, clip: bool = True) -> torch.Tensor:
    """Histogram matching of model outputs to (normalised) target images, pairwise over the batch; clipped to [0, 1] as
    scripts/infer.py:309 does."""
    src, ref = _images(output, "output", torch.float32), _images(reference, "reference", torch.float32)
    if src.shape[0] != ref.shape[0]:
        raise ValueError("one reference image per output image")
    n, h, w = src.shape
    hr, wr = ref.shape[1:]
    L = _native.lib()
    nbytes = L.mrisr_match_histograms_scratch_bytes(n, h, w, hr, wr)
    if nbytes == 0:
        raise ValueError("match_histograms: batch too large for one call (2^31 - 1 pixels)")
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=src.device)
    out = torch.empty_like(src)
    rc = L.mrisr_match_histograms(_native.ptr(src), _native.ptr(ref), _native.ptr(out), _native.ptr(scratch), nbytes, n, h, w, hr, wr,
                                  _native.stream_ptr())
    _native.check(rc, "mrisr_match_histograms")
    return out.clamp_(0.0, 1.0) if clip else out
