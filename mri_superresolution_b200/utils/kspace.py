"""Low-field LR synthesis on the device: utils/preprocessing.py:126-163,225-293 and the LR half of
utils/extraction_utils.py:139-165 for whole batches of slices, so LR inputs can be made from HR volumes on the fly instead of
from pre-extracted PNG pairs.  cuFFT does the transforms; mask, noise, magnitude and the rescale are fused kernels around them.
No CPU fallback."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from .. import _native

__all__ = ["robust_normalize", "simulate_low_field_mri", "area_half", "make_low_res"]


def _stack(x: torch.Tensor, name: str, dtype) -> torch.Tensor:
    if not x.is_cuda:
        raise RuntimeError(f"{name}: CUDA tensors only (no CPU fallback)")
    if x.dim() == 2:
        x = x[None]
    if x.dim() != 3:
        raise ValueError(f"{name}: expected N x H x W or H x W, got {tuple(x.shape)}")
    return x.to(dtype).contiguous()


def robust_normalize(slice_data: torch.Tensor, lower_percentile: float = 0.5, upper_percentile: float = 99.5,
                     target_range: Tuple[float, float] = (0, 1)) -> torch.Tensor:
    """N x H x W float64 slices (what nibabel's get_fdata yields) -> float32, percentile-clipped and scaled to target_range."""
    src = _stack(slice_data, "slice_data", torch.float64)
    n, h, w = src.shape
    L = _native.lib()
    nbytes = L.mrisr_robust_normalize_scratch_bytes(n, h, w)
    if nbytes == 0:
        raise ValueError("robust_normalize: batch too large for one call (2^31 - 1 pixels)")
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=src.device)
    dst = torch.empty((n, h, w), dtype=torch.float32, device=src.device)
    rc = L.mrisr_robust_normalize_f64(_native.ptr(src), _native.ptr(dst), None, _native.ptr(scratch), nbytes, n, h, w,
                                      float(lower_percentile), float(upper_percentile), float(target_range[0]), float(target_range[1]),
                                      _native.stream_ptr())
    _native.check(rc, "mrisr_robust_normalize_f64")
    return dst


def simulate_low_field_mri(data: torch.Tensor, kspace_crop_factor: float = 0.5, noise_std: float = 5,
                           noise: Optional[Tuple[torch.Tensor, torch.Tensor]] = None, seed: int = 0) -> torch.Tensor:
    """N x H x W float32 slices in [0, 1] -> simulated low-field magnitude images, float64 (numpy's result dtype).  ``noise``:
    optional (real, imag) float64 arrays on the shifted k-space grid (what the reference draws with np.random.normal); default =
    device-side Philox noise keyed by ``seed``."""
    src = _stack(data, "data", torch.float32)
    n, h, w = src.shape
    L = _native.lib()
    nbytes = L.mrisr_kspace_scratch_bytes(n, h, w)
    if nbytes == 0:
        raise RuntimeError("simulate_low_field_mri: cuFFT could not plan this shape: " + L.mrisr_last_error().decode())
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=src.device)
    out = torch.empty((n, h, w), dtype=torch.float64, device=src.device)
    nr = ni = None
    if noise is not None:
        nr, ni = (_stack(t, "noise", torch.float64) for t in noise)
        if nr.shape != src.shape or ni.shape != src.shape:
            raise ValueError("noise arrays must match the data shape")
    rc = L.mrisr_simulate_low_field(_native.ptr(src), _native.ptr(out), _native.ptr(nr), _native.ptr(ni), _native.ptr(scratch), nbytes,
                                    n, h, w, float(kspace_crop_factor), float(noise_std), int(seed) & (2 ** 64 - 1), _native.stream_ptr())
    _native.check(rc, "mrisr_simulate_low_field")
    return out


def area_half(images: torch.Tensor, clip01: bool = False, as_uint8: bool = False) -> torch.Tensor:
    """cv2.resize(INTER_AREA) to exactly half size on float64 slices (``as_uint8``: then clip(x * 255) truncated, the stored PNG)."""
    src = _stack(images, "images", torch.float64)
    n, h, w = src.shape
    dst = torch.empty((n, h // 2, w // 2), dtype=torch.uint8 if as_uint8 else torch.float64, device=src.device)
    rc = _native.lib().mrisr_area_half_f64(_native.ptr(src), None if as_uint8 else _native.ptr(dst), _native.ptr(dst) if as_uint8 else None,
                                           n, h, w, int(clip01), _native.stream_ptr())
    _native.check(rc, "mrisr_area_half_f64")
    return dst


def make_low_res(slice_data: torch.Tensor, kspace_crop_factor: float = 0.5, noise_std: float = 5, noise=None, seed: int = 0) -> torch.Tensor:
    """The LR branch of extract_slices (utils/extraction_utils.py:139-165) for square power-of-two-free HR slices already at the
    target size: robust_normalize -> simulate -> clip -> AREA 1/2 -> uint8."""
    norm = robust_normalize(slice_data)
    sim = simulate_low_field_mri(norm, kspace_crop_factor, noise_std, noise, seed)
    return area_half(sim, clip01=True, as_uint8=True)
