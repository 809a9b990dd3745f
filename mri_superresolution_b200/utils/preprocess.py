"""On-device input pipeline: the per-image host work of the reference's loaders as batched CUDA kernels.

  preprocess_images   scripts/infer.py:97-130   percentile clip (0.5 / 99.5) + min-max of uint8 slices -> N x 1 x H x W float32
  to_tensor           utils/dataset.py:124-125  transforms.ToTensor(): uint8 / 255
  PairAugmenter       utils/dataset.py:68-175   augment_pair: same defaults, same draw order from Python's ``random``, so the
                                                 reference's seed gives the reference's decisions; the pixels are transformed
                                                 on the device (mrisr_augment_u8), bit for bit what PIL produces

At 3.9 k train / 13 k inference slices per second and GPU the reference's PIL loader (one image at a time, <= 16 worker
processes, scripts/train.py:133-140,215-233) cannot feed the path; uint8 batches are 1/4 of the float32 bytes over PCIe, too.
No CPU fallback: tensors must be CUDA tensors.
"""
from __future__ import annotations

import math
import random
import struct
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from .. import _native

__all__ = ["preprocess_images", "to_tensor", "pil_rotate_matrix", "AugmentDecision", "PairAugmenter", "DEFAULT_AUG_PARAMS"]

AUG_INTS = 16        # MRISR_AUG_PARAM_INTS


def _u8(x: torch.Tensor, name: str) -> torch.Tensor:
    if not x.is_cuda:
        raise RuntimeError(f"{name}: the B200 input pipeline needs CUDA tensors (no CPU fallback)")
    if x.dtype != torch.uint8:
        raise ValueError(f"{name}: expected uint8 gray levels, got {x.dtype}")
    if x.dim() == 4 and x.shape[1] == 1:
        x = x[:, 0]
    if x.dim() == 2:
        x = x[None]
    if x.dim() != 3:
        raise ValueError(f"{name}: expected N x H x W (or N x 1 x H x W, or H x W), got {tuple(x.shape)}")
    return x.contiguous()


def to_tensor(images: torch.Tensor) -> torch.Tensor:
    """transforms.ToTensor() for a batch of uint8 slices: N x 1 x H x W float32 in [0, 1]."""
    src = _u8(images, "to_tensor")
    n, h, w = src.shape
    dst = torch.empty((n, 1, h, w), dtype=torch.float32, device=src.device)
    _native.check(_native.lib().mrisr_u8_to_float(_native.ptr(src), _native.ptr(dst), src.numel(), _native.stream_ptr()),
                  "mrisr_u8_to_float")
    return dst


def preprocess_images(images: torch.Tensor, min_percentile: float = 0.5, max_percentile: float = 99.5,
                      return_window: bool = False):
    """preprocess_image of scripts/infer.py:97-130 for a batch of uint8 slices (what ``Image.open(p).convert('L')`` yields):
    N x 1 x H x W float32.  ``return_window``: also the N x 2 tensor of (min_val, max_val) per slice."""
    src = _u8(images, "preprocess_images")
    n, h, w = src.shape
    dst = torch.empty((n, 1, h, w), dtype=torch.float32, device=src.device)
    mm = torch.empty((n, 2), dtype=torch.float32, device=src.device)
    rc = _native.lib().mrisr_percentile_normalize_u8(_native.ptr(src), _native.ptr(dst), _native.ptr(mm), n, h, w,
                                                     float(min_percentile), float(max_percentile), _native.stream_ptr())
    _native.check(rc, "mrisr_percentile_normalize_u8")
    return (dst, mm) if return_window else dst


def pil_rotate_matrix(angle: float, w: int, h: int) -> Tuple[int, int, int, int, int, int]:
    """Image.rotate(angle) (Pillow Image.py: rotate -> transform(AFFINE), Geometry.c: affine_fixed) as the six 16.16 fixed-point
    coefficients the nearest-neighbour loop steps with: xin = (a2 + a1*y + a0*x) >> 16, yin = (a5 + a4*y + a3*x) >> 16."""
    ang = -math.radians(angle % 360.0)                     # Image.rotate: angle = angle % 360.0 first
    m = [round(math.cos(ang), 15), round(math.sin(ang), 15), 0.0, round(-math.sin(ang), 15), round(math.cos(ang), 15), 0.0]
    cx, cy = w / 2.0, h / 2.0
    m[2] = m[0] * -cx + m[1] * -cy + m[2]
    m[5] = m[3] * -cx + m[4] * -cy + m[5]
    m[2] += cx
    m[5] += cy
    fix = lambda v: int(math.floor(v * 65536.0 + 0.5))
    return (fix(m[0]), fix(m[1]), fix(m[2] + m[0] * 0.5 + m[1] * 0.5), fix(m[3]), fix(m[4]), fix(m[5] + m[3] * 0.5 + m[4] * 0.5))


def _fbits(v: float) -> int:
    return struct.unpack("<i", struct.pack("<f", float(v)))[0]


def _i32(v: int) -> int:
    v &= 0xFFFFFFFF
    return v - (1 << 32) if v >= (1 << 31) else v


class AugmentDecision:
    """The random choices augment_pair makes for one pair (utils/dataset.py:142-171), drawn on the host."""
    __slots__ = ("flip", "angle", "brightness", "contrast", "noise", "seed")

    def __init__(self, flip=False, angle=None, brightness=None, contrast=None, noise=False, seed=0):
        self.flip, self.angle, self.brightness, self.contrast, self.noise, self.seed = flip, angle, brightness, contrast, noise, seed

    def row(self, w: int, h: int, noise_std_levels: float, with_noise: bool) -> List[int]:
        a = pil_rotate_matrix(self.angle, w, h) if self.angle is not None else (0,) * 6
        return [int(self.flip), int(self.angle is not None), *[_i32(v) for v in a],
                int(self.brightness is not None), _fbits(self.brightness or 0.0),
                int(self.contrast is not None), _fbits(self.contrast or 0.0),
                int(bool(self.noise and with_noise)), _fbits(noise_std_levels), _i32(self.seed), _i32(self.seed >> 32)]


DEFAULT_AUG_PARAMS: Dict = {"flip_prob": 0.5, "rotate_prob": 0.5, "rotate_range": (-5, 5), "brightness_prob": 0.3,
                            "brightness_range": (0.9, 1.1), "contrast_prob": 0.3, "contrast_range": (0.9, 1.1),
                            "noise_prob": 0.2, "noise_std": 0.01}            # utils/dataset.py:71-81


class PairAugmenter:
    """augment_pair + ToTensor for a whole batch: ``(full_u8 N x H2 x W2, low_u8 N x H x W) -> (low, full)`` float32
    N x 1 x . x . tensors, the order ``MRISuperResDataset.__getitem__`` returns (utils/dataset.py:136)."""

    def __init__(self, augmentation_params: Optional[Dict] = None, rng: Optional[random.Random] = None):
        self.aug_params = {**DEFAULT_AUG_PARAMS, **(augmentation_params or {})}
        self.rng = rng if rng is not None else random           # the reference draws from the module-level generator
        # keys of the device-side noise generator come from their own stream: the reference draws its noise from numpy's
        # generator (utils/dataset.py:168), never from `random`, so taking keys from self.rng would shift every later decision
        self._keys = random.Random(0x6D726973)

    def draw(self) -> AugmentDecision:
        """One pair's decisions in the reference's draw order (flip, rotate + angle, brightness + factor, contrast + factor, noise)."""
        p, r = self.aug_params, self.rng
        d = AugmentDecision()
        d.flip = r.random() < p["flip_prob"]
        if r.random() < p["rotate_prob"]:
            d.angle = r.uniform(*p["rotate_range"])
        if r.random() < p["brightness_prob"]:
            d.brightness = r.uniform(*p["brightness_range"])
        if r.random() < p["contrast_prob"]:
            d.contrast = r.uniform(*p["contrast_range"])
        d.noise = r.random() < p["noise_prob"]
        d.seed = self._keys.getrandbits(64) if d.noise else 0
        return d

    def _apply(self, src: torch.Tensor, decisions: Sequence[AugmentDecision], with_noise: bool, noise: Optional[torch.Tensor],
               as_uint8: bool):
        n, h, w = src.shape
        rows = [d.row(w, h, self.aug_params["noise_std"] * 255.0, with_noise) for d in decisions]
        params = torch.tensor(rows, dtype=torch.int32).to(src.device, non_blocking=True)
        tmp = torch.empty_like(src)
        sums = torch.empty(2 * n, dtype=torch.int64, device=src.device)
        dst = torch.empty((n, 1, h, w), dtype=torch.float32, device=src.device)
        dst8 = torch.empty_like(src) if as_uint8 else None
        if noise is not None:
            if noise.dtype != torch.float64 or tuple(noise.shape[-2:]) != (h, w) or noise.numel() != src.numel():
                raise ValueError("noise must be a float64 tensor with one value per low-resolution pixel")
            noise = noise.to(src.device).contiguous()
        rc = _native.lib().mrisr_augment_u8(_native.ptr(src), _native.ptr(dst), _native.ptr(dst8), _native.ptr(tmp),
                                            _native.ptr(sums), _native.ptr(params), _native.ptr(noise), n, h, w,
                                            _native.stream_ptr())
        _native.check(rc, "mrisr_augment_u8")
        return (dst, dst8) if as_uint8 else dst

    def __call__(self, full_u8: torch.Tensor, low_u8: torch.Tensor, decisions: Optional[Sequence[AugmentDecision]] = None,
                 noise: Optional[torch.Tensor] = None, as_uint8: bool = False):
        full, low = _u8(full_u8, "full_u8"), _u8(low_u8, "low_u8")
        if full.shape[0] != low.shape[0]:
            raise ValueError("full / low batches differ in length")
        if decisions is None:
            decisions = [self.draw() for _ in range(full.shape[0])]
        if len(decisions) != full.shape[0]:
            raise ValueError("one AugmentDecision per pair")
        f = self._apply(full, decisions, False, None, as_uint8)
        l = self._apply(low, decisions, True, noise, as_uint8)
        return l, f
