"""CPU restatement of the reference's per-image input pipeline (TEST INFRASTRUCTURE, see oracle/__init__.py).

  preprocess_u8        scripts/infer.py:97-130 (np.percentile + clip + min-max) -- numpy itself, as the reference calls it
  to_tensor_u8         utils/dataset.py:124-125 (torchvision ToTensor)
  draw_decision        utils/dataset.py:142-166: the order in which augment_pair consumes Python's `random` stream
  augment_u8           utils/dataset.py:138-175 with the PIL operations restated in integer / float32 arithmetic
                       (Pillow 12.2: Image.rotate -> Geometry.c affine_fixed, ImageEnhance -> Blend.c)
Pinned by tests/golden/preproc.npz, which oracle/make_golden.py writes by running the reference's own preprocess_image and
MRISuperResDataset.augment_pair (PIL, torchvision, numpy) on seeded inputs.
"""
from __future__ import annotations

import math

import numpy as np

F = np.float32


def preprocess_u8(img: np.ndarray, min_percentile: float = 0.5, max_percentile: float = 99.5):
    """scripts/infer.py:103-116 on one uint8 slice; returns (normalised float32 image, min_val, max_val)."""
    x = img.astype(np.float32)
    lo = np.percentile(x, min_percentile)
    hi = np.percentile(x, max_percentile)
    x = np.clip(x, lo, hi)
    if hi > lo:
        x = (x - lo) / (hi - lo)
    return x.astype(np.float32), F(lo), F(hi)


def to_tensor_u8(img: np.ndarray) -> np.ndarray:
    return (img.astype(np.float32) / F(255.0)).astype(np.float32)


DEFAULTS = {"flip_prob": 0.5, "rotate_prob": 0.5, "rotate_range": (-5, 5), "brightness_prob": 0.3, "brightness_range": (0.9, 1.1),
            "contrast_prob": 0.3, "contrast_range": (0.9, 1.1), "noise_prob": 0.2, "noise_std": 0.01}


def draw_decision(rng, p=DEFAULTS):
    """(flip, angle | None, brightness | None, contrast | None, noise) in the reference's draw order."""
    flip = rng.random() < p["flip_prob"]
    angle = rng.uniform(*p["rotate_range"]) if rng.random() < p["rotate_prob"] else None
    bri = rng.uniform(*p["brightness_range"]) if rng.random() < p["brightness_prob"] else None
    con = rng.uniform(*p["contrast_range"]) if rng.random() < p["contrast_prob"] else None
    noise = rng.random() < p["noise_prob"]
    return flip, angle, bri, con, noise


def _fix(v: float) -> int:
    return int(math.floor(v * 65536.0 + 0.5))


def rotate_nearest(img: np.ndarray, angle: float, fill: int) -> np.ndarray:
    h, w = img.shape
    ang = -math.radians(angle % 360.0)
    m = [round(math.cos(ang), 15), round(math.sin(ang), 15), 0.0, round(-math.sin(ang), 15), round(math.cos(ang), 15), 0.0]
    cx, cy = w / 2.0, h / 2.0
    m[2] = m[0] * -cx + m[1] * -cy + m[2]
    m[5] = m[3] * -cx + m[4] * -cy + m[5]
    m[2] += cx
    m[5] += cy
    a0, a1, a3, a4 = _fix(m[0]), _fix(m[1]), _fix(m[3]), _fix(m[4])
    a2, a5 = _fix(m[2] + m[0] * 0.5 + m[1] * 0.5), _fix(m[5] + m[3] * 0.5 + m[4] * 0.5)
    ys, xs = np.mgrid[0:h, 0:w].astype(np.int64)
    xin, yin = (a2 + a1 * ys + a0 * xs) >> 16, (a5 + a4 * ys + a3 * xs) >> 16
    ok = (xin >= 0) & (xin < w) & (yin >= 0) & (yin < h)
    out = np.full((h, w), fill, np.uint8)
    out[ok] = img[yin[ok], xin[ok]]
    return out


def pil_blend(degenerate: int, img: np.ndarray, factor: float) -> np.ndarray:
    t = F(degenerate) + F(factor) * (img.astype(np.int32) - int(degenerate)).astype(np.float32)
    return np.where(t <= 0, 0, np.where(t >= 255, 255, t.astype(np.int32))).astype(np.uint8)


def augment_u8(img: np.ndarray, decision, noise: np.ndarray | None = None, noise_std: float = 0.01) -> np.ndarray:
    """One image of the pair through augment_pair with the given decision; `noise` (float64, gray levels) only for the LR image."""
    flip, angle, bri, con, do_noise = decision
    x = img
    if flip:
        x = x[:, ::-1]
    if angle is not None:
        fill = int(float(F(x.astype(np.float64).sum() / (255.0 * x.size))) * 255)       # int(to_tensor(img).mean().item() * 255)
        x = rotate_nearest(np.ascontiguousarray(x), angle, fill)
    if bri is not None:
        x = pil_blend(0, x, bri)
    if con is not None:
        mean = int(x.astype(np.float64).sum() / x.size + 0.5)
        x = pil_blend(mean, x, con)
    if do_noise and noise is not None:
        x = np.clip(x.astype(np.float32) + noise, 0, 255).astype(np.uint8)
    return np.ascontiguousarray(x)
