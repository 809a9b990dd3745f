"""CPU restatement of the reference's evaluation and LR-synthesis code (TEST INFRASTRUCTURE, see oracle/__init__.py).

  resize_linear_x2 / resize_cubic_x2 / sharpen3x3   scripts/test_comparison.py:92-134 calls cv2.resize / cv2.filter2D; OpenCV
        (opencv-python, unpinned in requirements.txt; 4.13.0 here) is not under /root/reference, so its published fixed-point
        algorithm (modules/imgproc/src/resize.cpp: HResize* / VResize*, INTER_RESIZE_COEF_BITS = 11) is restated and pinned
        against cv2's own outputs (tests/golden/eval.npz)
  image_metrics        scripts/test_comparison.py:164-202 (PSNR = skimage.metrics.peak_signal_noise_ratio, data_range 1:
        10 log10(1 / mse); scikit-image is absent here, its one-line definition is restated)
  match_histograms     scripts/infer.py:303-307 calls skimage.exposure.match_histograms (absent here): the published algorithm
        (skimage/exposure/histogram_matching.py: _match_cumulative_cdf) restated with numpy
  robust_normalize, simulate_low_field, area_half   utils/preprocessing.py:126-163,225-293, extraction_utils.py:143-157: numpy
        as the reference calls it (the noise arrays are arguments so that a test can feed both sides the same draw)
"""
from __future__ import annotations

import numpy as np

F = np.float32


def _coords(dn):
    d = np.arange(dn)
    f = (d + 0.5) * 0.5 - 0.5
    s = np.floor(f).astype(np.int64)
    return s, (f - s).astype(np.float32)


def resize_linear_x2(img: np.ndarray) -> np.ndarray:
    h, w = img.shape
    sx, fx = _coords(2 * w)
    sy, fy = _coords(2 * h)
    sx, fx = sx.copy(), fx.copy()
    lo = sx < 0
    fx[lo] = 0
    sx[lo] = 0
    hi = sx >= w - 1
    fx[hi] = 0
    sx[hi] = w - 1
    sc = lambda v: np.rint(v * 2048).astype(np.int32)
    ax0, ax1, ay0, ay1 = sc(1 - fx), sc(fx), sc(1 - fy), sc(fy)
    s = img.astype(np.int32)
    rows = s[:, sx] * ax0[None, :] + s[:, np.minimum(sx + 1, w - 1)] * ax1[None, :]
    S0, S1 = rows[np.clip(sy, 0, h - 1)], rows[np.clip(sy + 1, 0, h - 1)]
    out = (((ay0[:, None] * (S0 >> 4)) >> 16) + ((ay1[:, None] * (S1 >> 4)) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


def _cubic_taps(x):
    a = -0.75
    x = F(x)
    c = np.zeros(4, np.float32)
    c[0] = ((a * (x + 1) - 5 * a) * (x + 1) + 8 * a) * (x + 1) - 4 * a
    c[1] = ((a + 2) * x - (a + 3)) * x * x + 1
    c[2] = ((a + 2) * (1 - x) - (a + 3)) * (1 - x) * (1 - x) + 1
    c[3] = F(1) - c[0] - c[1] - c[2]
    return np.rint(c * 2048).astype(np.int64)


def resize_cubic_x2(img: np.ndarray) -> np.ndarray:
    h, w = img.shape
    H, W = 2 * h, 2 * w
    sx, fx = _coords(W)
    sy, fy = _coords(H)
    s = img.astype(np.int64)
    rows = np.zeros((h, W), np.int64)
    for dx in range(W):
        c = _cubic_taps(fx[dx])
        for k in range(4):
            rows[:, dx] += s[:, min(max(sx[dx] - 1 + k, 0), w - 1)] * c[k]
    out = np.zeros((H, W), np.int64)
    nv = (W // 4) * 4                          # SIMD body of VResizeCubic: groups of four columns in float32
    scale = F(1.0) / F(2048 * 2048)
    for dy in range(H):
        c = _cubic_taps(fy[dy])
        S = [rows[min(max(sy[dy] - 1 + k, 0), h - 1)] for k in range(4)]
        out[dy] = (sum(S[k] * c[k] for k in range(4)) + (1 << 21)) >> 22
        b = [F(F(c[k]) * scale) for k in range(4)]
        f = [S[k][:nv].astype(np.float32) for k in range(4)]
        r = f[3] * b[3]
        for k in (2, 1, 0):
            r = (f[k] * b[k]).astype(np.float32) + r
        out[dy, :nv] = np.rint(r).astype(np.int64)
    return np.clip(out, 0, 255).astype(np.uint8)


def sharpen3x3(img: np.ndarray) -> np.ndarray:
    h, w = img.shape
    p = np.pad(img.astype(np.int32), 1, mode="reflect")
    acc = 10 * p[1:-1, 1:-1]
    for dy in range(3):
        for dx in range(3):
            acc = acc - p[dy:dy + h, dx:dx + w]
    return np.clip(acc, 0, 255).astype(np.uint8)


def upscale_with_interpolation(img: np.ndarray, method: str) -> np.ndarray:
    if method == "bicubic":
        up = resize_cubic_x2(img)
    elif method in ("bilinear", "sharp_bilinear"):
        up = resize_linear_x2(img)
        if method == "sharp_bilinear":
            up = sharpen3x3(up)
    else:
        raise ValueError(f"Unknown interpolation method: {method}")
    return up.astype(np.float32) / 255.0


def image_metrics(hr: np.ndarray, up: np.ndarray) -> dict:
    mse = ((hr - up) ** 2).mean()
    rmse = np.sqrt(mse)
    mae = np.abs(hr - up).mean()
    psnr = 100.0 if mse < 1e-10 else 10.0 * np.log10(1.0 / float(mse))
    return {"mse": float(mse), "rmse": float(rmse), "mae": float(mae), "psnr": float(psnr)}


def match_histograms(source: np.ndarray, template: np.ndarray) -> np.ndarray:
    src_values, src_idx, src_counts = np.unique(source.reshape(-1), return_inverse=True, return_counts=True)
    tmpl_values, tmpl_counts = np.unique(template.reshape(-1), return_counts=True)
    src_q = np.cumsum(src_counts) / source.size
    tmpl_q = np.cumsum(tmpl_counts) / template.size
    return np.interp(src_q, tmpl_q, tmpl_values)[src_idx].reshape(source.shape)


def robust_normalize(x: np.ndarray, lower=0.5, upper=99.5, target=(0, 1)) -> np.ndarray:
    if x.size == 0 or np.all(x == x.flat[0]):
        return np.zeros_like(x, dtype=np.float32)
    lo, hi = np.percentile(x, lower), np.percentile(x, upper)
    if hi == lo:
        return np.zeros_like(x, dtype=np.float32)
    n = (np.clip(x, lo, hi) - lo) / (hi - lo)
    return (n * (target[1] - target[0]) + target[0]).astype(np.float32)


def simulate_low_field(data: np.ndarray, noise_real: np.ndarray, noise_imag: np.ndarray, kspace_crop_factor=0.5) -> np.ndarray:
    """preprocessing.py:225-293 with the two np.random.normal draws (:275-276) passed in."""
    omin, omax = data.min(), data.max()
    k = np.fft.fftshift(np.fft.fft2(data))
    rows, cols = k.shape
    cr, cc = rows // 2, cols // 2
    sr, sc = int(rows * kspace_crop_factor), int(cols * kspace_crop_factor)
    mask = np.zeros((rows, cols), dtype=np.complex128)
    mask[cr - sr // 2:cr + sr // 2, cc - sc // 2:cc + sc // 2] = 1
    noisy = k * mask + noise_real + 1j * noise_imag
    mag = np.abs(np.fft.ifft2(np.fft.ifftshift(noisy)))
    sim = (mag - mag.min()) / (mag.max() - mag.min())
    return sim * (omax - omin) + omin


def noise_sigma(rows: int, cols: int, noise_std=5) -> float:
    return (noise_std / 255.0) * np.sqrt(rows * cols) / 10          # :271


def area_half(x: np.ndarray) -> np.ndarray:
    """cv2.resize(float64, half size, INTER_AREA): scalar ResizeAreaFast, ((s00 + s01) + s10 + s11) * 0.25 in double."""
    h, w = x.shape
    s = x.reshape(h // 2, 2, w // 2, 2)
    return (((s[:, 0, :, 0] + s[:, 0, :, 1]) + s[:, 1, :, 0]) + s[:, 1, :, 1]) * 0.25
