# Probe of the SURVEY 8(f) device ops at odd / tiny / large shapes against the numpy restatements (oracle/), run through gpurun.
# One line per (op, shape): ok / MISMATCH (count) / FAILED (exception).  The committed tests pin the same ops to the reference's
# own PIL / cv2 / numpy outputs at a few shapes; this sweeps the shapes.
import random
import sys
import numpy as np
import torch
sys.path.insert(0, '.')
from oracle import eval_oracle as EO
from oracle import preproc_oracle as PO
from mri_superresolution_b200.utils.preprocess import preprocess_images, to_tensor, AugmentDecision, PairAugmenter
from mri_superresolution_b200.utils.evaluation import upscale_with_interpolation, calculate_metrics, match_histograms
from mri_superresolution_b200.utils.kspace import area_half, robust_normalize, simulate_low_field_mri

dev = torch.device('cuda')
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
rng = np.random.default_rng(11)
bad = 0


def report(op, shape, fn):
    global bad
    try:
        msg = fn()
        if msg:
            bad += 1
        print(f"{op:22s} {str(shape):14s} {'MISMATCH ' + msg if msg else 'ok'}", flush=True)
    except Exception as ex:
        bad += 1
        print(f"{op:22s} {str(shape):14s} FAILED {type(ex).__name__}: {str(ex)[:160]}", flush=True)


def u8(n, h, w, kind=0):
    if kind == 1:
        return np.full((n, h, w), 77, np.uint8)
    return np.stack([rng.normal(90 + 10 * k, 45, (h, w)).clip(0, 255).astype(np.uint8) for k in range(n)])


SHAPES = [(1, 1), (1, 2), (2, 3), (5, 7), (11, 4), (17, 31), (63, 65), (130, 96), (255, 257), (300, 513), (1024, 768)]
for h, w in SHAPES:
    for kind in (0, 1):
        imgs = u8(2, h, w, kind)
        def f_pp():
            out = preprocess_images(t(imgs))
            for k in range(2):
                ref = PO.preprocess_u8(imgs[k])[0]
                if not np.array_equal(out[k, 0].cpu().numpy(), ref):
                    return f"{int((out[k, 0].cpu().numpy() != ref).sum())} px (slice {k})"
        report("preprocess_images" + ("/const" if kind else ""), (h, w), f_pp)
    imgs = u8(2, h, w)
    report("to_tensor", (h, w), lambda: None if np.array_equal(to_tensor(t(imgs))[:, 0].cpu().numpy(), np.stack([PO.to_tensor_u8(x) for x in imgs])) else "differs")

    def f_aug():
        n = 4
        lr, hr = u8(n, h, w), u8(n, 2 * h, 2 * w)
        r = random.Random(h * 1000 + w)
        decs = [AugmentDecision(flip=bool(k & 1), angle=r.uniform(-5, 5) if k != 2 else None, brightness=r.uniform(0.9, 1.1) if k < 3 else None,
                                contrast=r.uniform(0.9, 1.1) if k > 0 else None, noise=k % 2 == 0) for k in range(n)]
        noise = rng.normal(0, 2.55, lr.shape)
        (l, l8), (f, f8) = PairAugmenter()(t(hr), t(lr), decisions=decs, noise=torch.from_numpy(noise), as_uint8=True)
        for k, dc in enumerate(decs):
            tup = (dc.flip, dc.angle, dc.brightness, dc.contrast, dc.noise)
            a, b = PO.augment_u8(hr[k], tup), PO.augment_u8(lr[k], tup, noise[k])
            if not np.array_equal(f8[k].cpu().numpy(), a):
                return f"hr pair {k}: {int((f8[k].cpu().numpy() != a).sum())} px"
            if not np.array_equal(l8[k].cpu().numpy(), b):
                return f"lr pair {k}: {int((l8[k].cpu().numpy() != b).sum())} px"
    report("augment_pair", (h, w), f_aug)

    for m in ("bilinear", "sharp_bilinear", "bicubic"):
        def f_up(m=m):
            got = upscale_with_interpolation(t(imgs), m).cpu().numpy()
            for k in range(2):
                ref = EO.upscale_with_interpolation(imgs[k], m)
                if not np.array_equal(got[k], ref):
                    return f"{int((got[k] != ref).sum())} px, max {float(np.abs(got[k].astype(np.float64) - ref).max()):.3g}"
        report("upscale/" + m, (h, w), f_up)

    def f_met():
        hr = rng.random((3, 2 * h, 2 * w)).astype(np.float32)
        up = np.clip(hr + 0.05 * rng.standard_normal(hr.shape), 0, 1).astype(np.float32)
        mm = calculate_metrics(t(hr), t(up))
        for k in range(3):
            ref = EO.image_metrics(hr[k], up[k])
            for key in ("mse", "rmse", "mae", "psnr"):
                if abs(float(mm[key][k]) - ref[key]) > 5e-6 * max(abs(ref[key]), 1e-3):
                    return f"{key}[{k}] {float(mm[key][k])} vs {ref[key]}"
    report("calculate_metrics", (2 * h, 2 * w), f_met)

    def f_hm():
        out = (1 / (1 + np.exp(-rng.normal(0, 2, (2, h, w))))).astype(np.float32)
        tgt = rng.random((2, h + 3, max(1, w - 1))).astype(np.float32)
        got = match_histograms(t(out), t(tgt), clip=False).cpu().numpy()
        for k in range(2):
            ref = EO.match_histograms(out[k], tgt[k]).astype(np.float32)
            if not np.array_equal(got[k], ref):
                return f"{int((got[k] != ref).sum())} px, max {float(np.abs(got[k] - ref).max()):.3g}"
    report("match_histograms", (h, w), f_hm)

    def f_rn():
        vol = rng.normal(300, 120, (2, h, w))
        got = robust_normalize(t(vol)).cpu().numpy()
        ref = np.stack([EO.robust_normalize(v) for v in vol])          # per slice, as the device op batches it
        if not np.array_equal(got, ref):
            return f"{int((got != ref).sum())} px, max {float(np.abs(got - ref).max()):.3g}"
    report("robust_normalize", (h, w), f_rn)

    def f_ks():
        x = rng.random((2, h, w))
        nr, ni = rng.normal(0, EO.noise_sigma(h, w), (2, h, w)), rng.normal(0, EO.noise_sigma(h, w), (2, h, w))
        got = simulate_low_field_mri(t(x), noise=(t(nr), t(ni))).cpu().numpy()
        for k in range(2):
            ref = EO.simulate_low_field(x[k], nr[k], ni[k])
            e = float(np.abs(got[k] - ref).max())
            if h * w > 1 and not e < 2e-5:                    # one pixel: 0 / 0 in the reference's min-max
                return f"max err {e:.3g}"
    report("simulate_low_field", (h, w), f_ks)

    if h % 2 == 0 and w % 2 == 0:
        def f_ah():
            x = rng.random((2, h, w))
            got = area_half(t(x)).cpu().numpy()
            for k in range(2):
                ref = EO.area_half(x[k])
                if not np.array_equal(got[k], ref):
                    return f"{int((got[k] != ref).sum())} px"
        report("area_half", (h, w), f_ah)
print("problems:", bad)
