"""Generate tests/golden/*.npz by RUNNING THE REFERENCE ITSELF (imported read-only from
/root/reference) -- TEST INFRASTRUCTURE, see oracle/__init__.py.

Run in the build container only (the reference does not exist on the GPU box):

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden

The reference ships no golden vectors, so these files are what pins the
oracle (tests/test_oracle_golden.py) and, through it, the CUDA path.  Inputs and weights are
re-derived from seeds by the oracle's own generators (platform-stable CPU torch.Generator
streams); every file carries checksums of them so a mismatch is detected instead of silently
comparing different problems.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

REF = os.environ.get("MRISR_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")

BIG = 4096  # tensors above this many elements are stored as (head, norm, sum) instead of in full


def _import_reference():
    sys.dont_write_bytecode = True
    sys.path.insert(0, REF)
    from models.unet_model import UNetSuperRes          # noqa
    import utils.losses as ref_losses                   # noqa
    sys.path.pop(0)
    return UNetSuperRes, ref_losses


def _summ(t: torch.Tensor) -> dict:
    t = t.detach().to(torch.float64).flatten()
    return {"norm": float(t.norm()), "sum": float(t.sum()), "head": t[:512].to(torch.float32).numpy()}


def _store_tensor(d: dict, key: str, t: torch.Tensor) -> None:
    t = t.detach()
    if t.numel() <= BIG:
        d[key] = t.to(torch.float32).numpy()
    else:
        s = _summ(t)
        d[key + "@head"] = s["head"]
        d[key + "@norm"] = np.float64(s["norm"])
        d[key + "@sum"] = np.float64(s["sum"])


def _checksum_params(params) -> np.ndarray:
    return np.array([float(sum(v.double().sum() for v in params.values())),
                     float(sum((v.double() ** 2).sum() for v in params.values()))])


def unet_case(UNetSuperRes, ref_losses, name, f, n, h, w, affine, ssim_weight, full_out=True):
    from oracle.unet_oracle import init_params, randomize_affine
    from oracle.synth import make_pairs
    params = init_params(f, seed=1234)
    if affine:
        params = randomize_affine(params)
    x, target = make_pairs(n, h, w, seed=1234)
    model = UNetSuperRes(1, 1, base_filters=f)
    model.load_state_dict(params, strict=True)          # the reference's own strict key/shape check
    model.train()
    crit = ref_losses.CombinedLoss(ssim_weight=ssim_weight, perceptual_weight=0.0)
    metric = ref_losses.SSIM(device=torch.device("cpu"))
    out = model(x)
    out.retain_grad()
    loss = crit(out, target)
    loss.backward()
    d = {"meta": np.array([f, n, h, w, int(affine)]), "ssim_weight": np.float64(ssim_weight),
         "param_checksum": _checksum_params(params),
         "x_checksum": np.array([float(x.double().sum()), float(target.double().sum())]),
         "loss": np.float64(loss.item()),
         "l1": np.float64((out - target).abs().mean().item()),
         "ssim": np.float64(metric(out, target).item())}
    if full_out:
        d["out"] = out.detach().numpy()
        d["dout"] = out.grad.numpy()
    else:
        _store_tensor(d, "out", out)
        _store_tensor(d, "dout", out.grad)
    for k, v in model.named_parameters():
        _store_tensor(d, "grad:" + k, v.grad)
    # intermediate activations after each encoder/decoder block (hooked from the reference modules)
    feats = {}
    hooks = []
    for mod_name in ("inc", "down1", "down2", "down3", "up1", "up2", "up3"):
        hooks.append(getattr(model, mod_name).register_forward_hook(
            lambda m, i, o, k=mod_name: feats.__setitem__(k, o.detach())))
    with torch.no_grad():
        model(x)
    for hk in hooks:
        hk.remove()
    for k, v in feats.items():
        _store_tensor(d, "act:" + k, v)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
    print(name, "loss", d["loss"], "ssim", d["ssim"])


def loss_cases(ref_losses):
    """SSIM / CombinedLoss goldens on assorted shapes, incl. gradient wrt output."""
    g = torch.Generator("cpu").manual_seed(77)
    d = {}
    shapes = [(1, 1, 16, 16), (2, 1, 40, 56), (3, 1, 64, 64), (1, 1, 11, 23), (2, 1, 96, 72)]
    for i, shp in enumerate(shapes):
        base = torch.rand(shp, generator=g)
        base = torch.nn.functional.avg_pool2d(base, 5, 1, 2)
        base = (base - base.amin()) / (base.amax() - base.amin())
        pred = (base + 0.08 * torch.randn(shp, generator=g)).clamp(0, 1)
        d[f"s{i}:target"] = base.numpy()
        d[f"s{i}:pred"] = pred.numpy()
        d[f"s{i}:ssim"] = np.float64(ref_losses.ssim(pred, base).item())
        d[f"s{i}:ssim_per_sample"] = ref_losses.ssim(pred, base, size_average=False).numpy()
        for j, (sw,) in enumerate([(0.4,), (0.0,), (1.0,), (0.3,)]):
            p = pred.clone().requires_grad_(True)
            crit = ref_losses.CombinedLoss(ssim_weight=sw)
            loss = crit(p, base)
            loss.backward()
            d[f"s{i}:w{j}:sw"] = np.float64(sw)
            d[f"s{i}:w{j}:loss"] = np.float64(loss.item())
            d[f"s{i}:w{j}:grad"] = p.grad.numpy()
    # identical images -> ssim == 1, clamp edge
    x = torch.rand((1, 1, 32, 32), generator=g)
    d["ident:ssim"] = np.float64(ref_losses.ssim(x, x).item())
    # window taps as printed by the reference
    d["taps"] = ref_losses.gaussian_window(11, 1.5).numpy()
    d["window2d"] = ref_losses.create_window(11, 1, 1.5, torch.device("cpu")).numpy()
    np.savez_compressed(os.path.join(OUT, "losses.npz"), **d)
    print("losses: ", {k: float(v) for k, v in d.items() if k.endswith(":ssim")})


def perceptual_case(ref_losses):
    """CombinedLoss with the VGG19[:36] branch; ImageNet weights are not on disk / no network, so
    torchvision.models.vgg19 is patched to weights=None and the oracle's seeded tensors are
    loaded into it."""
    import torchvision.models as tvm
    from oracle.losses_oracle import make_vgg_params
    from oracle.synth import make_pairs
    vggp = make_vgg_params(seed=1235)
    orig = tvm.vgg19

    def patched(weights=None, **kw):
        m = orig(weights=None, **kw)
        convs = [l for l in m.features if isinstance(l, torch.nn.Conv2d)]
        with torch.no_grad():
            for l, (wt, b) in zip(convs, vggp):
                l.weight.copy_(wt)
                l.bias.copy_(b)
        return m

    tvm.vgg19 = patched
    try:
        d = {}
        _, hr = make_pairs(2, 32, 32, seed=99)             # HR 64x64
        g = torch.Generator("cpu").manual_seed(5)
        pred0 = (hr + 0.1 * torch.randn(hr.shape, generator=g)).clamp(0, 1)
        d["pred"] = pred0.numpy()
        d["target"] = hr.numpy()
        for tag, (sw, pw, typ) in {"a": (0.4, 0.1, "l1"), "b": (0.0, 1.0, "l2"), "c": (0.3, 0.2, "mse")}.items():
            crit = ref_losses.CombinedLoss(ssim_weight=sw, perceptual_weight=pw, perceptual_loss_type=typ)
            p = pred0.clone().requires_grad_(True)
            loss = crit(p, hr)
            loss.backward()
            with torch.no_grad():
                feats = crit.perceptual_loss.feature_extractor(pred0)
                perc = crit.perceptual_loss(pred0, hr)
            d[f"{tag}:cfg"] = np.array([sw, pw, {"l1": 1, "l2": 2, "mse": 2}[typ]])
            d[f"{tag}:loss"] = np.float64(loss.item())
            d[f"{tag}:perc"] = np.float64(perc.item())
            d[f"{tag}:grad"] = p.grad.numpy()
            if tag == "a":
                _store_tensor(d, "feat", feats)
        np.savez_compressed(os.path.join(OUT, "perceptual.npz"), **d)
        print("perceptual:", {k: float(v) for k, v in d.items() if k.endswith("loss") or k.endswith("perc")})
    finally:
        tvm.vgg19 = orig


def train_case(UNetSuperRes, ref_losses):
    """Three reference train steps (scripts/train.py:296-323 body) at f=16."""
    from oracle.unet_oracle import init_params, randomize_affine
    from oracle.synth import make_pairs
    params = randomize_affine(init_params(16, seed=1234))
    model = UNetSuperRes(1, 1, base_filters=16)
    model.load_state_dict(params, strict=True)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=1e-5)   # lr raised so 3 steps move visibly
    crit = ref_losses.CombinedLoss(ssim_weight=0.4)
    metric = ref_losses.SSIM(device=torch.device("cpu"))
    losses, ssims = [], []
    for step in range(3):
        x, t = make_pairs(2, 32, 32, seed=500 + step)
        opt.zero_grad(set_to_none=True)
        out = model(x)
        loss = crit(out, t)
        loss.backward()
        opt.step()
        losses.append(loss.item())
        with torch.no_grad():
            ssims.append(metric(out, t).item())
    d = {"losses": np.array(losses), "ssims": np.array(ssims)}
    for k, v in model.state_dict().items():
        _store_tensor(d, "param:" + k, v)
    np.savez_compressed(os.path.join(OUT, "train3.npz"), **d)
    print("train3 losses", losses)


def trained_case(UNetSuperRes, ref_losses):
    """A weight set that is NOT random init: the reference model trained for 60 steps of its own loop body
    (scripts/train.py:296-318: Adam, L1 + 0.4 SSIM) on synthetic pairs, then evaluated on two held-out pairs.  After those steps
    the final sigmoid is no longer saturating (SSIM 0.06 -> 0.78), so the literal north_star bounds for bf16 (output and loss
    within 1e-2 relative, SSIM within 1e-3, PSNR within 0.05 dB) can be asserted on it (tests/test_gpu_parity.py)."""
    from oracle.unet_oracle import init_params
    from oracle.synth import make_pairs
    f = 16
    model = UNetSuperRes(1, 1, base_filters=f)
    model.load_state_dict(init_params(f, seed=1234), strict=True)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=2e-3, weight_decay=1e-5)
    crit = ref_losses.CombinedLoss(ssim_weight=0.4)
    losses = []
    for step in range(60):
        x, t = make_pairs(4, 48, 48, seed=9000 + step % 16)
        opt.zero_grad(set_to_none=True)
        loss = crit(model(x), t)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    model.eval()
    x, t = make_pairs(2, 64, 64, seed=777)
    metric = ref_losses.SSIM(device=torch.device("cpu"))
    with torch.no_grad():
        out = model(x)
        d = {"meta": np.array([f, 2, 64, 64]), "train_losses": np.array(losses),
             "x_checksum": np.array([float(x.double().sum()), float(t.double().sum())]),
             "out": out.numpy(), "loss": np.float64(crit(out, t).item()), "ssim": np.float64(metric(out, t).item()),
             "mse": np.float64(((out - t) ** 2).mean().item())}
    for k, v in model.state_dict().items():
        d["param:" + k] = v.detach().numpy()
    np.savez_compressed(os.path.join(OUT, "trained_f16.npz"), **d)
    print("trained_f16: loss", losses[0], "->", losses[-1], "eval ssim", d["ssim"])


def perceptual_case_large(ref_losses):
    """The VGG19[:36] branch at HR 256 x 256 (the 64 x 64 case leaves 4 x 4 images at the deepest layers: no multi-tile launch of the
    512-channel kernels): features, perceptual value, total loss and the gradient wrt the prediction, stored as summaries;
    inputs are re-derived from seeds by the test (checksums stored)."""
    import torchvision.models as tvm
    from oracle.losses_oracle import make_vgg_params
    from oracle.synth import make_pairs
    vggp = make_vgg_params(seed=1235)
    orig = tvm.vgg19

    def patched(weights=None, **kw):
        m = orig(weights=None, **kw)
        convs = [l for l in m.features if isinstance(l, torch.nn.Conv2d)]
        with torch.no_grad():
            for l, (wt, b) in zip(convs, vggp):
                l.weight.copy_(wt)
                l.bias.copy_(b)
        return m

    tvm.vgg19 = patched
    try:
        _, hr = make_pairs(1, 128, 128, seed=98)            # HR 256 x 256
        g = torch.Generator("cpu").manual_seed(6)
        pred0 = (hr + 0.1 * torch.randn(hr.shape, generator=g)).clamp(0, 1)
        d = {"checksum": np.array([float(pred0.double().sum()), float(hr.double().sum())])}
        crit = ref_losses.CombinedLoss(ssim_weight=0.4, perceptual_weight=0.1, perceptual_loss_type="l1")
        p = pred0.clone().requires_grad_(True)
        loss = crit(p, hr)
        loss.backward()
        with torch.no_grad():
            feats = crit.perceptual_loss.feature_extractor(pred0)
            perc = crit.perceptual_loss(pred0, hr)
        d["loss"] = np.float64(loss.item())
        d["perc"] = np.float64(perc.item())
        _store_tensor(d, "feat", feats)
        _store_tensor(d, "grad", p.grad)
        d["grad_crop"] = p.grad[0, 0, 96:160, 96:160].numpy()
        d["feat_shape"] = np.array(feats.shape)
        np.savez_compressed(os.path.join(OUT, "perceptual_256.npz"), **d)
        print("perceptual_256: loss", d["loss"], "perc", d["perc"], "feat", tuple(feats.shape))
    finally:
        tvm.vgg19 = orig


def preproc_case():
    """Input-pipeline goldens from the reference's own code: scripts/infer.py preprocess_image (needs
    matplotlib / skimage at import: stubbed, they are not on its path) on PNG files, and utils/dataset.py
    MRISuperResDataset.augment_pair + transforms.ToTensor under seeded `random` / `np.random`."""
    import random
    import tempfile
    from unittest import mock
    from PIL import Image
    for name in ("matplotlib", "matplotlib.pyplot", "skimage", "skimage.exposure", "skimage.metrics", "nibabel"):
        sys.modules.setdefault(name, mock.MagicMock())
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(REF, "scripts"))
    from scripts.infer import preprocess_image
    from utils.dataset import MRISuperResDataset
    import torchvision.transforms as T
    sys.path.pop(0)
    sys.path.pop(0)
    rng = np.random.default_rng(2024)
    d = {}
    shapes = [(64, 48), (40, 72), (17, 23), (32, 32)]
    with tempfile.TemporaryDirectory() as tmp:
        for i, (h, w) in enumerate(shapes):
            if i == 3:
                img = np.full((h, w), 77, np.uint8)                      # constant image: max == min, left un-normalised
            elif i == 2:
                img = rng.integers(0, 256, (h, w)).astype(np.uint8)       # flat histogram: the percentiles fall between levels
            else:
                img = rng.normal(90 + 30 * i, 45, (h, w)).clip(0, 255).astype(np.uint8)
            path = os.path.join(tmp, f"s{i}.png")
            Image.fromarray(img).save(path)
            _, t = preprocess_image(path)
            d[f"pp{i}:img"] = img
            d[f"pp{i}:out"] = t[0, 0].numpy()
    ds = object.__new__(MRISuperResDataset)
    ds.aug_params = {"flip_prob": 0.5, "rotate_prob": 0.5, "rotate_range": (-5, 5), "brightness_prob": 0.3, "brightness_range": (0.9, 1.1),
                     "contrast_prob": 0.3, "contrast_range": (0.9, 1.1), "noise_prob": 0.2, "noise_std": 0.01}
    to_tensor = T.ToTensor()
    seeds = [7, 8, 4, 24, 736, 5, 10, 50]          # chosen so that every branch (and all / none together) occurs
    d["aug:seeds"] = np.array(seeds)
    for i, seed in enumerate(seeds):
        h, w = 24 + 8 * (i % 3), 40 - 8 * (i % 2)
        hr = rng.normal(110, 50, (2 * h, 2 * w)).clip(0, 255).astype(np.uint8)
        lr = rng.normal(110, 50, (h, w)).clip(0, 255).astype(np.uint8)
        random.seed(seed)
        np.random.seed(seed)
        f, l = ds.augment_pair(Image.fromarray(hr), Image.fromarray(lr))
        d[f"aug{i}:hr"], d[f"aug{i}:lr"] = hr, lr
        d[f"aug{i}:out_hr"], d[f"aug{i}:out_lr"] = np.array(f), np.array(l)
        d[f"aug{i}:tensor_lr"] = to_tensor(l)[0].numpy()
    np.savez_compressed(os.path.join(OUT, "preproc.npz"), **d)
    print("preproc: ", len(shapes), "preprocess cases,", len(seeds), "augmented pairs")


def eval_case():
    """Evaluation / LR-synthesis goldens: cv2's own x2 upscalers as scripts/test_comparison.py:92-134
    calls them, and the reference's utils/preprocessing.py robust_normalize / simulate_low_field_mri / letterbox_resize(AREA)
    under a seeded np.random (the two normal draws are stored so that the other side can be fed the same noise)."""
    import cv2
    sys.path.insert(0, REF)
    from utils.preprocessing import InterpolationMethod, letterbox_resize, robust_normalize, simulate_low_field_mri
    sys.path.pop(0)
    rng = np.random.default_rng(77)
    d = {}
    kern = np.array([[-1, -1, -1], [-1, 9, -1], [-1, -1, -1]])
    for i, (h, w) in enumerate([(48, 64), (33, 21), (7, 5)]):
        img = rng.normal(120, 60, (h, w)).clip(0, 255).astype(np.uint8)
        d[f"up{i}:img"] = img
        lin = cv2.resize(img, (2 * w, 2 * h), interpolation=cv2.INTER_LINEAR)
        d[f"up{i}:bilinear"] = lin
        d[f"up{i}:sharp_bilinear"] = np.clip(cv2.filter2D(lin, -1, kern), 0, 255)
        d[f"up{i}:bicubic"] = cv2.resize(img, (2 * w, 2 * h), interpolation=cv2.INTER_CUBIC)
    for i, (h, w) in enumerate([(64, 64), (48, 80)]):
        vol = rng.gamma(2.0, 150.0, (h, w))                                   # NIfTI-like intensities, float64
        norm = robust_normalize(vol)
        np.random.seed(100 + i)
        sim = simulate_low_field_mri(norm, kspace_crop_factor=0.5, noise_std=5)
        np.random.seed(100 + i)
        sd = (5 / 255.0) * np.sqrt(h * w) / 10
        nr, ni = np.random.normal(0, sd, (h, w)), np.random.normal(0, sd, (h, w))
        lr = letterbox_resize(np.clip(sim, 0, 1), (w // 2, h // 2), interpolation=InterpolationMethod.AREA)
        d[f"ks{i}:vol"], d[f"ks{i}:norm"], d[f"ks{i}:sim"] = vol, norm, sim
        d[f"ks{i}:noise_re"], d[f"ks{i}:noise_im"] = nr, ni
        d[f"ks{i}:lr_u8"] = np.clip(lr * 255, 0, 255).astype(np.uint8)
    np.savez_compressed(os.path.join(OUT, "eval.npz"), **d)
    print("eval: 3 upscaler cases, 2 k-space cases")


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(os.cpu_count() or 1)
    UNetSuperRes, ref_losses = _import_reference()
    unet_case(UNetSuperRes, ref_losses, "unet_f16_n2_h32", 16, 2, 32, 32, True, 0.4)
    unet_case(UNetSuperRes, ref_losses, "unet_f16_n1_h24x40", 16, 1, 24, 40, True, 0.3)
    unet_case(UNetSuperRes, ref_losses, "unet_f32_n1_h64", 32, 1, 64, 64, False, 0.4, full_out=True)
    loss_cases(ref_losses)
    perceptual_case(ref_losses)
    train_case(UNetSuperRes, ref_losses)
    trained_case(UNetSuperRes, ref_losses)
    perceptual_case_large(ref_losses)
    preproc_case()
    eval_case()


if __name__ == "__main__":
    main()
