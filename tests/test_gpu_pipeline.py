"""GPU parity of the adjacent rows of SURVEY.md 8(f): on-device input pipeline (rank 2), k-space LR synthesis (rank 3), batched
evaluation (rank 4).  Through the C ABI (via utils/preprocess.py, kspace.py, evaluation.py) against the committed goldens made
by the reference's own code and, at larger seeded sizes, against the oracle.  uint8 / index work: bit-exact; float normalise:
bit-exact too (same IEEE operations); FFT-based synthesis: 1e-5 of the [0, 1] range (cuFFT vs pocketfft in single precision)."""
import os
import random

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import eval_oracle as EO
from oracle import preproc_oracle as PO


def _dev():
    return torch.device("cuda:0")


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(_dev())


def test_preprocess_images_bit_exact(golden_dir):
    from mri_superresolution_b200.utils.preprocess import preprocess_images, to_tensor
    d = _load(golden_dir, "preproc")
    for i in range(4):
        out, win = preprocess_images(_t(d[f"pp{i}:img"]), return_window=True)
        assert out.shape == (1, 1) + d[f"pp{i}:img"].shape
        assert np.array_equal(out[0, 0].cpu().numpy(), d[f"pp{i}:out"]), i
    # a batch at the path's real size: every slice against numpy
    rng = np.random.default_rng(5)
    imgs = np.stack([rng.normal(60 + 20 * k, 30 + 5 * k, (256, 256)).clip(0, 255).astype(np.uint8) for k in range(6)])
    imgs[5] = rng.integers(0, 3, (256, 256)).astype(np.uint8)         # three gray levels only
    out, win = preprocess_images(_t(imgs), return_window=True)
    for k in range(6):
        ref, lo, hi = PO.preprocess_u8(imgs[k])
        assert np.array_equal(out[k, 0].cpu().numpy(), ref), k
        assert float(win[k, 0]) == float(lo) and float(win[k, 1]) == float(hi)
    assert np.array_equal(to_tensor(_t(imgs))[:, 0].cpu().numpy(), np.stack([PO.to_tensor_u8(x) for x in imgs]))
    with pytest.raises(RuntimeError):
        preprocess_images(torch.zeros(1, 8, 8, dtype=torch.uint8))
    with pytest.raises(ValueError):
        preprocess_images(torch.zeros(1, 8, 8, device=_dev()))


def test_augment_pair_bit_exact_against_reference_goldens(golden_dir):
    from mri_superresolution_b200.utils.preprocess import AugmentDecision, PairAugmenter
    d = _load(golden_dir, "preproc")
    for i, seed in enumerate(d["aug:seeds"]):
        aug = PairAugmenter(rng=random.Random(int(seed)))
        dec = aug.draw()                                               # the reference's decisions from the reference's seed
        lr, hr = d[f"aug{i}:lr"], d[f"aug{i}:hr"]
        np.random.seed(int(seed))
        noise = torch.from_numpy(np.random.normal(0, 0.01 * 255, lr.shape)) if dec.noise else None
        (l, l8), (f, f8) = aug(_t(hr), _t(lr), decisions=[dec], noise=noise, as_uint8=True)
        assert np.array_equal(f8[0].cpu().numpy(), d[f"aug{i}:out_hr"]), (i, "hr")
        assert np.array_equal(l8[0].cpu().numpy(), d[f"aug{i}:out_lr"]), (i, "lr")
        assert np.array_equal(l[0, 0].cpu().numpy(), d[f"aug{i}:tensor_lr"]), (i, "tensor")


def test_augment_batch_matches_oracle_at_full_size():
    """A 12-pair batch at 256 / 512 with every branch forced on somewhere; per-sample decisions differ inside one launch."""
    from mri_superresolution_b200.utils.preprocess import AugmentDecision, PairAugmenter
    rng = np.random.default_rng(9)
    n = 12
    hr = np.stack([rng.normal(100, 50, (512, 512)).clip(0, 255).astype(np.uint8) for _ in range(n)])
    lr = np.stack([rng.normal(100, 50, (256, 256)).clip(0, 255).astype(np.uint8) for _ in range(n)])
    r = random.Random(1)
    decs = []
    for k in range(n):
        decs.append(AugmentDecision(flip=k % 2 == 1, angle=r.uniform(-5, 5) if k % 3 else None,
                                    brightness=r.uniform(0.9, 1.1) if k % 4 < 2 else None,
                                    contrast=r.uniform(0.9, 1.1) if k % 5 < 3 else None, noise=k % 2 == 0))
    decs[7].angle = 180.0 - 1e-9
    noise = rng.normal(0, 2.55, lr.shape)
    aug = PairAugmenter()
    (l, l8), (f, f8) = aug(_t(hr), _t(lr), decisions=decs, noise=torch.from_numpy(noise), as_uint8=True)
    for k, dc in enumerate(decs):
        tup = (dc.flip, dc.angle, dc.brightness, dc.contrast, dc.noise)
        assert np.array_equal(f8[k].cpu().numpy(), PO.augment_u8(hr[k], tup)), (k, "hr")
        assert np.array_equal(l8[k].cpu().numpy(), PO.augment_u8(lr[k], tup, noise[k])), (k, "lr")
    # device-side noise (no explicit array): the right size and not the same twice
    decs2 = [AugmentDecision(noise=True, seed=1234 + k) for k in range(n)]
    (l1, a8), _ = aug(_t(hr), _t(lr), decisions=decs2, as_uint8=True)
    diff = a8.float() - _t(lr).float()
    interior = (_t(lr) > 8) & (_t(lr) < 247)                           # away from the clip
    sd = float(diff[interior].std())
    assert 0.6 < sd / 2.55 < 1.25, sd                                  # truncation of x + N(0, 2.55) widens / shifts it a little
    assert abs(float(diff[interior].mean()) + 0.5) < 0.1               # astype(uint8) truncates: mean shift of -0.5 gray levels
    assert not torch.equal(a8[0] - _t(lr)[0], a8[1] - _t(lr)[1])


def test_classical_upscalers_bit_exact_against_cv2_goldens(golden_dir):
    from mri_superresolution_b200.utils.evaluation import upscale_with_interpolation
    d = _load(golden_dir, "eval")
    for i in range(3):
        for m in ("bilinear", "sharp_bilinear", "bicubic"):
            got = upscale_with_interpolation(_t(d[f"up{i}:img"]), m, as_uint8=True)[0].cpu().numpy()
            assert np.array_equal(got, d[f"up{i}:{m}"]), (i, m, int((got != d[f"up{i}:{m}"]).sum()))
    rng = np.random.default_rng(3)
    imgs = np.stack([rng.normal(120, 70, (96, 130)).clip(0, 255).astype(np.uint8) for _ in range(3)])
    for m in ("bilinear", "sharp_bilinear", "bicubic"):
        got = upscale_with_interpolation(_t(imgs), m).cpu().numpy()
        for k in range(3):
            assert np.array_equal(got[k], EO.upscale_with_interpolation(imgs[k], m)), (k, m)
    with pytest.raises(ValueError):
        upscale_with_interpolation(_t(imgs), "lanczos")


def test_batched_metrics_match_the_reference_formulae():
    from mri_superresolution_b200.utils.evaluation import calculate_metrics
    from oracle import losses_oracle as LO
    rng = np.random.default_rng(4)
    hr = rng.random((5, 64, 80)).astype(np.float32)
    up = np.clip(hr + 0.05 * rng.standard_normal(hr.shape), 0, 1).astype(np.float32)
    up[4] = hr[4]                                                      # identical pair: PSNR guard -> 100
    m = calculate_metrics(_t(hr), _t(up))
    for k in range(5):
        ref = EO.image_metrics(hr[k], up[k])
        for key in ("mse", "rmse", "mae", "psnr"):
            assert abs(float(m[key][k]) - ref[key]) <= 2e-6 * max(abs(ref[key]), 1e-3), (k, key, float(m[key][k]), ref[key])
        s = float(LO.ssim(torch.from_numpy(up[k])[None, None], torch.from_numpy(hr[k])[None, None]))
        assert abs(float(m["ssim"][k]) - s) < 1e-5
    assert float(m["psnr"][4]) == 100.0


def test_match_histograms_matches_the_restated_algorithm():
    from mri_superresolution_b200.utils.evaluation import match_histograms
    from mri_superresolution_b200.utils.preprocess import preprocess_images
    rng = np.random.default_rng(6)
    out = (1 / (1 + np.exp(-rng.normal(0, 2, (3, 48, 40))))).astype(np.float32)
    out[1, :10] = 0.25                                                  # a long run of equal values
    tgt_u8 = np.stack([rng.normal(100, 40, (64, 56)).clip(0, 255).astype(np.uint8) for _ in range(3)])
    tgt = preprocess_images(_t(tgt_u8))[:, 0]                           # the reference normalises the target the same way (:292-299)
    got = match_histograms(_t(out), tgt, clip=False).cpu().numpy()
    for k in range(3):
        ref = EO.match_histograms(out[k], tgt[k].cpu().numpy())
        assert np.array_equal(got[k], ref.astype(np.float32)), (k, float(np.abs(got[k] - ref).max()))
    same = match_histograms(_t(out), _t(out)).cpu().numpy()
    assert np.allclose(same, out, atol=1e-7)


def test_kspace_simulation_matches_reference_goldens(golden_dir):
    from mri_superresolution_b200.utils.kspace import area_half, robust_normalize, simulate_low_field_mri
    d = _load(golden_dir, "eval")
    for i in range(2):
        norm = robust_normalize(_t(d[f"ks{i}:vol"]))
        assert np.array_equal(norm[0].cpu().numpy(), d[f"ks{i}:norm"]), i
        sim = simulate_low_field_mri(norm, noise=(_t(d[f"ks{i}:noise_re"]), _t(d[f"ks{i}:noise_im"])))
        assert sim.dtype == torch.float64
        err = float(np.abs(sim[0].cpu().numpy() - d[f"ks{i}:sim"]).max())
        assert err < 1e-5, (i, err)
        lr = area_half(sim, clip01=True, as_uint8=True)[0].cpu().numpy().astype(np.int32)
        # uint8 truncation of a value that differs by 1e-6: at most one gray level, on at most a handful of pixels
        delta = np.abs(lr - d[f"ks{i}:lr_u8"].astype(np.int32))
        assert delta.max() <= 1 and (delta > 0).mean() < 5e-3, (i, int(delta.max()), float((delta > 0).mean()))
        exact = area_half(_t(np.clip(d[f"ks{i}:sim"], 0, 1)), as_uint8=True)[0].cpu().numpy()
        assert np.array_equal(exact, d[f"ks{i}:lr_u8"])                 # fed the reference's own simulation: bit-exact
    # batch at the path's size with device-side noise: statistics of the reference's construction
    hr = torch.rand(4, 256, 256, device=_dev())
    a = simulate_low_field_mri(hr, seed=1)
    b = simulate_low_field_mri(hr, seed=2)
    assert a.shape == (4, 256, 256) and not torch.equal(a, b)
    assert float(a.amin()) >= float(hr.amin()) - 1e-9 and float(a.amax()) <= float(hr.amax()) + 1e-9
    clean = simulate_low_field_mri(hr, noise_std=0)
    ref = EO.simulate_low_field(hr[0].cpu().numpy(), np.zeros((256, 256)), np.zeros((256, 256)))
    assert float(np.abs(clean[0].cpu().numpy() - ref).max()) < 1e-5


def _png_pairs(root, n, h, w, seed=0):
    from PIL import Image
    from oracle.synth import make_pairs
    hr_dir, lr_dir = os.path.join(root, "hr"), os.path.join(root, "lr")
    os.makedirs(hr_dir), os.makedirs(lr_dir)
    lr, hr = make_pairs(n, h, w, seed=seed)
    for i in range(n):
        name = f"sub-S{i // 4:02d}_s{i:03d}.png"
        Image.fromarray((hr[i, 0].numpy() * 255).astype(np.uint8)).save(os.path.join(hr_dir, name))
        Image.fromarray((lr[i, 0].numpy() * 255).astype(np.uint8)).save(os.path.join(lr_dir, name))
    return hr_dir, lr_dir


def test_train_and_infer_command_lines_on_png_pairs(tmp_path):
    """The whole caller side on real files: PNG pairs -> `python -m mri_superresolution_b200.train` (uint8 batches, device-side
    ToTensor + augmentation, graph replay, ragged last batch, JSON-lines, checkpoints) -> `python -m mri_superresolution_b200.infer`
    on one file with a target (histogram matching + metrics) and on a directory."""
    import json
    import subprocess
    import sys
    from PIL import Image
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hr_dir, lr_dir = _png_pairs(str(tmp_path), 22, 32, 32)
    ck = str(tmp_path / "ck")
    env = dict(os.environ, PYTHONPATH=root)
    cmd = [sys.executable, "-m", "mri_superresolution_b200.train", "--full_res_dir", hr_dir, "--low_res_dir", lr_dir, "--base_filters", "16",
           "--batch_size", "4", "--epochs", "3", "--learning_rate", "2e-3", "--ssim_weight", "0.4", "--seed", "7", "--augmentation",
           "--checkpoint_dir", ck, "--num_workers", "2"]
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=str(tmp_path), timeout=600)
    assert res.returncode == 0, res.stderr[-3000:]
    lines = [json.loads(l) for l in res.stdout.splitlines() if l.startswith("{")]
    summ = [l for l in lines if l["type"] == "epoch_summary"]
    assert len(summ) == 3 and summ[-1]["train_loss"] < summ[0]["train_loss"]
    assert os.path.exists(os.path.join(ck, "best_model_unet.pth")) and os.path.exists(os.path.join(ck, "final_model_unet.pth"))
    name = sorted(os.listdir(lr_dir))[0]
    out_png = str(tmp_path / "sr.png")
    cmd = [sys.executable, "-m", "mri_superresolution_b200.infer", "--input", os.path.join(lr_dir, name), "--output", out_png,
           "--target", os.path.join(hr_dir, name), "--checkpoint_dir", ck, "--base_filters", "16"]
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=str(tmp_path), timeout=300)
    assert res.returncode == 0, res.stderr[-3000:]
    rec = [json.loads(l) for l in res.stdout.splitlines() if l.startswith("{") and '"inference"' in l][0]
    assert {"ssim", "psnr", "mse", "mae", "rmse"} <= set(rec) and 0 < rec["ssim"] <= 1
    img = np.asarray(Image.open(out_png))
    assert img.shape == (64, 64) and img.dtype == np.uint8
    out_dir = str(tmp_path / "sr_dir")
    cmd = [sys.executable, "-m", "mri_superresolution_b200.infer", "--input", lr_dir, "--output", out_dir, "--checkpoint_dir", ck,
           "--base_filters", "16", "--batch_size", "8"]
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=str(tmp_path), timeout=300)
    assert res.returncode == 0, res.stderr[-3000:]
    assert len(os.listdir(out_dir)) == 22
    # the single-file result without a target equals the directory run's file for the same slice (batch position is immaterial)
    a = np.asarray(Image.open(os.path.join(out_dir, name))).astype(int)
    cmd = [sys.executable, "-m", "mri_superresolution_b200.infer", "--input", os.path.join(lr_dir, name), "--output", out_png,
           "--checkpoint_dir", ck, "--base_filters", "16"]
    assert subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=str(tmp_path), timeout=300).returncode == 0
    b = np.asarray(Image.open(out_png)).astype(int)
    assert np.abs(a - b).max() <= 1


def test_train_command_line_under_torchrun_two_gpus(tmp_path):
    import json
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hr_dir, lr_dir = _png_pairs(str(tmp_path), 40, 32, 32)
    ck = str(tmp_path / "ck")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29581", "-m", "mri_superresolution_b200.train", "--full_res_dir", hr_dir, "--low_res_dir", lr_dir,
           "--base_filters", "16", "--batch_size", "4", "--epochs", "2", "--seed", "3", "--checkpoint_dir", ck, "--num_workers", "2"]
    res = subprocess.run(cmd, capture_output=True, text=True, env=dict(os.environ, PYTHONPATH=root), cwd=str(tmp_path), timeout=170)
    assert res.returncode == 0, res.stderr[-3000:]
    summ = [json.loads(l) for l in res.stdout.splitlines() if l.startswith("{") and "epoch_summary" in l]
    assert len(summ) == 2                                            # rank 0 alone logs
    assert os.path.exists(os.path.join(ck, "final_model_unet.pth"))
