"""Pins the oracle (oracle/*.py) against outputs of the REFERENCE modules themselves, generated
by oracle/make_golden.py in the build container and committed under tests/golden/."""
import os

import numpy as np
import pytest
import torch

from oracle import losses_oracle as LO
from oracle import unet_oracle as UO
from oracle.synth import make_pairs
from oracle.train_oracle import OracleTrainer


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


def _check_tensor(d, key, t, rtol, atol):
    t = t.detach()
    if key in d:
        np.testing.assert_allclose(t.numpy(), d[key], rtol=rtol, atol=atol, err_msg=key)
    else:
        flat = t.double().flatten()
        np.testing.assert_allclose(flat[:512].float().numpy(), d[key + "@head"], rtol=rtol, atol=atol, err_msg=key)
        assert abs(float(flat.norm()) - float(d[key + "@norm"])) <= rtol * float(d[key + "@norm"]) + atol, key
        assert abs(float(flat.sum()) - float(d[key + "@sum"])) <= 1e-3 * float(d[key + "@norm"]) + atol, key


@pytest.mark.parametrize("name", ["unet_f16_n2_h32", "unet_f16_n1_h24x40", "unet_f32_n1_h64"])
def test_unet_oracle_matches_reference(golden_dir, name):
    d = _load(golden_dir, name)
    f, n, h, w, affine = [int(v) for v in d["meta"]]
    params = UO.init_params(f, seed=1234)
    if affine:
        params = UO.randomize_affine(params)
    cs = np.array([float(sum(v.double().sum() for v in params.values())),
                   float(sum((v.double() ** 2).sum() for v in params.values()))])
    np.testing.assert_allclose(cs, d["param_checksum"], rtol=1e-12)
    x, target = make_pairs(n, h, w, seed=1234)
    np.testing.assert_allclose([float(x.double().sum()), float(target.double().sum())], d["x_checksum"], rtol=1e-12)
    p = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    taps = {}
    out = UO.unet_forward(p, x, taps)
    out.retain_grad()
    comps = {}
    loss = LO.combined_loss(out, target, ssim_weight=float(d["ssim_weight"]), components=comps)
    loss.backward()
    assert out.shape == (n, 1, 2 * h, 2 * w)
    _check_tensor(d, "out", out, 1e-5, 1e-6)
    _check_tensor(d, "dout", out.grad, 1e-4, 2e-8)
    assert abs(loss.item() - float(d["loss"])) < 1e-6
    assert abs(comps["l1"].item() - float(d["l1"])) < 1e-6
    assert abs(LO.ssim(out, target).item() - float(d["ssim"])) < 1e-6
    for k, v in p.items():
        _check_tensor(d, "grad:" + k, v.grad, 2e-3, 2e-7)
    ref_names = {"inc": "inc", "down1": "down1.maxpool_conv.1", "down2": "down2.maxpool_conv.1",
                 "down3": "down3.maxpool_conv.1", "up1": "up1.conv", "up2": "up2.conv", "up3": "up3.conv"}
    for k, ok in ref_names.items():
        _check_tensor(d, "act:" + k, taps[ok], 1e-4, 1e-5)


def test_param_specs_match_reference_layout():
    specs = UO.param_specs(32)
    assert len(specs) == 64
    assert sum(int(np.prod(s)) if s else 1 for _, s in specs) == 1823122


def test_losses_oracle_matches_reference(golden_dir):
    d = _load(golden_dir, "losses")
    np.testing.assert_allclose(LO.gaussian_taps().numpy(), d["taps"], rtol=0, atol=0)
    np.testing.assert_allclose(LO.gaussian_window2d().numpy(), d["window2d"][0, 0], rtol=0, atol=0)
    i = 0
    while f"s{i}:pred" in d:
        pred = torch.from_numpy(d[f"s{i}:pred"])
        tgt = torch.from_numpy(d[f"s{i}:target"])
        assert abs(LO.ssim(pred, tgt).item() - float(d[f"s{i}:ssim"])) < 1e-6
        np.testing.assert_allclose(LO.ssim(pred, tgt, size_average=False).numpy(), d[f"s{i}:ssim_per_sample"], atol=1e-6)
        j = 0
        while f"s{i}:w{j}:sw" in d:
            p = pred.clone().requires_grad_(True)
            loss = LO.combined_loss(p, tgt, ssim_weight=float(d[f"s{i}:w{j}:sw"]))
            loss.backward()
            assert abs(loss.item() - float(d[f"s{i}:w{j}:loss"])) < 1e-6
            np.testing.assert_allclose(p.grad.numpy(), d[f"s{i}:w{j}:grad"], rtol=1e-4, atol=2e-8)
            j += 1
        i += 1
    assert i == 5
    x = torch.rand(1, 1, 32, 32)
    assert abs(LO.ssim(x, x).item() - 1.0) < 1e-6


def test_loss_argument_validation():
    x = torch.rand(1, 1, 16, 16)
    for sw, pw in ((-0.1, 0.0), (1.1, 0.0), (0.5, -0.1), (0.7, 0.5)):
        with pytest.raises(ValueError):
            LO.combined_loss(x, x, ssim_weight=sw, perceptual_weight=pw)
    with pytest.raises(ValueError):
        LO.combined_loss(x, x, 0.1, 0.5, LO.make_vgg_params(), perceptual_loss_type="huber")


def test_perceptual_oracle_matches_reference(golden_dir):
    d = _load(golden_dir, "perceptual")
    vggp = LO.make_vgg_params(seed=1235)
    pred = torch.from_numpy(d["pred"])
    tgt = torch.from_numpy(d["target"])
    with torch.no_grad():
        _check_tensor(d, "feat", LO.vgg_features(vggp, pred), 1e-4, 1e-6)
    for tag in "abc":
        sw, pw, typ = d[f"{tag}:cfg"]
        p = pred.clone().requires_grad_(True)
        comps = {}
        loss = LO.combined_loss(p, tgt, float(sw), float(pw), vggp, "l1" if typ == 1 else "mse", components=comps)
        loss.backward()
        assert abs(loss.item() - float(d[f"{tag}:loss"])) < 1e-6
        assert abs(comps["perceptual"].item() - float(d[f"{tag}:perc"])) < 1e-7
        np.testing.assert_allclose(p.grad.numpy(), d[f"{tag}:grad"], rtol=2e-3, atol=2e-8)


def test_train_oracle_matches_reference(golden_dir):
    d = _load(golden_dir, "train3")
    params = UO.randomize_affine(UO.init_params(16, seed=1234))
    tr = OracleTrainer(params, lr=1e-3, weight_decay=1e-5, ssim_weight=0.4)
    for step in range(3):
        x, t = make_pairs(2, 32, 32, seed=500 + step)
        loss, s = tr.step(x, t)
        assert abs(loss - d["losses"][step]) < 2e-6
        assert abs(s - d["ssims"][step]) < 2e-6
    for k, v in tr.params.items():
        _check_tensor(d, "param:" + k, v, 1e-4, 2e-6)


def test_psnr_definition():
    a = torch.zeros(4, 4)
    assert LO.psnr(a, a) == 100.0
    assert abs(LO.psnr(a, a + 0.1) - 20.0) < 1e-4


def test_oracle_matches_reference_on_trained_weights(golden_dir):
    """tests/golden/trained_f16.npz: the reference model after 60 of its own training steps (oracle/make_golden.py:trained_case);
    the oracle forward / loss / SSIM on those weights must reproduce the reference's evaluation."""
    d = _load(golden_dir, "trained_f16")
    f, n, h, w = [int(v) for v in d["meta"]]
    params = {k[len("param:"):]: torch.from_numpy(v) for k, v in d.items() if k.startswith("param:")}
    assert list(params.keys()) == [name for name, _ in UO.param_specs(f)]
    x, t = make_pairs(n, h, w, seed=777)
    np.testing.assert_allclose([float(x.double().sum()), float(t.double().sum())], d["x_checksum"], rtol=1e-12)
    with torch.no_grad():
        out = UO.unet_forward(params, x)
    np.testing.assert_allclose(out.numpy(), d["out"], rtol=1e-5, atol=1e-6)
    assert abs(float(LO.combined_loss(out, t, ssim_weight=0.4)) - float(d["loss"])) < 1e-6
    assert abs(float(LO.ssim(out, t)) - float(d["ssim"])) < 1e-6
    assert float(d["ssim"]) > 0.7 and d["train_losses"][-1] < 0.25 * d["train_losses"][0]       # really trained


def test_vgg_oracle_matches_reference_at_256(golden_dir):
    """tests/golden/perceptual_256.npz: reference CombinedLoss with the VGG19[:36] branch at HR 256 x 256 (seeded weights)."""
    d = _load(golden_dir, "perceptual_256")
    _, hr = make_pairs(1, 128, 128, seed=98)
    g = torch.Generator("cpu").manual_seed(6)
    pred = (hr + 0.1 * torch.randn(hr.shape, generator=g)).clamp(0, 1)
    np.testing.assert_allclose([float(pred.double().sum()), float(hr.double().sum())], d["checksum"], rtol=1e-12)
    vggp = LO.make_vgg_params(seed=1235)
    p = pred.clone().requires_grad_(True)
    loss = LO.combined_loss(p, hr, ssim_weight=0.4, perceptual_weight=0.1, vgg_params=vggp)
    loss.backward()
    assert abs(loss.item() - float(d["loss"])) < 1e-5 * float(d["loss"])
    with torch.no_grad():
        feat = LO.vgg_features(vggp, pred)
    assert tuple(feat.shape) == tuple(int(v) for v in d["feat_shape"])
    _check_tensor(d, "feat", feat, 1e-4, 1e-6)
    _check_tensor(d, "grad", p.grad, 1e-3, 1e-9)
    np.testing.assert_allclose(p.grad[0, 0, 96:160, 96:160].numpy(), d["grad_crop"], rtol=1e-3, atol=1e-9)
