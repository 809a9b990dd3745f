"""GPU parity tests: the CUDA path (through the C ABI, via the drop-in modules) against the oracle and the
committed golden vectors.  Tolerances (BASELINE.json north_star): activations and loss within 1e-4
relative in fp32 mode and 1e-2 in bf16 mode; PSNR within 0.05 dB, SSIM within 1e-3.
'relative' is measured against the tensor's max magnitude: |a-b|_inf <= tol * |b|_inf."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import losses_oracle as LO
from oracle import unet_oracle as UO
from oracle.synth import make_pairs
from oracle.train_oracle import OracleTrainer


def _dev():
    return torch.device("cuda:0")


def _golden(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


def rel_inf(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-30))


def _model(f, precision, params=None):
    from mri_superresolution_b200.models.unet_model import UNetSuperRes
    m = UNetSuperRes(base_filters=f, precision=precision).to(_dev())
    if params is not None:
        m.load_state_dict(params, strict=True)
    return m


def test_native_library_loaded():
    from mri_superresolution_b200 import _native
    assert _native.lib().mrisr_version() == 100


# ------------------------------------------------------------------ loss kernel
def test_loss_kernel_matches_golden(golden_dir):
    from mri_superresolution_b200.utils.losses import CombinedLoss, ssim
    d = _golden(golden_dir, "losses")
    i = 0
    while f"s{i}:pred" in d:
        pred = torch.from_numpy(d[f"s{i}:pred"]).to(_dev())
        tgt = torch.from_numpy(d[f"s{i}:target"]).to(_dev())
        assert abs(ssim(pred, tgt).item() - float(d[f"s{i}:ssim"])) < 1e-5
        np.testing.assert_allclose(ssim(pred, tgt, size_average=False).cpu().numpy(), d[f"s{i}:ssim_per_sample"], atol=1e-5)
        j = 0
        while f"s{i}:w{j}:sw" in d:
            p = pred.clone().requires_grad_(True)
            crit = CombinedLoss(ssim_weight=float(d[f"s{i}:w{j}:sw"]))
            loss = crit(p, tgt)
            loss.backward()
            assert abs(loss.item() - float(d[f"s{i}:w{j}:loss"])) < 1e-4 * max(1.0, abs(float(d[f"s{i}:w{j}:loss"])))
            g_ref = torch.from_numpy(d[f"s{i}:w{j}:grad"])
            assert rel_inf(p.grad, g_ref) < 2e-4, (i, j, rel_inf(p.grad, g_ref))
            j += 1
        i += 1
    assert i == 5
    x = torch.rand(1, 1, 32, 32, device=_dev())
    assert abs(ssim(x, x).item() - 1.0) < 1e-5


def test_loss_kernel_scaled_backward_and_clamp():
    from mri_superresolution_b200.utils.losses import CombinedLoss
    g = torch.Generator("cpu").manual_seed(3)
    t = torch.rand(2, 1, 48, 40, generator=g).to(_dev())
    p = (1.0 - t).clone().requires_grad_(True)          # anti-correlated: mean SSIM < 0 -> clamp kills the SSIM gradient
    crit = CombinedLoss(ssim_weight=0.4)
    loss = crit(p, t)
    (loss * 8.0).backward()                              # upstream scale (GradScaler-style), scripts/train.py:309
    pr = (1.0 - t).cpu().clone().requires_grad_(True)
    lr = LO.combined_loss(pr, t.cpu(), ssim_weight=0.4)
    (lr * 8.0).backward()
    assert LO.ssim(pr, t.cpu()).item() < 0
    assert abs(loss.item() - lr.item()) < 1e-5
    assert rel_inf(p.grad, pr.grad) < 2e-4


def test_loss_full_size_properties():
    """BASELINE config-2 size (64 x 512^2): size-independent properties instead of an oracle run."""
    from mri_superresolution_b200.utils.losses import CombinedLoss, ssim
    _, hr = make_pairs(8, 256, 256, seed=7)
    hr = hr.to(_dev())
    assert abs(ssim(hr, hr).item() - 1.0) < 1e-5
    crit = CombinedLoss(ssim_weight=0.4)
    assert crit(hr, hr).item() < 1e-5
    p = (hr * 0.9 + 0.05).requires_grad_(True)
    loss = crit(p, hr)
    loss.backward()
    # directional derivative check of the analytic gradient
    v = torch.sign(p.grad)                               # a direction along which the loss really changes
    eps = 1e-3
    with torch.no_grad():
        lp = crit(p + eps * v, hr).item()
        lm = crit(p - eps * v, hr).item()
    num = (lp - lm) / (2 * eps)
    ana = float((p.grad * v).sum())
    assert abs(num - ana) < 2e-2 * max(abs(num), abs(ana), 1e-6), (num, ana)


# ------------------------------------------------------------------ U-Net forward/backward
CASES = ["unet_f16_n2_h32", "unet_f16_n1_h24x40", "unet_f32_n1_h64"]


def _run_case(golden_dir, name, precision):
    from mri_superresolution_b200.utils.losses import CombinedLoss, SSIM
    d = _golden(golden_dir, name)
    f, n, h, w, affine = [int(v) for v in d["meta"]]
    params = UO.init_params(f, seed=1234)
    if affine:
        params = UO.randomize_affine(params)
    x, tgt = make_pairs(n, h, w, seed=1234)
    model = _model(f, precision, params)
    model.train()
    crit = CombinedLoss(ssim_weight=float(d["ssim_weight"]))
    out = model(x.to(_dev()))
    loss = crit(out, tgt.to(_dev()))
    loss.backward()
    metric = SSIM()(out.detach(), tgt.to(_dev())).item()
    return d, model, out, loss, metric, tgt


@pytest.mark.parametrize("name", CASES)
def test_unet_fp32_matches_golden(golden_dir, name):
    d, model, out, loss, metric, tgt = _run_case(golden_dir, name, "fp32")
    assert out.shape == tuple(d["out"].shape) and out.dtype == torch.float32
    ref_out = torch.from_numpy(d["out"])
    assert rel_inf(out, ref_out) < 1e-4, rel_inf(out, ref_out)
    assert abs(loss.item() - float(d["loss"])) < 1e-4 * abs(float(d["loss"]))
    assert abs(metric - float(d["ssim"])) < 1e-4
    assert abs(LO.psnr(out.cpu(), tgt) - LO.psnr(ref_out, tgt)) < 0.05
    for k, p in model.named_parameters():
        if "grad:" + k in d:
            e = rel_inf(p.grad, torch.from_numpy(d["grad:" + k]).view(p.shape))
        else:
            e = abs(float(p.grad.double().norm()) - float(d["grad:" + k + "@norm"])) / float(d["grad:" + k + "@norm"])
            head = torch.from_numpy(d["grad:" + k + "@head"])
            e = max(e, rel_inf(p.grad.flatten()[:512], head) * float(head.abs().max()) / float(p.grad.abs().max()))
        assert e < 2e-3, (k, e)


def rel_l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / max(float(b.norm()), 1e-30))


@pytest.mark.parametrize("name", CASES)
def test_unet_bf16_matches_golden(golden_dir, name):
    """bf16 mode against the fp32 golden output.  Criteria (BASELINE.json north_star): loss within 1e-2 relative, SSIM
    within 1e-3, PSNR within 0.05 dB.  The output tensor itself is bounded by the bf16-operand noise floor of this
    21-conv network: a random-init net turns bf16 rounding of weights/activations into ~1.5-3 % relative-L2 output
    noise whatever computes it (oracle.unet_forward_bf16_storage, fp32 or fp64 arithmetic between the rounding points,
    differs from itself by ~1 %), so the test requires the CUDA path to be no further from fp32 than 2x that floor and
    <= 5e-2 relative L2; per-layer parity at bf16 resolution is pinned by tests/test_gpu_conv_tc.py."""
    d, model, out, loss, metric, tgt = _run_case(golden_dir, name, "bf16")
    ref_out = torch.from_numpy(d["out"])
    f, n, h, w, affine = [int(v) for v in d["meta"]]
    params = UO.init_params(f, seed=1234)
    if affine:
        params = UO.randomize_affine(params)
    x, _ = make_pairs(n, h, w, seed=1234)
    pq = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    out_q = UO.unet_forward_bf16_storage(pq, x)
    LO.combined_loss(out_q, tgt, ssim_weight=float(d["ssim_weight"])).backward()
    floor = rel_l2(out_q, ref_out)
    e = rel_l2(out, ref_out)
    assert e < 5e-2 and e < 2.0 * floor + 2e-3, (e, floor)
    assert abs(loss.item() - float(d["loss"])) < 1e-2 * abs(float(d["loss"]))
    assert abs(metric - float(d["ssim"])) < 1e-3
    assert abs(LO.psnr(out.cpu(), tgt) - LO.psnr(ref_out, tgt)) < 0.05
    # gradients against autograd of the bf16-storage restatement (same operand rounding, fp32 gradient arithmetic)
    gmax = max(float(v.grad.norm()) for v in pq.values())
    for k, p in model.named_parameters():
        r = pq[k].grad.double()
        g = p.grad.double().cpu().view(r.shape)
        if r.numel() > 1 and float(r.norm()) > 1e-3 * gmax:
            cos = float((g * r).sum() / (g.norm() * r.norm() + 1e-30))
            assert cos > 0.9, (k, cos)
        tol = 0.15 * float(r.norm()) + 2e-3 * gmax
        # bf16 noise floor of this gradient: distance between the bf16-storage restatement and the fp32 golden gradient.
        # alpha's gradient is a sum of (xb - xp) * g over every output pixel (cancellation leaves a value far below its
        # terms) and the first layers sit at the end of a 20-layer bf16 chain: on these tiny random-init cases operand
        # rounding alone moves their norms by 20-30 %, so the CUDA path is allowed the same distance.
        if r.numel() == 1 and ("grad:" + k) in d:
            tol += 2.0 * abs(float(r) - float(d["grad:" + k]))
        elif ("grad:" + k + "@norm") in d:
            tol += abs(float(r.norm()) - float(d["grad:" + k + "@norm"]))
        elif ("grad:" + k) in d:
            tol += abs(float(r.norm()) - float(np.linalg.norm(d["grad:" + k].astype(np.float64))))
        assert abs(float(g.norm() if r.numel() > 1 else g) - float(r.norm() if r.numel() > 1 else r)) < tol, k


def test_inference_matches_training_forward_and_clamp(golden_dir):
    d = _golden(golden_dir, "unet_f16_n2_h32")
    params = UO.randomize_affine(UO.init_params(16, seed=1234))
    x, _ = make_pairs(2, 32, 32, seed=1234)
    model = _model(16, "fp32", params).eval()
    with torch.no_grad():
        y = torch.clamp(model(x.to(_dev())), 0, 1)           # scripts/infer.py:275-276
    assert rel_inf(y, torch.from_numpy(d["out"])) < 1e-4
    assert float(y.min()) >= 0 and float(y.max()) <= 1


def test_backward_after_newer_forward_raises():
    model = _model(16, "fp32")
    x, _ = make_pairs(1, 16, 16)
    y1 = model(x.to(_dev()))
    _ = model(x.to(_dev()))
    with pytest.raises(RuntimeError):
        y1.sum().backward()


def test_gradient_accumulation_matches_sum():
    model = _model(16, "fp32", UO.randomize_affine(UO.init_params(16)))
    xa, _ = make_pairs(1, 16, 16, seed=1)
    xb, _ = make_pairs(1, 16, 16, seed=2)
    model(xa.to(_dev())).sum().backward()
    ga = [p.grad.clone() for p in model.parameters()]
    model.zero_grad(set_to_none=True)
    model(xb.to(_dev())).sum().backward()
    gb = [p.grad.clone() for p in model.parameters()]
    model.zero_grad(set_to_none=True)
    model(xa.to(_dev())).sum().backward()
    model(xb.to(_dev())).sum().backward()             # accumulates into .grad
    for p, a, b in zip(model.parameters(), ga, gb):
        assert torch.allclose(p.grad, a + b, rtol=1e-5, atol=1e-7)


def test_gradient_marks_fire_only_when_the_bucket_is_final():
    """mrisr_unet_backward_marked: a side stream that waits on mark k and copies grads[offset_k:] must see exactly the final
    values (the arena still holds another batch's gradients until the kernels of this backward overwrite them)."""
    import ctypes
    from mri_superresolution_b200 import _native
    from mri_superresolution_b200.parallel import SCHEDULE_CUTS, schedule_bounds
    L = _native.lib()
    model = _model(32, "bf16", UO.randomize_affine(UO.init_params(32)))
    xa, _ = make_pairs(4, 64, 64, seed=1)
    xb, _ = make_pairs(4, 64, 64, seed=2)
    model(xa.to(_dev())).square().sum().backward()                      # fills the arena with batch A's gradients
    total = L.mrisr_unet_param_elems(32)
    bounds = schedule_bounds(total, [L.mrisr_unet_grads_final_from(32, l) for l in SCHEDULE_CUTS])
    events = [L.mrisr_event_create() for _ in bounds]
    assert all(events)
    model.set_gradient_marks([s for s, _ in bounds], events)
    arena_a = model.flat_grads()
    model.zero_grad(set_to_none=True)
    side = torch.cuda.Stream()
    model(xb.to(_dev())).square().sum().backward()
    assert model.gradient_marks_recorded() and model.flat_grads().data_ptr() == arena_a.data_ptr()
    flat = model.flat_grads()
    snaps = []
    with torch.cuda.stream(side):
        for (s, e), ev in zip(bounds, events):
            assert L.mrisr_stream_wait_event(ctypes.c_void_p(side.cuda_stream), ctypes.c_void_p(ev)) == 0
            snaps.append(flat[s:e].clone())
    torch.cuda.synchronize()
    for (s, e), snap in zip(bounds, snaps):
        assert torch.equal(snap, flat[s:e])
    # accumulation on top of .grad: the marks no longer describe .grad, the exchange must fall back to a stream wait
    model(xa.to(_dev())).square().sum().backward()
    assert not model.gradient_marks_recorded()
    model.set_gradient_marks(None, None)
    for ev in events:
        assert L.mrisr_event_destroy(ctypes.c_void_p(ev)) == 0


def test_cpu_tensor_raises():
    model = _model(16, "fp32")
    with pytest.raises(RuntimeError):
        model(torch.zeros(1, 1, 16, 16))
    with pytest.raises(ValueError):
        model(torch.zeros(1, 1, 12, 16, device=_dev()))       # H not a multiple of 8


# ------------------------------------------------------------------ optimiser / train step
def test_fused_adam_matches_torch():
    from mri_superresolution_b200.optim import FusedAdam
    model = _model(16, "fp32", UO.randomize_affine(UO.init_params(16)))
    ref = [p.detach().clone().requires_grad_(True) for p in model.parameters()]
    opt_ref = torch.optim.Adam(ref, lr=1e-3, weight_decay=1e-5)
    opt = FusedAdam(model, lr=1e-3, weight_decay=1e-5)
    g = torch.Generator("cpu").manual_seed(0)
    for step in range(3):
        x, _ = make_pairs(1, 16, 16, seed=10 + step)
        opt.zero_grad()
        model(x.to(_dev())).square().mean().backward()
        for r, p in zip(ref, model.parameters()):
            r.grad = p.grad.detach().clone()
        opt.step()
        opt_ref.step()
    for r, p in zip(ref, model.parameters()):
        assert torch.allclose(p, r, rtol=1e-5, atol=1e-7)


def test_three_train_steps_match_golden(golden_dir):
    from mri_superresolution_b200.optim import FusedAdam
    from mri_superresolution_b200.utils.losses import CombinedLoss, SSIM
    d = _golden(golden_dir, "train3")
    model = _model(16, "fp32", UO.randomize_affine(UO.init_params(16, seed=1234)))
    opt = FusedAdam(model, lr=1e-3, weight_decay=1e-5)
    crit, metric = CombinedLoss(ssim_weight=0.4), SSIM()
    for step in range(3):
        x, t = make_pairs(2, 32, 32, seed=500 + step)
        x, t = x.to(_dev()), t.to(_dev())
        opt.zero_grad(set_to_none=True)
        out = model(x)
        loss = crit(out, t)
        loss.backward()
        opt.step()
        assert abs(loss.item() - d["losses"][step]) < 2e-4 * d["losses"][step], step
        with torch.no_grad():
            assert abs(metric(out, t).item() - d["ssims"][step]) < 2e-4
    sd = model.state_dict()
    for k, v in sd.items():
        if "param:" + k in d:
            assert rel_inf(v, torch.from_numpy(d["param:" + k]).view(v.shape)) < 5e-3, k


def test_stock_torch_adam_works_on_arena_views():
    """scripts/train.py:186 builds torch.optim.Adam(model.parameters()): must keep working (drop-in)."""
    from mri_superresolution_b200.utils.losses import CombinedLoss
    params = UO.randomize_affine(UO.init_params(16, seed=1234))
    model = _model(16, "fp32", params)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=1e-5)
    tr = OracleTrainer(params, lr=1e-3, weight_decay=1e-5, ssim_weight=0.4)
    crit = CombinedLoss(ssim_weight=0.4)
    for step in range(2):
        x, t = make_pairs(2, 32, 32, seed=500 + step)
        opt.zero_grad(set_to_none=True)
        loss = crit(model(x.to(_dev())), t.to(_dev()))
        loss.backward()
        opt.step()
        lo, _ = tr.step(x, t)
        assert abs(loss.item() - lo) < 2e-4 * lo
    assert model._is_flat()


def test_graphed_train_step_matches_eager():
    """utils/graph.py: the step replayed as one CUDA graph = the same step launched eagerly (same kernels, same order;
    only the Adam bias correction is computed on the device instead of the host)."""
    from mri_superresolution_b200.optim import FusedAdam
    from mri_superresolution_b200.utils.graph import GraphedTrainStep
    from mri_superresolution_b200.utils.losses import CombinedLoss
    params = UO.randomize_affine(UO.init_params(16, seed=5))
    batches = [make_pairs(2, 32, 48, seed=40 + i) for i in range(4)]
    results = []
    for use_graph in (False, True):
        model = _model(16, "bf16", params)
        crit = CombinedLoss(ssim_weight=0.4)
        opt = FusedAdam(model, lr=1e-3, weight_decay=1e-5, capturable=use_graph)   # eager leg: the host-stepped Adam
        losses = []
        if use_graph:
            x0, t0 = batches[0]
            stepper = GraphedTrainStep(model, crit, opt, x0.to(_dev()), t0.to(_dev()))
            assert opt.step_count == 0                       # warm-up and capture left no trace
        for x, t in batches:
            x, t = x.to(_dev()), t.to(_dev())
            if use_graph:
                losses.append(float(stepper(x, t).item()))
            else:
                opt.zero_grad(set_to_none=True)
                loss = crit(model(x), t)
                loss.backward()
                opt.step()
                losses.append(float(loss.item()))
        assert opt.step_count == len(batches)
        results.append((losses, model.flat_parameters().clone(), opt.exp_avg.clone()))
    (l0, p0, m0), (l1, p1, m1) = results
    for a, b in zip(l0, l1):
        assert abs(a - b) <= 1e-6 * abs(a), (l0, l1)
    assert rel_inf(p1, p0) < 1e-5 and rel_inf(m1, m0) < 1e-5
    with pytest.raises(ValueError):
        GraphedTrainStep(model, crit, FusedAdam(model), batches[0][0].to(_dev()), batches[0][1].to(_dev()))


def test_graphed_forward_matches_eager(golden_dir):
    from mri_superresolution_b200.utils.graph import GraphedForward
    model = _model(16, "bf16", UO.randomize_affine(UO.init_params(16, seed=5))).eval()
    xs = [make_pairs(2, 32, 48, seed=60 + i)[0].to(_dev()) for i in range(3)]
    g = GraphedForward(model, xs[0])
    for x in xs:
        with torch.no_grad():
            ref = torch.clamp(model(x), 0, 1)
        assert torch.equal(g(x), ref)
    with pytest.raises(ValueError):
        g(xs[0][:1])


def test_full_size_forward_properties():
    """Config-2 shape (256^2 input) at a small batch: output range/shape and batch-independence
    (GroupNorm is per-sample, so sample i of a batch equals the same sample run alone)."""
    model = _model(32, "bf16", UO.init_params(32, seed=1234)).eval()
    x, _ = make_pairs(3, 256, 256, seed=11)
    x = x.to(_dev())
    with torch.no_grad():
        y = model(x)
        y1 = model(x[1:2].contiguous())
    assert y.shape == (3, 1, 512, 512)
    assert float(y.min()) >= 0 and float(y.max()) <= 1
    # not bit-equal: the GroupNorm partial sums are grouped by 128-pixel tile, and a sample's tile alignment
    # depends on its position in the batch.  A last-bit change of a (mean, rstd) pair moves bf16 roundings downstream, and
    # this random-init net's steep final sigmoid amplifies that at isolated pixels (as in __graft_entry__.smoke), so the
    # yardstick is the bf16 rounding floor itself: the distance of the bf16 result from the fp32-mode result of the same
    # sample.  Batch position must matter less than bf16 storage does.
    ref = _model(32, "fp32", UO.init_params(32, seed=1234)).eval()
    with torch.no_grad():
        y32 = ref(x[1:2].contiguous())
    d = (y[1:2] - y1).float()
    floor = (y1.float() - y32).norm() / y32.norm()
    rel = d.norm() / y1.float().norm()
    print("batch-independence: rel L2 %.3e, max %.3e; bf16-vs-fp32 floor rel L2 %.3e" % (float(rel), float(d.abs().max()), float(floor)))
    assert float(rel) < 0.5 * float(floor)
    assert float((y[1:2].float() - y32).norm() / y32.norm()) < 1.5 * float(floor)


# ------------------------------------------------------------------ VGG19 perceptual branch (bf16 tensor-core path)
def _feat_check(d, feat):
    """feat against the golden summary (head values, norm, sum) written by oracle/make_golden.py:_store_tensor."""
    f = feat.detach().double().cpu().reshape(-1)
    head = torch.from_numpy(d["feat@head"]).double()
    assert abs(float(f.norm()) - float(d["feat@norm"])) < 2e-2 * float(d["feat@norm"])
    assert abs(float(f.sum()) - float(d["feat@sum"])) < 2e-2 * abs(float(d["feat@sum"])) + 1e-3 * float(d["feat@norm"])
    assert float((f[: head.numel()] - head).abs().max()) < 3e-2 * max(float(head.abs().max()), float(f.abs().max()) * 0.1)


def test_vgg_features_match_golden(golden_dir):
    from mri_superresolution_b200.utils.perceptual import VGGFeatureExtractor
    d = _golden(golden_dir, "perceptual")
    vggp = LO.make_vgg_params(seed=1235)
    ext = VGGFeatureExtractor(35, weights=vggp)
    pred = torch.from_numpy(d["pred"]).to(_dev())
    feat = ext(pred)
    ref = LO.vgg_features(vggp, torch.from_numpy(d["pred"]))
    assert feat.shape == ref.shape
    # bf16 operands through 16 convolutions: relative L2 of the feature tensor
    e = float((feat.cpu().double() - ref.double()).norm() / ref.double().norm())
    assert e < 3e-2, e
    _feat_check(d, feat)


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_combined_loss_with_perceptual_matches_golden(golden_dir, tag):
    """CombinedLoss with the perceptual term (utils/losses.py:229-236) against the reference run recorded in
    tests/golden/perceptual.npz (same seeded VGG19 weights): loss within 1e-2 relative (bf16 mode tolerance), gradient
    direction and size against the reference autograd gradient."""
    from mri_superresolution_b200.utils.losses import CombinedLoss
    d = _golden(golden_dir, "perceptual")
    sw, pw, typ = [float(v) for v in d[f"{tag}:cfg"]]
    crit = CombinedLoss(ssim_weight=sw, perceptual_weight=pw, perceptual_loss_type="l1" if typ == 1 else "mse",
                        vgg_weights=LO.make_vgg_params(seed=1235))
    p = torch.from_numpy(d["pred"]).to(_dev()).requires_grad_(True)
    tgt = torch.from_numpy(d["target"]).to(_dev())
    loss = crit(p, tgt)
    loss.backward()
    assert abs(loss.item() - float(d[f"{tag}:loss"])) < 1e-2 * abs(float(d[f"{tag}:loss"]))
    g, r = p.grad.double().cpu().reshape(-1), torch.from_numpy(d[f"{tag}:grad"]).double().reshape(-1)
    cos = float((g * r).sum() / (g.norm() * r.norm()))
    # bf16 noise floor of this gradient: the same loss with bf16-rounded VGG operands (oracle restatement, fp32 arithmetic)
    vggp = LO.make_vgg_params(seed=1235)
    pq = torch.from_numpy(d["pred"]).clone().requires_grad_(True)
    tq = torch.from_numpy(d["target"])
    fq, ftq = LO.vgg_features_bf16_storage(vggp, pq), LO.vgg_features_bf16_storage(vggp, tq).detach()
    perc_q = (fq - ftq).abs().mean() if typ == 1 else ((fq - ftq) ** 2).mean()
    ((1 - sw - pw) * (pq - tq).abs().mean() + sw * (1 - LO.ssim(pq, tq).clamp(0, 1)) + pw * perc_q).backward()
    gq = pq.grad.double().reshape(-1)
    floor = float((gq * r).sum() / (gq.norm() * r.norm()))
    assert cos > min(0.97, floor - 0.02), (cos, floor)
    cos_q = float((g * gq).sum() / (g.norm() * gq.norm()))
    assert cos_q > 0.9, cos_q                      # same operand rounding: only accumulation order / tie routing differ
    assert abs(float(g.norm()) - float(r.norm())) < 0.1 * float(r.norm())


def test_perceptual_target_branch_has_no_gradient_and_l2_alias(golden_dir):
    from mri_superresolution_b200.utils.perceptual import PerceptualLoss
    d = _golden(golden_dir, "perceptual")
    vggp = LO.make_vgg_params(seed=1235)
    pl2, pmse = PerceptualLoss(35, "l2", vgg_weights=vggp), PerceptualLoss(35, "mse", vgg_weights=vggp)
    p = torch.from_numpy(d["pred"]).to(_dev()).requires_grad_(True)
    t = torch.from_numpy(d["target"]).to(_dev()).requires_grad_(True)
    a = pl2(p, t)
    a.backward()
    assert t.grad is None and p.grad is not None                  # utils/losses.py:146-147
    assert abs(a.item() - pmse(p.detach(), t.detach()).item()) < 1e-6 * abs(a.item())   # :130 'l2' == 'mse'
    assert pl2(p.detach(), p.detach()).item() == 0.0


@pytest.mark.gpu
def test_vgg_fp32_parity_mode_features_and_loss(golden_dir):
    """precision='fp32' of the perceptual term (mrisr_vgg_*_fp32, CUDA cores): the north star's fp32-mode bound, 1e-4, for L6 /
    L7 -- features against the float32 restatement, then CombinedLoss values against the reference run recorded in
    tests/golden/perceptual.npz and the gradient against the reference autograd gradient."""
    from mri_superresolution_b200.utils.losses import CombinedLoss
    from mri_superresolution_b200.utils.perceptual import VGGFeatureExtractor
    d = _golden(golden_dir, "perceptual")
    vggp = LO.make_vgg_params(seed=1235)
    ext = VGGFeatureExtractor(35, weights=vggp, precision="fp32")
    feat = ext(torch.from_numpy(d["pred"]).to(_dev()))
    ref = LO.vgg_features(vggp, torch.from_numpy(d["pred"]))
    e = float((feat.cpu().double() - ref.double()).norm() / ref.double().norm())
    assert e < 1e-4, e
    for tag in ("a", "b", "c"):
        sw, pw, typ = [float(v) for v in d[f"{tag}:cfg"]]
        crit = CombinedLoss(ssim_weight=sw, perceptual_weight=pw, perceptual_loss_type="l1" if typ == 1 else "mse",
                            vgg_weights=vggp, precision="fp32")
        p = torch.from_numpy(d["pred"]).to(_dev()).requires_grad_(True)
        loss = crit(p, torch.from_numpy(d["target"]).to(_dev()))
        loss.backward()
        assert abs(loss.item() - float(d[f"{tag}:loss"])) < 1e-4 * abs(float(d[f"{tag}:loss"])), tag
        g, r = p.grad.double().cpu().reshape(-1), torch.from_numpy(d[f"{tag}:grad"]).double().reshape(-1)
        # fp32 accumulation order differs from cuDNN/MKL's; a ReLU or pooling tie may route differently at a handful of pixels
        assert float((g - r).norm() / r.norm()) < 2e-3, (tag, float((g - r).norm() / r.norm()))
        assert float((g * r).sum() / (g.norm() * r.norm())) > 0.99999


@pytest.mark.gpu
def test_vgg_fp32_mode_agrees_with_bf16_mode_and_rejects_bad_precision(golden_dir):
    from mri_superresolution_b200.utils.perceptual import PerceptualLoss
    d = _golden(golden_dir, "perceptual")
    vggp = LO.make_vgg_params(seed=1235)
    p, t = torch.from_numpy(d["pred"]).to(_dev()), torch.from_numpy(d["target"]).to(_dev())
    a = PerceptualLoss(35, "l1", vgg_weights=vggp, precision="fp32")(p, t).item()
    b = PerceptualLoss(35, "l1", vgg_weights=vggp, precision="bf16")(p, t).item()
    assert abs(a - b) < 2e-2 * abs(a)
    # twice through the same workspace (zero-border invariant survives a call), and a shallower prefix with a pooling at the end of a block
    crit = PerceptualLoss(8, "mse", vgg_weights=vggp[:4], precision="fp32")
    assert crit(p, t).item() == crit(p, t).item()
    with pytest.raises(ValueError):
        PerceptualLoss(35, "l1", vgg_weights=vggp, precision="fp16")


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(32, 1, 8, 8), (32, 3, 16, 8), (16, 2, 8, 24), (64, 1, 8, 8)])
def test_smallest_slices_forward_and_train_step(shape):
    """8-pixel edges: one pixel at the deepest level (every layer falls back to the kernels without tile-size floors).  fp32
    mode against the oracle at 1e-4; bf16 mode runs a whole train step and stays finite."""
    from mri_superresolution_b200.models.unet_model import UNetSuperRes
    from mri_superresolution_b200.utils.losses import CombinedLoss
    f, n, h, w = shape
    params = UO.init_params(f, seed=77)
    x, t = make_pairs(n, h, w, seed=78)
    x, t = x.to(_dev()), t.to(_dev())
    m = UNetSuperRes(base_filters=f, precision="fp32").to(_dev())
    m.load_state_dict(params)
    m.train()
    y = m(x)
    ref = UO.unet_forward(params, x.cpu())
    assert float((y.detach().cpu() - ref).abs().max()) < 1e-4 * max(1.0, float(ref.abs().max()))
    CombinedLoss(ssim_weight=0.4)(y, t).backward()
    assert all(torch.isfinite(p.grad).all() for p in m.parameters())
    mb = UNetSuperRes(base_filters=f).to(_dev())
    mb.load_state_dict(params)
    mb.train()
    yb = mb(x)
    CombinedLoss(ssim_weight=0.4)(yb, t).backward()
    assert torch.isfinite(yb).all() and all(torch.isfinite(p.grad).all() for p in mb.parameters())
    assert float((yb.detach().cpu() - ref).norm() / ref.norm()) < 0.1


@pytest.mark.gpu
def test_device_prefetcher_order_and_reuse():
    """DevicePrefetcher hands out every host batch once, in order, while the next one is being staged."""
    from mri_superresolution_b200.utils.prefetch import DevicePrefetcher
    dev = torch.device("cuda:0")
    host = [(torch.full((4, 1, 8, 8), float(i)).pin_memory(), torch.full((4, 1, 16, 16), float(-i)).pin_memory()) for i in range(7)]
    seen = []
    acc = torch.zeros((), device=dev)
    for x, t in DevicePrefetcher(host, dev):
        assert x.is_cuda and t.is_cuda
        # a long-running consumer on the current stream: the slot must not be overwritten under it
        for _ in range(50):
            acc = acc + (x.sum() + t.sum()) * 0.0
        seen.append((float(x.mean()), float(t.mean())))
    assert seen == [(float(i), float(-i)) for i in range(7)]
    with pytest.raises(ValueError):
        DevicePrefetcher(host, dev, slots=1)


# ------------------------------------------------------------------ drivers
def test_train_driver_two_epochs_match_the_oracle_loop(tmp_path, capsys):
    """drivers.train (graph replay, device-side running sums, prefetched CPU batches) against the reference loop restated over
    the oracle (oracle/train_oracle.py = scripts/train.py:296-323) on the same batches: per-epoch train loss / SSIM, the
    validation numbers, the JSON-lines schema and the checkpoint files."""
    import json
    from mri_superresolution_b200.drivers import TrainConfig, find_best_checkpoint, load_model, train
    f = 16
    params = UO.randomize_affine(UO.init_params(f, seed=21))
    tr = [make_pairs(2, 32, 32, seed=300 + i) for i in range(3)]
    va = [make_pairs(2, 32, 32, seed=400 + i) for i in range(2)]
    cfg = TrainConfig(epochs=2, learning_rate=2e-4, weight_decay=1e-5, ssim_weight=0.4, base_filters=f, patience=5,
                      checkpoint_dir=str(tmp_path), precision="fp32", batch_size=2)
    res = train(cfg, lambda: tr, lambda: va, device=_dev(), model=_model(f, "fp32", params))
    oracle = OracleTrainer(params, lr=2e-4, weight_decay=1e-5, ssim_weight=0.4)
    for ep in range(2):
        ls, ss = zip(*[oracle.step(x, t) for x, t in tr])
        with torch.no_grad():
            vl = [float(LO.combined_loss(UO.unet_forward(oracle.params, x), t, ssim_weight=0.4)) for x, t in va]
            vs = [float(LO.ssim(UO.unet_forward(oracle.params, x), t)) for x, t in va]
        h = res["history"][ep]
        # fp32 mode; Adam's m / sqrt(v) amplifies last-bit gradient differences where a gradient is near zero, hence 1e-3
        assert abs(h["train_loss"] - sum(ls) / 3) < 1e-3 * (sum(ls) / 3), (ep, h, ls)
        assert abs(h["train_ssim"] - sum(ss) / 3) < 1e-3, (ep, h, ss)
        assert abs(h["val_loss"] - sum(vl) / 2) < 1e-3 * (sum(vl) / 2), (ep, h, vl)
        assert abs(h["val_ssim"] - sum(vs) / 2) < 1e-3, (ep, h, vs)
    lines = [json.loads(l) for l in capsys.readouterr().out.strip().splitlines() if l.startswith("{")]
    kinds = [l["type"] for l in lines]
    assert kinds[0] == "params" and kinds.count("epoch_summary") == 2 and "batch_update" in kinds
    summ = [l for l in lines if l["type"] == "epoch_summary"][0]
    assert set(summ) == {"type", "epoch", "total_epochs", "train_loss", "val_loss", "train_ssim", "val_ssim", "elapsed", "lr"}
    best = torch.load(res["best_path"], map_location="cpu", weights_only=False)
    final = torch.load(res["final_path"], map_location="cpu", weights_only=False)
    assert set(best) == {"epoch", "model_state_dict", "optimizer_state_dict", "scheduler_state_dict", "val_loss", "val_ssim"}
    assert set(final) == {"epoch", "model_state_dict", "optimizer_state_dict", "val_loss", "val_ssim"}
    assert list(final["model_state_dict"].keys()) == list(params.keys())
    torch.optim.Adam([torch.nn.Parameter(v.clone()) for v in final["model_state_dict"].values()]).load_state_dict(
        final["optimizer_state_dict"])                                     # stock Adam layout
    assert find_best_checkpoint(str(tmp_path), "unet") == res["best_path"]
    m = load_model("unet", res["final_path"], _dev(), base_filters=f, precision="fp32")
    assert not m.training
    for k, v in oracle.params.items():
        assert rel_inf(m.state_dict()[k], v) < 5e-3, k                     # six sign-like Adam updates of 2e-4 each


def test_super_resolve_shards_a_slice_stack_without_exchange():
    """BASELINE.json config 5 in miniature: ranks take contiguous slice blocks; the blocks concatenate to the single-process
    result, which equals the per-slice reference loop (scripts/infer.py:268-276) restated over the oracle."""
    from mri_superresolution_b200.drivers import super_resolve
    params = UO.randomize_affine(UO.init_params(16, seed=9))
    model = _model(16, "fp32", params).eval()
    vol, _ = make_pairs(10, 40, 40, seed=77)                               # 10 slices of a 40 x 40 volume (40 = 8 * 5)
    whole = super_resolve(model, vol, batch=3)
    parts = [super_resolve(model, vol, batch=2, rank=r, world=4, use_graph=(r % 2 == 0)) for r in range(4)]
    assert [p.shape[0] for p in parts] == [2, 3, 2, 3]
    assert rel_inf(torch.cat(parts), whole) < 1e-5                         # batch position only moves summation order
    from oracle.train_oracle import infer_step
    ref = infer_step(params, vol)
    assert rel_inf(whole, ref) < 1e-4
    with pytest.raises(ValueError):
        super_resolve(model, vol[:, 0])


@pytest.mark.parametrize("hw", [(320, 320), (512, 512)])
def test_config4_config5_input_sizes_forward_parity(hw):
    """BASELINE.json configs 4 (512 x 512 slices) and 5 (320 x 320 volume slices; 320 = 8 * 40 is not a multiple of the
    128-pixel MMA tile or the 256-pixel ring block) at base_filters = 32, one slice: the inference forward against the
    oracle run on the host in fp32.  fp32 mode within 1e-4; bf16 mode within the north_star tolerances (loss 1e-2 relative,
    SSIM 1e-3, PSNR 0.05 dB) and at the bf16 rounding floor in relative L2."""
    h, w = hw
    params = UO.init_params(32, seed=1234)
    x, tgt = make_pairs(1, h, w, seed=5)
    with torch.no_grad():
        ref = UO.unet_forward(params, x)
        y32 = _model(32, "fp32", params).eval()(x.to(_dev())).cpu()
        y16 = _model(32, "bf16", params).eval()(x.to(_dev())).cpu()
    assert y16.shape == (1, 1, 2 * h, 2 * w)
    assert rel_inf(y32, ref) < 1e-4
    lref = float(LO.combined_loss(ref, tgt, ssim_weight=0.4))
    l16 = float(LO.combined_loss(y16, tgt, ssim_weight=0.4))
    assert abs(l16 - lref) < 1e-2 * lref
    assert abs(float(LO.ssim(y16, tgt)) - float(LO.ssim(ref, tgt))) < 1e-3
    assert abs(LO.psnr(y16, tgt) - LO.psnr(ref, tgt)) < 0.05
    assert float((y16 - ref).norm() / ref.norm()) < 5e-2


# ------------------------------------------------------------------ round 2: full-size gradients, trained weights, VGG at 256^2
def test_full_size_train_step_gradients_match_oracle_autograd():
    """The shape BENCH times (f = 32, 256 x 256 -> 512 x 512; batch 2): every parameter gradient of the bf16 path against the
    oracle's autograd run on the same GPU in fp32 (TF32 off).  Yardstick = the bf16-storage restatement of the same algorithm
    (oracle.unet_forward_bf16_storage: identical rounding points, fp32 arithmetic): per tensor the CUDA path may be at most
    twice as far from the fp32 gradient as that restatement is (relative L2), and its norm must be within 3 % of the
    restatement's."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    from mri_superresolution_b200.utils.losses import CombinedLoss
    f, n, h, w = 32, 2, 256, 256
    params = UO.randomize_affine(UO.init_params(f, seed=1234))
    x, t = make_pairs(n, h, w, seed=31)
    x, t = x.to(_dev()), t.to(_dev())
    model = _model(f, "bf16", params)
    model.train()
    loss = CombinedLoss(ssim_weight=0.4)(model(x), t)
    loss.backward()
    res = {}
    for tag, fwd in (("fp32", UO.unet_forward), ("bf16q", UO.unet_forward_bf16_storage)):
        p = {k: v.to(_dev()).clone().requires_grad_(True) for k, v in params.items()}
        l = LO.combined_loss(fwd(p, x), t, ssim_weight=0.4)
        l.backward()
        res[tag] = ({k: v.grad.double() for k, v in p.items()}, float(l))
    (g32, l32), (gq, lq) = res["fp32"], res["bf16q"]
    assert abs(loss.item() - l32) < 1e-2 * l32
    assert abs(loss.item() - lq) < 2e-3 * lq
    gmax = max(float(v.norm()) for v in g32.values())
    worst = []
    for k, pn in model.named_parameters():
        g = pn.grad.double().view(g32[k].shape)
        r32, rq = g32[k], gq[k]
        if r32.numel() == 1:
            # alpha: a sum of (xb - xp) * g over every output pixel, cancellation leaves a value far below its terms
            floor = abs(float(rq) - float(r32))
            assert abs(float(g) - float(r32)) <= 2.0 * floor + 2e-3 * abs(float(r32)) + 1e-6 * gmax, (k, float(g), float(r32), float(rq))
            continue
        nr = float(r32.norm())
        e = float((g - r32).norm()) / max(nr, 1e-30)
        floor = float((rq - r32).norm()) / max(nr, 1e-30)
        eq = float((g - rq).norm()) / max(float(rq.norm()), 1e-30)
        worst.append((e / max(floor, 1e-9), k, e, floor, eq))
        assert e <= 2.0 * floor + 1e-3, (k, e, floor)
        if nr > 1e-4 * gmax:
            assert abs(float(g.norm()) / float(rq.norm()) - 1.0) < 0.03, (k, float(g.norm()), float(rq.norm()))
    worst.sort(reverse=True)
    print("full-size gradients: worst (e / floor, tensor, e vs fp32, floor, e vs bf16 restatement):", worst[:4])


def test_trained_weights_bf16_meets_the_literal_north_star_bounds(golden_dir):
    """tests/golden/trained_f16.npz: the REFERENCE model after 60 of its own training steps, evaluated by the reference on two
    held-out pairs.  On weights that are not random init the final sigmoid is not saturating and the bf16 path meets the
    north_star bounds as written: output within 1e-2 relative (norm-wise; the worst single pixel is reported and bounded by
    5e-2), loss within 1e-2 relative, SSIM within 1e-3, PSNR within 0.05 dB; fp32 mode within 1e-4 in the max norm."""
    from mri_superresolution_b200.utils.losses import CombinedLoss, SSIM
    d = _golden(golden_dir, "trained_f16")
    f, n, h, w = [int(v) for v in d["meta"]]
    params = {k[len("param:"):]: torch.from_numpy(v) for k, v in d.items() if k.startswith("param:")}
    x, t = make_pairs(n, h, w, seed=777)
    ref = torch.from_numpy(d["out"])
    crit, metric = CombinedLoss(ssim_weight=0.4), SSIM()
    for precision in ("fp32", "bf16"):
        model = _model(f, precision, params).eval()
        with torch.no_grad():
            y = model(x.to(_dev()))
            loss = float(crit(y, t.to(_dev())))
            s = float(metric(y, t.to(_dev())))
        e_inf, e_l2 = rel_inf(y, ref), rel_l2(y, ref)
        dpsnr = abs(LO.psnr(y.cpu(), t) - LO.psnr(ref, t))
        print(f"trained weights [{precision}]: out rel max {e_inf:.2e} l2 {e_l2:.2e}, loss {loss:.6f} vs {float(d['loss']):.6f}, "
              f"ssim {s:.5f} vs {float(d['ssim']):.5f}, dPSNR {dpsnr:.4f} dB")
        if precision == "fp32":
            assert e_inf < 1e-4 and abs(loss - float(d["loss"])) < 1e-4 * float(d["loss"]) and abs(s - float(d["ssim"])) < 1e-4
        else:
            assert e_l2 < 1e-2 and e_inf < 5e-2, (e_l2, e_inf)
            assert abs(loss - float(d["loss"])) < 1e-2 * float(d["loss"])
            assert abs(s - float(d["ssim"])) < 1e-3
        assert dpsnr < 0.05


def test_vgg_perceptual_matches_reference_at_256(golden_dir):
    """tests/golden/perceptual_256.npz (the reference run at HR 256 x 256: 16 x 16 images at the deepest layers, multi-tile
    launches of the 256- and 512-channel kernels, two-block Cout = 512 epilogue, ReLU-mask dgrad at every level)."""
    from mri_superresolution_b200.utils.losses import CombinedLoss
    from mri_superresolution_b200.utils.perceptual import VGGFeatureExtractor
    d = _golden(golden_dir, "perceptual_256")
    _, hr = make_pairs(1, 128, 128, seed=98)
    g = torch.Generator("cpu").manual_seed(6)
    pred = (hr + 0.1 * torch.randn(hr.shape, generator=g)).clamp(0, 1)
    np.testing.assert_allclose([float(pred.double().sum()), float(hr.double().sum())], d["checksum"], rtol=1e-12)
    vggp = LO.make_vgg_params(seed=1235)
    feat = VGGFeatureExtractor(35, weights=vggp)(pred.to(_dev()))
    assert tuple(feat.shape) == tuple(int(v) for v in d["feat_shape"])
    _feat_check(d, feat)
    ref_feat = LO.vgg_features(vggp, pred)
    assert rel_l2(feat, ref_feat) < 3e-2, rel_l2(feat, ref_feat)
    crit = CombinedLoss(ssim_weight=0.4, perceptual_weight=0.1, perceptual_loss_type="l1", vgg_weights=vggp)
    p = pred.to(_dev()).requires_grad_(True)
    loss = crit(p, hr.to(_dev()))
    loss.backward()
    assert abs(loss.item() - float(d["loss"])) < 1e-2 * float(d["loss"])
    gr = p.grad.double().cpu()
    crop, rc = gr[0, 0, 96:160, 96:160], torch.from_numpy(d["grad_crop"]).double()
    cos = float((crop * rc).sum() / (crop.norm() * rc.norm()))
    assert cos > 0.97, cos
    assert abs(float(gr.norm()) - float(d["grad@norm"])) < 0.05 * float(d["grad@norm"])


# ------------------------------------------------------------------ drivers under the conditions ADVICE (round 1) named
@pytest.mark.parametrize("case", ["ragged", "perceptual"])
def test_train_driver_graph_survives_other_shapes_and_lr_changes(tmp_path, case):
    """drivers.train with the CUDA-graph step against the same loop run eagerly (use_graph=False), three epochs:
      ragged      the last batch of every epoch is smaller (the reference's DataLoader has no drop_last): it runs eagerly with
                  its own workspace, which used to evict the workspace the recorded graph points into
      perceptual  perceptual_weight > 0 with validation under no_grad: the forward-only VGG workspace used to evict the
                  training one of the recorded graph
    Same kernels in both legs, so losses / metrics agree to accumulation order; a replay on freed memory does not."""
    from mri_superresolution_b200.drivers import TrainConfig, train
    f = 16
    params = UO.randomize_affine(UO.init_params(f, seed=21))
    sizes = [2, 2, 1] if case == "ragged" else [2, 2, 2]
    tr = [make_pairs(b, 32, 32, seed=300 + i) for i, b in enumerate(sizes)]
    va = [make_pairs(3 if case == "perceptual" else 2, 32, 32, seed=400 + i) for i in range(2)]
    hist = []
    for use_graph in (True, False):
        cfg = TrainConfig(epochs=3, learning_rate=2e-4, weight_decay=1e-5, ssim_weight=0.4,
                          perceptual_weight=0.1 if case == "perceptual" else 0.0, base_filters=f,
                          patience=6, checkpoint_dir=str(tmp_path / f"g{int(use_graph)}"),
                          precision="bf16", batch_size=2, use_graph=use_graph,
                          vgg_weights=LO.make_vgg_params(seed=1235) if case == "perceptual" else None)
        res = train(cfg, lambda: tr, lambda: va, device=_dev(), model=_model(f, "bf16", params))
        hist.append(res["history"])
    assert len(hist[0]) == len(hist[1]) >= 2
    for hg, he in zip(*hist):
        for key in ("train_loss", "val_loss", "train_ssim", "val_ssim"):
            assert abs(hg[key] - he[key]) <= 2e-3 * max(abs(he[key]), 1e-3), (case, key, hg, he)
        assert hg["lr"] == he["lr"]


def test_graphed_step_follows_a_scheduler_between_replays():
    """ReduceLROnPlateau (scripts/train.py:189,397) changes param_groups[0]['lr'] between steps: the recorded step reads the
    learning rate from device memory (FusedAdam.sync_lr), so replays after the change must equal eager steps after it."""
    from mri_superresolution_b200.optim import FusedAdam
    from mri_superresolution_b200.utils.graph import GraphedTrainStep
    from mri_superresolution_b200.utils.losses import CombinedLoss
    params = UO.randomize_affine(UO.init_params(16, seed=5))
    batches = [tuple(v.to(_dev()) for v in make_pairs(2, 32, 32, seed=70 + i)) for i in range(5)]
    finals = []
    for use_graph in (False, True):
        model = _model(16, "bf16", params)
        crit = CombinedLoss(ssim_weight=0.4)
        opt = FusedAdam(model, lr=1e-3, weight_decay=1e-5, capturable=use_graph)
        sched = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode="min", factor=0.5, patience=0)
        stepper = GraphedTrainStep(model, crit, opt, *batches[0]) if use_graph else None
        for i, (x, t) in enumerate(batches):
            if use_graph:
                stepper(x, t)
            else:
                opt.zero_grad(set_to_none=True)
                crit(model(x), t).backward()
                opt.step()
            sched.step(1.0 + i)                                # a metric that only gets worse: lr halves after every step
        assert opt.param_groups[0]["lr"] < 1e-3 / 8
        finals.append(model.flat_parameters().clone())
        if stepper is not None:
            stepper.close()
            with pytest.raises(RuntimeError):
                stepper(*batches[0])
    assert rel_inf(finals[1], finals[0]) < 1e-5


def test_bench_two_gpu_run_exits_cleanly():
    """bench.py --gpus 2 must print its JSON line AND exit (round 1: the process hung in NCCL teardown while the recorded
    step was still alive).  Skipped on a single-GPU box."""
    import json
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29571", os.path.join(root, "bench.py"), "--gpus", "2", "--steps", "3", "--warmup", "3", "--batch", "8",
           "--no-cpu-baseline"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=170, cwd=root)
    assert res.returncode == 0, res.stderr[-3000:]
    line = json.loads([l for l in res.stdout.splitlines() if l.startswith("{")][-1])
    assert line["n_gpus"] == 2 and line["value"] > 0
