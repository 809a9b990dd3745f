"""Per-layer parity of the CUDA path, forward AND backward, without the 21-layer amplification: every internal tensor the
library keeps (read through mrisr_unet_debug_tensor) is compared with the reference operation of THAT layer
(models/unet_model.py:29-37,52,69-75,93,101-113,150-158,166-173,206-211 and its autograd) applied in fp32 to the library's own
input tensors of the layer.  A wrong scale, a wrong tap, a wrong routing in any single kernel (conv / dgrad / wgrad engines,
gn_apply, max-pool, bilinear x2 and its adjoint, head_blend, final_head, every gn_bwd gradient-source kind, d(alpha)) shows up
at its own layer at the resolution of one bf16 rounding instead of being averaged into an end-to-end cosine.

Tolerances (north_star: forward activations within 1e-2 relative in bf16 mode, 1e-4 in fp32 mode; 'relative' = max-norm error
over the tensor's max magnitude):
  bf16 mode  forward tensors <= 1e-2 (north_star); backward tensors stored in bf16 <= 8e-3 (half an ulp of bf16 is 3.9e-3 of the
             element, plus GroupNorm statistics taken from the stored instead of the fp32 accumulator values); fp32 results
             (parameter gradients) <= 2e-3
  fp32 mode  forward <= 1e-4, backward <= 2e-3 (the tolerance of tests/test_gpu_parity.py for gradients)
LeakyReLU's derivative jumps at z = 0: elements whose pre-activation is within 1e-2 of zero may take the other slope when the
GroupNorm statistics differ in the last bits; they are compared against the looser bound |upstream| * |gamma| * rstd.
"""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from oracle import unet_oracle as UO
from oracle.synth import make_pairs

KINDS = {"R": 0, "A": 1, "P": 2, "Q": 3, "dR": 4, "G": 5, "dQ": 6, "UB": 7, "XM": 8, "DY3": 9}
# conv index -> (weight, gamma, beta) positions in the reference's parameter order (csrc/unet.cu conv_table)
TAB = [(1, 2, 3), (4, 5, 6), (7, 8, 9), (10, 11, 12), (13, 14, 15), (16, 17, 18), (19, 20, 21), (22, 23, 24),
       (25, 26, 27), (28, 29, 30), (31, 32, 33), (34, 35, 36), (37, 38, 39), (40, 41, 42), (43, 44, 45),
       (46, 47, 48), (49, 50, 51), (52, 53, 54), (55, 57, 58), (59, 60, 61)]
UP1X1 = (8, 11, 14)
SKIP = {9: 5, 12: 3, 15: 1}


def _dev():
    return torch.device("cuda:0")


def _tap(model, n, h, w, kind, layer=0):
    from mri_superresolution_b200 import _native
    ws = model._workspace(n, h, w, True)
    prec = _native.BF16 if model.precision == "bf16" else _native.FP32
    dims = (ctypes.c_int * 4)()
    L = _native.lib()
    _native.check(L.mrisr_unet_debug_tensor(_native.ptr(ws), n, h, w, model.base_filters, prec, KINDS[kind], layer, None,
                                            ctypes.cast(dims, ctypes.c_void_p), _native.stream_ptr()), "debug_tensor")
    out = torch.empty(tuple(dims), dtype=torch.float32, device=ws.device)
    _native.check(L.mrisr_unet_debug_tensor(_native.ptr(ws), n, h, w, model.base_filters, prec, KINDS[kind], layer,
                                            _native.ptr(out), None, _native.stream_ptr()), "debug_tensor")
    return out


def _gn_lrelu(r, gamma, beta):
    return F.leaky_relu(F.group_norm(r, 8, gamma, beta, 1e-5), 0.2)


def _preact(r, gamma, beta):
    return F.group_norm(r, 8, gamma, beta, 1e-5)


def _up(t):
    return F.interpolate(t, scale_factor=2, mode="bilinear", align_corners=True)


class Checker:
    def __init__(self):
        self.rows = []
        self.bad = []

    def check(self, name, got, want, tol, near_zero=None, loose=None, slack=None):
        got, want = got.double(), want.double()
        scale = max(float(want.abs().max()), 1e-30)
        diff = (got - want).abs()
        if slack is not None:                       # per-element absolute allowance (kink elements inside a parameter sum)
            diff = (diff - slack.double()).clamp_min(0.0)
        if near_zero is not None and bool(near_zero.any()):
            # elements on LeakyReLU's kink: the looser bound; everything else: the tight one
            worst_kink = float((diff[near_zero] / loose.double()[near_zero].clamp_min(1e-30)).max())
            if worst_kink > 1.0 + 1e-3:
                self.bad.append((name + " (kink elements)", worst_kink, 1.0))
            diff = diff.masked_fill(near_zero, 0.0)
        e = float(diff.max()) / scale
        self.rows.append((name, e, tol, scale))
        if not e <= tol:
            self.bad.append((name, e, tol))

    def report(self):
        for name, e, tol, scale in self.rows:
            print(f"{name:14s} rel err {e:.3e} (tol {tol:.0e}, scale {scale:.3e})")
        assert not self.bad, self.bad


def _run_layers(precision, f, n, h, w):
    from mri_superresolution_b200.models.unet_model import UNetSuperRes
    from mri_superresolution_b200.utils.losses import CombinedLoss
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = _dev()
    bf = precision == "bf16"
    q = (lambda t: t.to(torch.bfloat16).to(torch.float32)) if bf else (lambda t: t)
    tol_f = 1e-2 if bf else 1e-4
    tol_b = 8e-3 if bf else 2e-3
    tol_p = 2e-3
    params = UO.randomize_affine(UO.init_params(f, seed=1234))
    names = [k for k, _ in UO.param_specs(f)]
    p = {k: v.to(dev) for k, v in params.items()}
    Wt = lambda l: q(p[names[TAB[l][0]]])
    gam = lambda l: p[names[TAB[l][1]]]
    bet = lambda l: p[names[TAB[l][2]]]
    x, tgt = make_pairs(n, h, w, seed=4321)
    x, tgt = x.to(dev), tgt.to(dev)
    model = UNetSuperRes(base_filters=f, precision=precision).to(dev)
    model.load_state_dict(params)
    model.train()
    y = model(x)
    y.retain_grad()
    CombinedLoss(ssim_weight=0.4)(y, tgt).backward()
    dy = y.grad.detach()
    y = y.detach()
    T = lambda kind, l=0: _tap(model, n, h, w, kind, l)
    R = {l: T("R", l) for l in range(20)}
    A = {l: T("A", l) for l in range(17)}
    Pq = {l: T("P", l) for l in (1, 3, 5)}
    Q = {l: T("Q", l) for l in UP1X1}
    UB, XM = T("UB"), T("XM")
    dR = {l: T("dR", l) for l in range(20)}
    G = {l: T("G", l) for l in range(1, 20)}
    dQ = {l: T("dQ", l) for l in UP1X1}
    DY3 = T("DY3")
    grads = {k: v.grad.detach() for k, v in model.named_parameters()}
    ck = Checker()

    # ------------------------------------------------------------------ forward, layer by layer
    def conv_in(l):
        if l == 0:
            return q(x)
        if l in (2, 4, 6):
            return Pq[l - 1]
        if l in SKIP:
            return torch.cat([A[SKIP[l]], A[l - 1]], 1)
        if l == 17:
            return UB
        if l == 18:
            return A[16]
        if l == 19:
            return XM
        return A[l - 1]

    for l in range(20):
        if l in UP1X1:
            ck.check(f"Q{l}", Q[l], F.conv2d(A[l - 1], Wt(l)), tol_f)
            ck.check(f"R{l}", R[l], _up(Q[l]), tol_f)
        else:
            bias = p[names[56]] if l == 18 else None
            ck.check(f"R{l}", R[l], F.conv2d(conv_in(l), Wt(l), bias, padding=1), tol_f)
        if l < 17:
            ck.check(f"A{l}", A[l], _gn_lrelu(R[l], gam(l), bet(l)), tol_f)
    for l in (1, 3, 5):
        ck.check(f"P{l}", Pq[l], F.max_pool2d(A[l], 2), 0.0 if bf else 1e-7)       # a selection: exact
    ck.check("UB", UB, _up(A[16]), tol_f)
    al = torch.sigmoid(p["alpha"])
    xb = _gn_lrelu(R[17], gam(17), bet(17))
    xp = _gn_lrelu(F.pixel_shuffle(R[18], 2), gam(18), bet(18))
    ck.check("XM", XM, al * xb + (1 - al) * xp, tol_f)
    a19 = _gn_lrelu(R[19], gam(19), bet(19))
    zfin = F.conv2d(a19, p[names[62]], p[names[63]])
    ck.check("y", y, torch.sigmoid(zfin), tol_f)

    # ------------------------------------------------------------------ backward, kernel by kernel
    def gn_bwd(l, upstream, shuffle=False):
        r = R[l].clone().requires_grad_(True)
        g_, b_ = gam(l).clone().requires_grad_(True), bet(l).clone().requires_grad_(True)
        src = F.pixel_shuffle(r, 2) if shuffle else r
        z = _preact(src, g_, b_)
        F.leaky_relu(z, 0.2).backward(upstream)
        near = z.detach().abs() < 1e-2
        # an element on the kink may take the other slope: |d(dr)| <= 0.8 |upstream| |gamma| rstd (+ the group-mean terms, tiny)
        var = src.detach().reshape(src.shape[0], 8, -1).var(2, unbiased=False)
        rstd = torch.rsqrt(var + 1e-5).repeat_interleave(src.shape[1] // 8, 1)[:, :, None, None]
        loose = 0.8 * upstream.abs() * g_.detach().abs().view(1, -1, 1, 1) * rstd + tol_b * float(r.grad.abs().max())
        # the same elements inside the per-channel parameter sums: each may move d(beta) by 0.8 |upstream| and d(gamma) by that
        # times |xhat| (deep layers of small cases have only a few hundred elements per channel, so it does not average out)
        kink_up = 0.8 * (upstream.abs() * near)
        xhat = (z.detach() - b_.detach().view(1, -1, 1, 1)) / g_.detach().view(1, -1, 1, 1)
        slack_b, slack_g = kink_up.sum((0, 2, 3)), (kink_up * xhat.abs()).sum((0, 2, 3))
        if shuffle:
            near, loose = F.pixel_unshuffle(near.float(), 2) > 0, F.pixel_unshuffle(loose, 2)
        return r.grad, g_.grad, b_.grad, near, loose, slack_g, slack_b

    def pool_route(l, gpool):
        a = A[l].clone().requires_grad_(True)
        F.max_pool2d(a, 2).backward(gpool)
        return a.grad

    def up_adjoint(g_hi, lo_shape):
        t = torch.zeros(lo_shape, device=dev, requires_grad=True)
        _up(t).backward(g_hi)
        return t.grad

    def check_block(l, upstream, shuffle=False):
        dr, dg, db, near, loose, slack_g, slack_b = gn_bwd(l, upstream, shuffle)
        ck.check(f"dR{l}", dR[l], dr, tol_b, near, loose)
        ck.check(f"dgamma{l}", grads[names[TAB[l][1]]], dg, 5 * tol_p, slack=slack_g)
        ck.check(f"dbeta{l}", grads[names[TAB[l][2]]], db, 5 * tol_p, slack=slack_b)

    def check_conv_bwd(l, src, dy_native, want_g=True, ks=3):
        pad = 1 if ks == 3 else 0
        wgt = Wt(l)
        ck.check(f"dW{l}", grads[names[TAB[l][0]]], torch.nn.grad.conv2d_weight(src, wgt.shape, dy_native, padding=pad), tol_p)
        if want_g:
            ck.check(f"G{l}", G[l], torch.nn.grad.conv2d_input(src.shape, wgt, dy_native, padding=pad), tol_b)

    # final_conv.1..3 + sigmoid (gn_bwd kind SRC_OUTER) ------------------------------------------------------
    dl = dy * y * (1 - y)
    w1 = p[names[62]]
    check_block(19, dl * w1.view(1, -1, 1, 1))
    ck.check("dW_1x1", grads[names[62]].flatten(), (dl * a19).sum((0, 2, 3)), tol_p)
    ck.check("db_1x1", grads[names[63]].flatten(), dl.sum().view(1), tol_p)
    check_conv_bwd(19, XM, dR[19])
    # blend + pixel shuffle (kinds SRC_BLEND_BIL, SRC_BLEND_PS) and d(alpha) --------------------------------
    check_block(17, al * G[19])
    check_block(18, (1 - al) * G[19], shuffle=True)
    ck.check("dalpha", grads["alpha"].view(1), ((xb - xp) * G[19]).sum().view(1) * al * (1 - al), 5 * tol_p)
    ck.check("dbias18", grads[names[56]], dR[18].sum((0, 2, 3)), tol_p)
    check_conv_bwd(17, UB, dR[17])
    check_conv_bwd(18, A[16], dR[18])
    ck.check("DY3", DY3, up_adjoint(G[17], A[16].shape) + G[18], tol_b)
    # decoder / encoder chain (kinds SRC_DIRECT, SRC_DIRECT_POOL; bilinear adjoint; concat split) -----------
    check_block(16, DY3)
    check_conv_bwd(16, A[15], dR[16])
    up_g = {15: G[16], 12: G[13], 9: G[10]}
    for l in (15, 14, 13, 12, 11, 10, 9, 8, 7, 6, 5, 4, 3, 2, 1, 0):
        if l in SKIP:
            check_block(l, up_g[l])
            check_conv_bwd(l, torch.cat([A[SKIP[l]], A[l - 1]], 1), dR[l])
        elif l in UP1X1:
            cat = G[l + 1]
            check_block(l, cat[:, cat.shape[1] // 2:].contiguous())
            ck.check(f"dQ{l}", dQ[l], up_adjoint(dR[l], Q[l].shape), tol_b)
            check_conv_bwd(l, A[l - 1], dQ[l], ks=1)
        elif l in (1, 3, 5):
            cat = G[{1: 15, 3: 12, 5: 9}[l]]
            check_block(l, cat[:, :cat.shape[1] // 2].contiguous() + pool_route(l, G[l + 1]))
            check_conv_bwd(l, A[l - 1], dR[l])
        else:
            check_block(l, G[l + 1])
            check_conv_bwd(l, conv_in(l), dR[l], want_g=(l != 0))
    ck.report()


def test_every_layer_forward_and_backward_bf16_tensor_core_shapes():
    """f = 32, 128 x 128, batch 2: the tcgen05 engines (conv_tc / conv_ring / wgrad_*) and the cluster-fused gn_bwd with more
    than one cluster rank are the kernels that run (checked below through mrisr_unet_layer_engine)."""
    from mri_superresolution_b200 import _native
    engines = [_native.lib().mrisr_unet_layer_engine(32, l, 128, 128) for l in range(20)]
    assert engines.count(0) <= 1 and 1 in engines and 2 in engines, engines       # only the 1-channel first conv on CUDA cores
    _run_layers("bf16", 32, 2, 128, 128)


def test_every_layer_forward_and_backward_bf16_ragged_shape():
    """f = 16, 24 x 40, batch 3: widths that are not multiples of any tile, odd cluster splits."""
    _run_layers("bf16", 16, 3, 24, 40)


def test_every_layer_forward_and_backward_bf16_base_filters_64():
    """f = 64 is the default of the reference's inference command (scripts/infer.py:473) and of its load_model (:50): 512
    channels at the bottleneck (conv_tc in two column blocks with a separate statistics pass, CUDA-core weight gradient:
    unet_plan.cuh layer_engine), NHWC at every level.  128 x 128 so that the 16 x 16 bottleneck is on the tensor-core kernels."""
    _run_layers("bf16", 64, 2, 128, 128)


def test_every_layer_forward_and_backward_fp32_mode():
    _run_layers("fp32", 16, 2, 32, 48)
