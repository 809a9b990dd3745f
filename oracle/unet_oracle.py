"""Functional fp32 restatement of the reference U-Net (TEST INFRASTRUCTURE, see oracle/__init__.py).

Follows /root/reference/models/unet_model.py; parameters are addressed by the reference's own
``state_dict`` keys so that a reference checkpoint drives the oracle unchanged.

Pinned against the reference modules by ``oracle/make_golden.py`` -> ``tests/golden/*.npz``.
"""
from __future__ import annotations

import math
from typing import Dict, List, Tuple

import torch
import torch.nn.functional as F

GN_GROUPS = 8        # models/unet_model.py:30,36,73,103,153,169
GN_EPS = 1e-5        # torch.nn.GroupNorm default used by the reference
LRELU_SLOPE = 0.2    # models/unet_model.py:31


def param_specs(base_filters: int = 32, in_channels: int = 1, out_channels: int = 1) -> List[Tuple[str, Tuple[int, ...]]]:
    """(name, shape) of every parameter in the reference's registration order.

    Order = nn.Module attribute registration order of UNetSuperRes.__init__
    (models/unet_model.py:129-175): inc, down1..3, up1..3, final_up_bilinear,
    final_up_pixelshuffle, alpha, final_conv.  ``state_dict()`` iterates parameters of the
    root module first (alpha) and then the children; ``parameters()`` does the same.
    """
    f = base_filters
    specs: List[Tuple[str, Tuple[int, ...]]] = [("alpha", ())]

    def double_conv(prefix: str, cin: int, cout: int) -> None:
        specs.append((f"{prefix}.double_conv.0.weight", (cout, cin, 3, 3)))
        specs.append((f"{prefix}.double_conv.1.weight", (cout,)))
        specs.append((f"{prefix}.double_conv.1.bias", (cout,)))
        specs.append((f"{prefix}.double_conv.3.weight", (cout, cout, 3, 3)))
        specs.append((f"{prefix}.double_conv.4.weight", (cout,)))
        specs.append((f"{prefix}.double_conv.4.bias", (cout,)))

    double_conv("inc", in_channels, f)
    double_conv("down1.maxpool_conv.1", f, 2 * f)
    double_conv("down2.maxpool_conv.1", 2 * f, 4 * f)
    double_conv("down3.maxpool_conv.1", 4 * f, 8 * f)
    for name, cup, cskip, cout in (("up1", 8 * f, 4 * f, 4 * f), ("up2", 4 * f, 2 * f, 2 * f), ("up3", 2 * f, f, f)):
        specs.append((f"{name}.up.1.weight", (cup // 2, cup, 1, 1)))
        specs.append((f"{name}.up.2.weight", (cup // 2,)))
        specs.append((f"{name}.up.2.bias", (cup // 2,)))
        double_conv(f"{name}.conv", cskip + cup // 2, cout)
    specs.append(("final_up_bilinear.1.weight", (f // 2, f, 3, 3)))
    specs.append(("final_up_bilinear.2.weight", (f // 2,)))
    specs.append(("final_up_bilinear.2.bias", (f // 2,)))
    specs.append(("final_up_pixelshuffle.conv.weight", (2 * f, f, 3, 3)))
    specs.append(("final_up_pixelshuffle.conv.bias", (2 * f,)))
    specs.append(("final_up_pixelshuffle.norm.weight", (f // 2,)))
    specs.append(("final_up_pixelshuffle.norm.bias", (f // 2,)))
    specs.append(("final_conv.0.weight", (f // 2, f // 2, 3, 3)))
    specs.append(("final_conv.1.weight", (f // 2,)))
    specs.append(("final_conv.1.bias", (f // 2,)))
    specs.append(("final_conv.3.weight", (out_channels, f // 2, 1, 1)))
    specs.append(("final_conv.3.bias", (out_channels,)))
    return specs


def init_params(base_filters: int = 32, seed: int = 1234, initial_alpha: float = 0.0) -> Dict[str, torch.Tensor]:
    """Reference initialisation (models/unet_model.py:177-187) restated functionally.

    Conv weights: kaiming_normal_(mode='fan_out', nonlinearity='leaky_relu') i.e.
    N(0, std) with std = sqrt(2 / (1 + 0.01**2)) / sqrt(Cout*kh*kw)  (torch's default
    leaky slope 0.01 when `a` is not passed); conv biases 0; GroupNorm weight 1 / bias 0;
    alpha = initial_alpha / 100 (:162-163).  NOTE: this does not reproduce the reference's
    RNG stream draw-for-draw (its modules consume extra draws for default inits and ICNR),
    parity tests therefore always copy the SAME tensors into both sides.
    """
    g = torch.Generator("cpu").manual_seed(seed)
    out: Dict[str, torch.Tensor] = {}
    gain = math.sqrt(2.0 / (1.0 + 0.01 ** 2))
    for name, shape in param_specs(base_filters):
        if name == "alpha":
            out[name] = torch.tensor(initial_alpha / 100.0, dtype=torch.float32)
        elif len(shape) == 4:
            fan_out = shape[0] * shape[2] * shape[3]
            out[name] = torch.randn(shape, generator=g, dtype=torch.float32) * (gain / math.sqrt(fan_out))
        elif name.endswith("conv.bias") or name == "final_conv.3.bias":
            out[name] = torch.zeros(shape, dtype=torch.float32)
        elif name.endswith(".weight"):
            out[name] = torch.ones(shape, dtype=torch.float32)
        else:
            out[name] = torch.zeros(shape, dtype=torch.float32)
    return out


def randomize_affine(params: Dict[str, torch.Tensor], seed: int = 4321) -> Dict[str, torch.Tensor]:
    """Perturb GroupNorm affine, biases and alpha so parity tests exercise them
    (the reference init leaves them at 1/0 where several gradient paths are degenerate)."""
    g = torch.Generator("cpu").manual_seed(seed)
    out = {}
    for k, v in params.items():
        if v.dim() == 4:
            out[k] = v.clone()
        elif k == "alpha":
            out[k] = torch.tensor(0.3, dtype=torch.float32)
        elif k.endswith(".weight"):
            out[k] = (1.0 + 0.2 * torch.randn(v.shape, generator=g)).to(torch.float32)
        else:
            out[k] = (0.1 * torch.randn(v.shape, generator=g)).to(torch.float32)
    return out


def _gn_lrelu(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    return F.leaky_relu(F.group_norm(x, GN_GROUPS, w, b, GN_EPS), LRELU_SLOPE)


def _double_conv(p: Dict[str, torch.Tensor], prefix: str, x: torch.Tensor, taps: Dict[str, torch.Tensor] | None) -> torch.Tensor:
    """DoubleConv.forward, models/unet_model.py:17-45 (dilation is always 1 in this net)."""
    w0 = p[f"{prefix}.double_conv.0.weight"]
    w3 = p[f"{prefix}.double_conv.3.weight"]
    y = _gn_lrelu(F.conv2d(x, w0, None, padding=1),
                  p[f"{prefix}.double_conv.1.weight"], p[f"{prefix}.double_conv.1.bias"])
    y = _gn_lrelu(F.conv2d(y, w3, None, padding=1),
                  p[f"{prefix}.double_conv.4.weight"], p[f"{prefix}.double_conv.4.bias"])
    if w0.shape[1] == w3.shape[0]:            # use_residual, :25,:43-44 (never true in UNetSuperRes)
        y = y + x
    if taps is not None:
        taps[prefix] = y
    return y


def _up(p: Dict[str, torch.Tensor], name: str, x1: torch.Tensor, x2: torch.Tensor, taps) -> torch.Tensor:
    """Up.forward, models/unet_model.py:80-94."""
    x1 = F.interpolate(x1, scale_factor=2, mode="bilinear", align_corners=True)       # :71
    x1 = F.conv2d(x1, p[f"{name}.up.1.weight"], None)                                   # :72
    x1 = _gn_lrelu(x1, p[f"{name}.up.2.weight"], p[f"{name}.up.2.bias"])               # :73-74
    dy = x2.size(2) - x1.size(2)
    dx = x2.size(3) - x1.size(3)
    if dy != 0 or dx != 0:                                                              # :86-90
        x1 = F.pad(x1, [dx // 2, dx - dx // 2, dy // 2, dy - dy // 2])
    if taps is not None:
        taps[f"{name}.up"] = x1
    return _double_conv(p, f"{name}.conv", torch.cat([x2, x1], dim=1), taps)            # :93-94


def unet_forward(p: Dict[str, torch.Tensor], x: torch.Tensor, taps: Dict[str, torch.Tensor] | None = None) -> torch.Tensor:
    """UNetSuperRes.forward, models/unet_model.py:189-211.  x: N x 1 x H x W -> N x 1 x 2H x 2W."""
    x1 = _double_conv(p, "inc", x, taps)                                                # :191
    x2 = _double_conv(p, "down1.maxpool_conv.1", F.max_pool2d(x1, 2), taps)            # :192
    x3 = _double_conv(p, "down2.maxpool_conv.1", F.max_pool2d(x2, 2), taps)            # :193
    x4 = _double_conv(p, "down3.maxpool_conv.1", F.max_pool2d(x3, 2), taps)            # :194
    y = _up(p, "up1", x4, x3, taps)                                                     # :197
    y = _up(p, "up2", y, x2, taps)                                                      # :198
    y = _up(p, "up3", y, x1, taps)                                                      # :199
    xb = F.interpolate(y, scale_factor=2, mode="bilinear", align_corners=True)          # :151
    xb = _gn_lrelu(F.conv2d(xb, p["final_up_bilinear.1.weight"], None, padding=1),
                   p["final_up_bilinear.2.weight"], p["final_up_bilinear.2.bias"])      # :152-154
    xp = F.conv2d(y, p["final_up_pixelshuffle.conv.weight"], p["final_up_pixelshuffle.conv.bias"], padding=1)
    xp = F.pixel_shuffle(xp, 2)                                                         # :110-111
    xp = _gn_lrelu(xp, p["final_up_pixelshuffle.norm.weight"], p["final_up_pixelshuffle.norm.bias"])
    a = torch.sigmoid(p["alpha"])                                                       # :206
    z = a * xb + (1 - a) * xp                                                           # :207
    if taps is not None:
        taps["xb"], taps["xp"], taps["blend"] = xb, xp, z
    z = _gn_lrelu(F.conv2d(z, p["final_conv.0.weight"], None, padding=1),
                  p["final_conv.1.weight"], p["final_conv.1.bias"])                     # :168-170
    z = F.conv2d(z, p["final_conv.3.weight"], p["final_conv.3.bias"])                   # :172
    return torch.sigmoid(z)                                                             # :211


# ----------------------------------------------------------------------------------------------
# bf16-storage restatement: the SAME algorithm with every operand the CUDA bf16 mode stores in
# bfloat16 rounded at exactly that point (input image, conv weights, raw conv outputs, activations,
# the low-resolution 1x1 output of Up.up, the blended head tensor); all arithmetic in fp32.
# Rounding is straight-through for autograd, so .backward() gives the gradient of the bf16
# forward with fp32 gradient arithmetic.  It separates "the kernels compute the bf16-operand
# algorithm exactly" (tight tolerance against THIS function) from the bf16-vs-fp32 noise floor of
# a 21-conv network (measured against unet_forward and reported in DESIGN.md).
class _RoundBF16(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).to(x.dtype)

    @staticmethod
    def backward(ctx, g):
        return g


def _q(x: torch.Tensor) -> torch.Tensor:
    return _RoundBF16.apply(x)


def _gn_lrelu_q(r: torch.Tensor, w: torch.Tensor, b: torch.Tensor, stats_from_rounded: bool = False) -> torch.Tensor:
    """GroupNorm statistics from the fp32 conv result (the conv epilogue accumulates them before
    rounding), normalisation applied to the stored bf16 tensor, result stored in bf16."""
    n, c = r.shape[:2]
    rq = _q(r)
    src = rq if stats_from_rounded else r
    g = src.reshape(n, GN_GROUPS, -1)
    mean = g.mean(2, keepdim=True)
    var = g.var(2, unbiased=False, keepdim=True)
    xn = ((rq.reshape(n, GN_GROUPS, -1) - mean) * torch.rsqrt(var + GN_EPS)).reshape(r.shape)
    return _q(F.leaky_relu(xn * w.view(1, c, 1, 1) + b.view(1, c, 1, 1), LRELU_SLOPE))


def unet_forward_bf16_storage(p: Dict[str, torch.Tensor], x: torch.Tensor, first_layer_fp32: bool = False) -> torch.Tensor:
    def dconv(prefix: str, t: torch.Tensor, first: bool = False) -> torch.Tensor:
        w0, w3 = p[f"{prefix}.double_conv.0.weight"], p[f"{prefix}.double_conv.3.weight"]
        r = F.conv2d(t, w0 if first and first_layer_fp32 else _q(w0), None, padding=1)
        t = _gn_lrelu_q(r, p[f"{prefix}.double_conv.1.weight"], p[f"{prefix}.double_conv.1.bias"])
        r = F.conv2d(t, _q(w3), None, padding=1)
        return _gn_lrelu_q(r, p[f"{prefix}.double_conv.4.weight"], p[f"{prefix}.double_conv.4.bias"])

    def up(name: str, x1: torch.Tensor, x2: torch.Tensor) -> torch.Tensor:
        # conv1x1 commuted below the bilinear interpolation (both linear, no bias): models/unet_model.py:70-72
        ql = _q(F.conv2d(x1, _q(p[f"{name}.up.1.weight"]), None))
        r = F.interpolate(ql, scale_factor=2, mode="bilinear", align_corners=True)
        u = _gn_lrelu_q(r, p[f"{name}.up.2.weight"], p[f"{name}.up.2.bias"], stats_from_rounded=True)
        return dconv(f"{name}.conv", torch.cat([x2, u], dim=1))

    x1 = dconv("inc", x if first_layer_fp32 else _q(x), first=True)
    x2 = dconv("down1.maxpool_conv.1", F.max_pool2d(x1, 2))
    x3 = dconv("down2.maxpool_conv.1", F.max_pool2d(x2, 2))
    x4 = dconv("down3.maxpool_conv.1", F.max_pool2d(x3, 2))
    y = up("up1", x4, x3)
    y = up("up2", y, x2)
    y = up("up3", y, x1)
    ub = _q(F.interpolate(y, scale_factor=2, mode="bilinear", align_corners=True))
    r17 = F.conv2d(ub, _q(p["final_up_bilinear.1.weight"]), None, padding=1)
    r18 = F.conv2d(y, _q(p["final_up_pixelshuffle.conv.weight"]), p["final_up_pixelshuffle.conv.bias"], padding=1)

    def gn_only(r, w, b):      # the head blends the two normalised branches in registers: only XM is stored
        n, c = r.shape[:2]
        g = r.reshape(n, GN_GROUPS, -1)
        mean, var = g.mean(2, keepdim=True), g.var(2, unbiased=False, keepdim=True)
        xn = ((_q(r).reshape(n, GN_GROUPS, -1) - mean) * torch.rsqrt(var + GN_EPS)).reshape(r.shape)
        return F.leaky_relu(xn * w.view(1, c, 1, 1) + b.view(1, c, 1, 1), LRELU_SLOPE)

    xb = gn_only(r17, p["final_up_bilinear.2.weight"], p["final_up_bilinear.2.bias"])
    xp = gn_only(F.pixel_shuffle(r18, 2), p["final_up_pixelshuffle.norm.weight"], p["final_up_pixelshuffle.norm.bias"])
    a = torch.sigmoid(p["alpha"])
    xm = _q(a * xb + (1 - a) * xp)
    r19 = F.conv2d(xm, _q(p["final_conv.0.weight"]), None, padding=1)
    z = gn_only(r19, p["final_conv.1.weight"], p["final_conv.1.bias"])
    z = F.conv2d(z, p["final_conv.3.weight"], p["final_conv.3.bias"])
    return torch.sigmoid(z)
