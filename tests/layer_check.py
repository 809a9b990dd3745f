"""Layer-by-layer comparison of the CUDA path against a torch fp32 replica of the oracle's forward and
autograd (run on the GPU box: ``python -m tests.layer_check [fp32|bf16] [f] [n] [h] [w]``).

Test infrastructure: localises a parity failure to one internal tensor through mrisr_unet_debug_tensor.
"""
from __future__ import annotations

import ctypes
import sys

import torch
import torch.nn.functional as F

from mri_superresolution_b200 import _native
from mri_superresolution_b200.models.unet_model import UNetSuperRes
from mri_superresolution_b200.utils.losses import CombinedLoss
from oracle import losses_oracle as LO
from oracle import unet_oracle as UO
from oracle.synth import make_pairs

KINDS = {"R": 0, "A": 1, "P": 2, "Q": 3, "dR": 4, "G": 5, "dQ": 6, "UB": 7, "XM": 8, "DY3": 9}


def native_tensor(model: UNetSuperRes, n, h, w, kind: str, layer: int = 0) -> torch.Tensor:
    ws = model._workspace(n, h, w, True)
    prec = _native.BF16 if model.precision == "bf16" else _native.FP32
    dims = (ctypes.c_int * 4)()
    L = _native.lib()
    rc = L.mrisr_unet_debug_tensor(_native.ptr(ws), n, h, w, model.base_filters, prec, KINDS[kind], layer, None,
                                   ctypes.cast(dims, ctypes.c_void_p), _native.stream_ptr())
    _native.check(rc, "debug_tensor")
    out = torch.empty(tuple(dims), dtype=torch.float32, device=ws.device)
    rc = L.mrisr_unet_debug_tensor(_native.ptr(ws), n, h, w, model.base_filters, prec, KINDS[kind], layer,
                                   _native.ptr(out), None, _native.stream_ptr())
    _native.check(rc, "debug_tensor")
    return out


def replica_forward(p, x):
    """Oracle forward (oracle/unet_oracle.py) with every internal tensor kept: returns (y, dict)."""
    t = {}

    def keep(name, v):
        v.retain_grad()
        t[name] = v
        return v

    gn = lambda v, l, g, b: F.leaky_relu(F.group_norm(v, 8, p[g], p[b], 1e-5), 0.2)
    names = [n for n, _ in UO.param_specs(p["final_conv.0.weight"].shape[0] * 2)]
    # conv table mirrors csrc/unet.cu conv_table: (weight idx, gamma idx, beta idx)
    tab = [(1, 2, 3), (4, 5, 6), (7, 8, 9), (10, 11, 12), (13, 14, 15), (16, 17, 18), (19, 20, 21), (22, 23, 24),
           (25, 26, 27), (28, 29, 30), (31, 32, 33), (34, 35, 36), (37, 38, 39), (40, 41, 42), (43, 44, 45),
           (46, 47, 48), (49, 50, 51), (52, 53, 54), (55, 57, 58), (59, 60, 61)]
    W = lambda l: p[names[tab[l][0]]]

    def block(l, src, pad=1):
        r = keep(f"R{l}", F.conv2d(src, W(l), None, padding=pad))
        return keep(f"A{l}", gn(r, l, names[tab[l][1]], names[tab[l][2]]))

    def up(l, src):
        u = F.interpolate(src, scale_factor=2, mode="bilinear", align_corners=True)
        return block(l, u, pad=0)

    a0 = block(0, x); a1 = block(1, a0)
    a2 = block(2, keep("P1", F.max_pool2d(a1, 2))); a3 = block(3, a2)
    a4 = block(4, keep("P3", F.max_pool2d(a3, 2))); a5 = block(5, a4)
    a6 = block(6, keep("P5", F.max_pool2d(a5, 2))); a7 = block(7, a6)
    a8 = up(8, a7); a9 = block(9, keep("CAT9", torch.cat([a5, a8], 1))); a10 = block(10, a9)
    a11 = up(11, a10); a12 = block(12, keep("CAT12", torch.cat([a3, a11], 1))); a13 = block(13, a12)
    a14 = up(14, a13); a15 = block(15, keep("CAT15", torch.cat([a1, a14], 1))); a16 = block(16, a15)
    ub = keep("UB", F.interpolate(a16, scale_factor=2, mode="bilinear", align_corners=True))
    r17 = keep("R17", F.conv2d(ub, W(17), None, padding=1))
    xb = gn(r17, 17, names[53], names[54])
    r18 = keep("R18", F.conv2d(a16, W(18), p[names[56]], padding=1))
    xp = gn(F.pixel_shuffle(r18, 2), 18, names[57], names[58])
    al = torch.sigmoid(p["alpha"])
    xm = keep("XM", al * xb + (1 - al) * xp)
    r19 = keep("R19", F.conv2d(xm, W(19), None, padding=1))
    a19 = gn(r19, 19, names[60], names[61])
    y = torch.sigmoid(F.conv2d(a19, p[names[62]], p[names[63]]))
    return y, t


def err(a, b):
    d = (a - b).abs().max().item()
    s = b.abs().max().item()
    return d / max(s, 1e-30), s


def main(precision="fp32", f=16, n=2, h=32, w=32, ssim_w=0.4):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda")
    params = UO.randomize_affine(UO.init_params(f, seed=1234))
    x, tgt = make_pairs(n, h, w, seed=1234)
    x, tgt = x.to(dev), tgt.to(dev)
    model = UNetSuperRes(base_filters=f, precision=precision).to(dev)
    model.load_state_dict(params)
    crit = CombinedLoss(ssim_weight=ssim_w)
    y = model(x)
    loss = crit(y, tgt)
    loss.backward()
    p = {k: v.to(dev).clone().requires_grad_(True) for k, v in params.items()}
    yr, t = replica_forward(p, x)
    lossr = LO.combined_loss(yr, tgt, ssim_weight=ssim_w)
    lossr.backward()
    print(f"precision={precision} f={f} n={n} h={h} w={w}")
    print(f"loss native {loss.item():.7f} oracle {lossr.item():.7f}")
    print("y      rel err %.3e (scale %.3e)" % err(y.detach(), yr.detach()))
    rows = []
    for l in range(20):
        rows.append((f"R{l}", err(native_tensor(model, n, h, w, "R", l), t[f"R{l}"].detach())))
        if l < 17:
            rows.append((f"A{l}", err(native_tensor(model, n, h, w, "A", l), t[f"A{l}"].detach())))
    rows.append(("UB", err(native_tensor(model, n, h, w, "UB"), t["UB"].detach())))
    rows.append(("XM", err(native_tensor(model, n, h, w, "XM"), t["XM"].detach())))
    for l in (1, 3, 5):
        rows.append((f"P{l}", err(native_tensor(model, n, h, w, "P", l), t[f"P{l}"].detach())))
    for l in range(19, -1, -1):
        if l in (8, 11, 14):
            continue     # native dR[l] is taken after the (commuted) upsample: different tensor, checked through G/dQ
        rows.append((f"dR{l}", err(native_tensor(model, n, h, w, "dR", l), t[f"R{l}"].grad)))
    gin = {19: "XM", 17: "UB", 16: "A15", 15: "CAT15", 13: "A12", 12: "CAT12", 10: "A9", 9: "CAT9", 7: "A6", 6: "P5",
           5: "A4", 4: "P3", 3: "A2", 2: "P1", 1: "A0", 14: "A13", 11: "A10", 8: "A7"}
    for l, name in gin.items():
        g = t[name].grad
        if g is None:
            continue
        nat = native_tensor(model, n, h, w, "G", l)
        if name in ("A13", "A10", "A7", "A15", "A12", "A9", "A6", "A4", "A2", "A0"):
            pass                                      # single consumer: grad of the tensor == G[l]
        rows.append((f"G{l}({name})", err(nat, g)))
    for name, e in rows:
        print(f"{name:12s} rel err {e[0]:.3e} (scale {e[1]:.3e})")
    print("--- parameter gradients")
    for (k, v), pn in zip(p.items(), model._param_list()):
        e = err(pn.grad, v.grad)
        flag = " <<<" if e[0] > (2e-3 if precision == "fp32" else 8e-2) else ""
        print(f"{k:45s} rel err {e[0]:.3e} (scale {e[1]:.3e}){flag}")


if __name__ == "__main__":
    a = sys.argv[1:]
    main(a[0] if a else "fp32", *[int(v) for v in a[1:5]])
