"""Prints the bf16-mode error budget on the GPU box (test infrastructure): CUDA bf16 path vs the fp32 golden output and vs
the bf16-storage oracle (oracle/unet_oracle.py: unet_forward_bf16_storage).  Run: PYTHONPATH=. python tests/bf16_metrics.py"""
import sys
import numpy as np
import torch
from oracle import losses_oracle as LO, unet_oracle as UO
from oracle.synth import make_pairs
from mri_superresolution_b200.models.unet_model import UNetSuperRes
from mri_superresolution_b200.utils.losses import CombinedLoss, SSIM

dev = torch.device('cuda:0')
for name in ["unet_f16_n2_h32", "unet_f16_n1_h24x40", "unet_f32_n1_h64"]:
    d = dict(np.load(f"tests/golden/{name}.npz"))
    f, n, h, w, affine = [int(v) for v in d["meta"]]
    params = UO.init_params(f, seed=1234)
    if affine:
        params = UO.randomize_affine(params)
    x, tgt = make_pairs(n, h, w, seed=1234)
    m = UNetSuperRes(base_filters=f, precision="bf16").to(dev)
    m.load_state_dict(params)
    m.train()
    sw = float(d["ssim_weight"])
    crit = CombinedLoss(ssim_weight=sw)
    out = m(x.to(dev))
    loss = crit(out, tgt.to(dev))
    loss.backward()
    metric = SSIM()(out.detach(), tgt.to(dev)).item()
    ref = torch.from_numpy(d["out"])
    o = out.detach().cpu()
    pq = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    oq = UO.unet_forward_bf16_storage(pq, x)
    lq = LO.combined_loss(oq, tgt, ssim_weight=sw)
    lq.backward()
    oqd = oq.detach()
    print(name, "vs fp32 golden: out relL2 %.3e max %.3e loss rel %.3e ssim d %.3e psnr d %.4f" % (
        float((o - ref).norm() / ref.norm()), float((o - ref).abs().max()), abs(loss.item() - float(d["loss"])) / float(d["loss"]),
        abs(metric - float(d["ssim"])), abs(LO.psnr(o, tgt) - LO.psnr(ref, tgt))))
    print("   vs bf16-storage oracle: out relL2 %.3e max %.3e loss rel %.3e" % (
        float((o - oqd).norm() / oqd.norm()), float((o - oqd).abs().max()), abs(loss.item() - lq.item()) / lq.item()))
    worst = (1.0, None)
    worstn = (0.0, None)
    for k, p in m.named_parameters():
        r = pq[k].grad.double()
        g = p.grad.double().cpu().view(r.shape)
        if r.numel() > 1:
            cos = float((g * r).sum() / (g.norm() * r.norm() + 1e-30))
            if cos < worst[0]:
                worst = (cos, k)
        nd = abs(float(g.norm()) - float(r.norm())) / (float(r.norm()) + 1e-12)
        if nd > worstn[0]:
            worstn = (nd, k)
    print("   grads vs bf16-storage oracle: worst cosine", worst, "worst norm dev", worstn)
