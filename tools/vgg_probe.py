# VGG perceptual term at several feature layers and shapes, bf16 and fp32 mode, against the oracle (run through gpurun).
import sys
import torch
sys.path.insert(0, '.')
from oracle import losses_oracle as LO
from mri_superresolution_b200.utils.perceptual import PerceptualLoss, VGGFeatureExtractor
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device('cuda')
vggp = LO.make_vgg_params(seed=1235)
full = [c for c in LO.VGG19_CFG]
bad = 0
# feature_layer_idx (a ReLU of vgg19.features) -> how many cfg items it covers
def cfg_prefix(idx):
    out, pos = [], -1
    for item in full:
        if item == "M":
            pos += 1
            if pos > idx: break
            out.append(item)
        else:
            pos += 2                      # conv, relu
            if pos > idx: break
            out.append(item)
    return out
for idx in (1, 3, 8, 17, 26, 35):
    cfg = cfg_prefix(idx)
    nconv = sum(1 for c in cfg if c != "M")
    div = 2 ** sum(1 for c in cfg if c == "M")
    for n, h, w in ((1, 16, 16), (2, 48, 80), (1, 160, 96)):
        if h % div or w % div:
            continue
        g = torch.Generator().manual_seed(idx * 100 + h)
        p = torch.rand(n, 1, h, w, generator=g); t = (p + 0.2 * torch.randn(n, 1, h, w, generator=g)).clamp(0, 1)
        for typ in ("l1", "mse"):
            prc = p.clone().requires_grad_(True)
            fg = LO.vgg_features(vggp[:nconv], prc, cfg); ft = LO.vgg_features(vggp[:nconv], t, cfg).detach()
            ref = (fg - ft).abs().mean() if typ == "l1" else ((fg - ft) ** 2).mean()
            ref.backward()
            # what bf16 storage alone does to this loss (oracle restatement with rounded operands, fp32 arithmetic): the floor
            fq, ftq = LO.vgg_features_bf16_storage(vggp[:nconv], p, cfg), LO.vgg_features_bf16_storage(vggp[:nconv], t, cfg)
            lq = (fq - ftq).abs().mean() if typ == "l1" else ((fq - ftq) ** 2).mean()
            floor = abs(float(lq) - float(ref.detach())) / max(abs(float(ref.detach())), 1e-12)
            for prec, tol_l, tol_g in (("fp32", 1e-4, 5e-3), ("bf16", max(2e-2, 2 * floor), 0.45)):
                try:
                    crit = PerceptualLoss(idx, typ, vgg_weights=vggp[:nconv], precision=prec)
                    pd = p.clone().to(dev).requires_grad_(True)
                    loss = crit(pd, t.to(dev)); loss.backward()
                    e_l = abs(float(loss.detach()) - float(ref.detach())) / max(abs(float(ref.detach())), 1e-12)
                    gd, gr = pd.grad.cpu().double().reshape(-1), prc.grad.double().reshape(-1)
                    e_g = float((gd - gr).norm() / max(float(gr.norm()), 1e-30))
                    ok = e_l < tol_l and e_g < tol_g
                    bad += not ok
                    print(f"idx {idx:2d} ({nconv:2d} convs) {n}x{h}x{w} {typ:3s} {prec}: loss rel err {e_l:.2e} grad rel L2 {e_g:.2e}{f' (bf16-storage floor of the loss {floor:.1e})' if prec == 'bf16' else ''} {'ok' if ok else 'MISMATCH'}", flush=True)
                except Exception as ex:
                    bad += 1
                    print(f"idx {idx:2d} {n}x{h}x{w} {typ} {prec}: FAILED {type(ex).__name__}: {str(ex)[:200]}", flush=True)
# shapes the kernels refuse must raise, not crash
for h, w in ((24, 24), (8, 40)):
    try:
        PerceptualLoss(35, "l1", vgg_weights=vggp)(torch.rand(1, 1, h, w, device=dev), torch.rand(1, 1, h, w, device=dev))
        print(f"35 at {h}x{w}: accepted"); 
    except (ValueError, RuntimeError) as ex:
        print(f"35 at {h}x{w}: refused with {type(ex).__name__}: {str(ex)[:120]}")
print("problems:", bad)
