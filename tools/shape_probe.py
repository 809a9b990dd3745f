# Inference at large, non-square and tiny shapes against the oracle on the same GPU (run through gpurun; prints one line per shape).
import torch, sys
sys.path.insert(0, '.')
from oracle import unet_oracle as UO
from mri_superresolution_b200.models.unet_model import UNetSuperRes
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device('cuda')
for f, n, h, w in [(32, 1, 1024, 1024), (32, 2, 768, 1280), (32, 1, 8, 8), (32, 3, 16, 8), (64, 1, 512, 512), (16, 2, 264, 520), (32, 1, 2048, 1024)]:
    params = UO.init_params(f, seed=3)
    m = UNetSuperRes(base_filters=f).to(dev); m.load_state_dict(params); m.eval()
    x = torch.rand(n, 1, h, w, device=dev)
    try:
        with torch.no_grad():
            y = m(x)
            ref = UO.unet_forward({k: v.to(dev) for k, v in params.items()}, x)
        torch.cuda.synchronize()
        if ref is not None:
            if isinstance(ref, tuple): ref = ref[0]
            e = float((y - ref).norm() / ref.norm())
        else:
            e = float('nan')
        print(f"f={f} {n}x{h}x{w}: ok out {tuple(y.shape)} rel l2 vs oracle {e:.3e} finite {bool(torch.isfinite(y).all())}")
    except Exception as ex:
        print(f"f={f} {n}x{h}x{w}: FAILED {type(ex).__name__}: {str(ex)[:200]}")
    del m; torch.cuda.empty_cache()

# ---- one training step (forward, L1 + 0.4 SSIM, backward) at odd shapes: every parameter gradient against the oracle's autograd
from mri_superresolution_b200.utils.losses import CombinedLoss
from oracle import losses_oracle as LO
for f, n, h, w in [(32, 2, 136, 200), (32, 1, 520, 264), (64, 1, 264, 136), (16, 2, 72, 328), (32, 5, 8, 8), (32, 1, 1024, 512)]:
    params = UO.randomize_affine(UO.init_params(f, seed=5))
    try:
        m = UNetSuperRes(base_filters=f).to(dev); m.load_state_dict(params); m.train()
        x = torch.rand(n, 1, h, w, device=dev); t = torch.rand(n, 1, 2 * h, 2 * w, device=dev)
        loss = CombinedLoss(ssim_weight=0.4)(m(x), t); loss.backward()
        p = {k: v.to(dev).requires_grad_(True) for k, v in params.items()}
        y = UO.unet_forward(p, x)
        lref = 0.6 * (y - t).abs().mean() + 0.4 * (1 - LO.ssim(y, t).clamp(0, 1)); lref.backward()
        worst, wname, wn = 2.0, "", 0.0
        for k, v in m.named_parameters():
            g, r = v.grad.double().reshape(-1), p[k].grad.double().reshape(-1)
            c = float((g * r).sum() / (g.norm() * r.norm() + 1e-30))
            if c < worst: worst, wname, wn = c, k, float(g.norm() / (r.norm() + 1e-30))
        ok = all(bool(torch.isfinite(v.grad).all()) for v in m.parameters())
        print(f"train f={f} {n}x{h}x{w}: loss {float(loss):.5f} vs {float(lref):.5f}; worst gradient cosine {worst:.4f} ({wname}, norm ratio {wn:.3f}); finite {ok}")
    except Exception as ex:
        print(f"train f={f} {n}x{h}x{w}: FAILED {type(ex).__name__}: {str(ex)[:200]}")
    del m; torch.cuda.empty_cache()
