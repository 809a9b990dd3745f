# L1 + SSIM loss, its gradient and the SSIM metric at odd shapes / window sizes against the oracle's autograd (run through gpurun).
import sys
import torch
sys.path.insert(0, '.')
from oracle import losses_oracle as LO
from mri_superresolution_b200.utils.losses import CombinedLoss, SSIM
dev = torch.device('cuda')
bad = 0
for ws, sigma in ((11, 1.5), (7, 1.0), (5, 0.8), (3, 1.5)):
    for n, h, w in ((1, 1, 1), (2, 3, 5), (1, 11, 23), (3, 33, 31), (2, 64, 100), (1, 257, 129), (2, 512, 512)):
        g = torch.Generator().manual_seed(h * 7 + w + ws)
        p = torch.rand(n, 1, h, w, generator=g)
        t = (p + 0.1 * torch.randn(n, 1, h, w, generator=g)).clamp(0, 1)
        for sw in (0.4, 1.0):
            try:
                pr = p.clone().requires_grad_(True)
                ref = (1 - sw) * (pr - t).abs().mean() + sw * (1 - LO.ssim(pr, t, window_size=ws, sigma=sigma).clamp(0, 1))
                ref.backward()
                pd = p.to(dev).requires_grad_(True)
                crit = CombinedLoss(ssim_weight=sw, window_size=ws, sigma=sigma)
                loss = crit(pd, t.to(dev))
                loss.backward()
                e_l = abs(float(loss) - float(ref))
                e_g = float((pd.grad.cpu() - pr.grad).abs().max()) / max(float(pr.grad.abs().max()), 1e-12)
                m = float(SSIM(window_size=ws, sigma=sigma)(p.to(dev), t.to(dev)))
                e_m = abs(m - float(LO.ssim(p, t, window_size=ws, sigma=sigma)))
                ok = e_l < 2e-5 and e_g < 5e-4 and e_m < 2e-5
                bad += not ok
                print(f"win {ws} sigma {sigma} {n}x{h}x{w} ssim_w {sw}: loss err {e_l:.2e} grad rel err {e_g:.2e} metric err {e_m:.2e} {'ok' if ok else 'MISMATCH'}", flush=True)
            except Exception as ex:
                bad += 1
                print(f"win {ws} sigma {sigma} {n}x{h}x{w} ssim_w {sw}: FAILED {type(ex).__name__}: {str(ex)[:200]}", flush=True)
print("problems:", bad)
