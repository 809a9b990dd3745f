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
