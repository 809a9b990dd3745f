"""Helper of test_gpu_knobs.py: one U-Net training-loss forward + backward at a mid-size shape, gradients and outputs
written to an .npz.  Run in a fresh process per kernel-selection knob (the library reads its MRISR_* knobs once)."""
import sys

import numpy as np
import torch


def main(path, f, n, h, w):
    from mri_superresolution_b200.models.unet_model import UNetSuperRes
    from mri_superresolution_b200.utils.losses import CombinedLoss
    from mri_superresolution_b200.utils.synthetic import make_pairs
    dev = torch.device("cuda:0")
    torch.manual_seed(1234)
    model = UNetSuperRes(base_filters=f, precision="bf16").to(dev)
    x, t = make_pairs(n, h, w, seed=7)
    x, t = x.to(dev), t.to(dev)
    crit = CombinedLoss(ssim_weight=0.4)
    out = model(x)
    loss = crit(out, t)
    loss.backward()
    res = {"out": out.detach().float().cpu().numpy(), "loss": np.array(float(loss))}
    for k, p in model.named_parameters():
        res["grad:" + k] = p.grad.detach().float().cpu().numpy()
    np.savez(path, **res)


if __name__ == "__main__":
    main(sys.argv[1], *[int(v) for v in sys.argv[2:6]])
