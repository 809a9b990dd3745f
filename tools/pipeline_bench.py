"""Throughput of the adjacent rows at the path's batch size on one GPU, CUDA events, inputs resident:
  python tools/pipeline_bench.py          (run on the GPU box)"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from mri_superresolution_b200.utils import evaluation as EV
from mri_superresolution_b200.utils import kspace as KS
from mri_superresolution_b200.utils import preprocess as PP


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    dev = torch.device("cuda:0")
    n = 64
    g = torch.Generator(device="cpu").manual_seed(0)
    lr = torch.randint(0, 256, (n, 256, 256), dtype=torch.uint8, generator=g).to(dev)
    hr = torch.randint(0, 256, (n, 512, 512), dtype=torch.uint8, generator=g).to(dev)
    aug = PP.PairAugmenter()
    decs = [aug.draw() for _ in range(n)]
    rows = []
    rows.append(("preprocess_images 64 x 256^2 (u8 -> f32)", timed(lambda: PP.preprocess_images(lr)), n, lr.numel() * 5))
    rows.append(("to_tensor 64 x 512^2", timed(lambda: PP.to_tensor(hr)), n, hr.numel() * 5))
    rows.append(("augment pair 64 x (512^2 + 256^2)", timed(lambda: aug(hr, lr, decisions=decs)), n, (hr.numel() + lr.numel()) * 8))
    for m in ("bilinear", "sharp_bilinear", "bicubic"):
        rows.append((f"upscale {m} 64 x 256^2 -> 512^2", timed(lambda: EV.upscale_with_interpolation(lr, m)), n, lr.numel() * (1 + 4 + 4 + 16)))
    a = torch.rand(n, 512, 512, device=dev)
    b = (a + 0.05 * torch.randn_like(a)).clamp(0, 1)
    rows.append(("calculate_metrics 64 x 512^2 (incl. SSIM)", timed(lambda: EV.calculate_metrics(a, b)), n, a.numel() * 16))
    rows.append(("match_histograms 64 x 512^2", timed(lambda: EV.match_histograms(b, a), reps=3), n, 0))
    vol = torch.rand(n, 512, 512, device=dev, dtype=torch.float64) * 900
    rows.append(("robust_normalize 64 x 512^2 f64", timed(lambda: KS.robust_normalize(vol), reps=3), n, 0))
    x = torch.rand(n, 512, 512, device=dev)
    rows.append(("simulate_low_field 64 x 512^2", timed(lambda: KS.simulate_low_field_mri(x, seed=1), reps=5), n, 0))
    rows.append(("make_low_res 64 x 512^2 -> u8 256^2", timed(lambda: KS.make_low_res(vol, seed=1), reps=3), n, 0))
    print("| stage | ms per batch of 64 | slices/s | GB/s (algorithmic) |")
    print("|---|---|---|---|")
    for name, ms, cnt, nbytes in rows:
        print(f"| {name} | {ms:.3f} | {cnt / ms * 1e3:,.0f} | {nbytes / ms / 1e6:.0f} |" if nbytes else f"| {name} | {ms:.3f} | {cnt / ms * 1e3:,.0f} | - |")
    # the reference's CPU code for the same stages, one image at a time (its own loop structure), on this box's host
    from oracle import eval_oracle as EO
    from oracle import preproc_oracle as PO
    lr_np = lr[:8].cpu().numpy()
    t0 = time.perf_counter()
    for k in range(8):
        PO.preprocess_u8(lr_np[k])
    t = (time.perf_counter() - t0) / 8
    print(f"\nCPU (numpy, as the reference calls it) preprocess_image: {t * 1e3:.2f} ms per 256^2 slice = {1 / t:,.0f} slices/s per core")


if __name__ == "__main__":
    main()
