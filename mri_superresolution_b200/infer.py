"""Command line of the inference driver: the argument surface of the reference's scripts/infer.py:452-490 over the B200 path.

    python -m mri_superresolution_b200.infer --input lr.png --output sr.png [--target hr.png] [--checkpoint_dir ./checkpoints]

process_single_image (scripts/infer.py:230-396) with every array operation on the device: percentile clip + min-max of the input
(and of the target), the forward pass, clamp(0, 1), histogram matching against the normalised target when one is given
(:285-314), SSIM / PSNR / MSE / MAE (:147-228), ``(x * 255).astype(uint8)`` and the PNG write.  ``--input`` may also be a
directory: every PNG in it is processed in batches (the reference loops one file at a time) into the ``--output`` directory.
The matplotlib views (``--show_comparison``, ``--show_diff``) are outside the path (no display stack): accepted, noted."""
from __future__ import annotations

import argparse
import os
import sys


def parse_args(argv=None):
    p = argparse.ArgumentParser(description="MRI quality enhancement inference (B200 path)")
    p.add_argument("--input", type=str, required=True)
    p.add_argument("--output", type=str, required=True)
    p.add_argument("--target", type=str, default=None)
    p.add_argument("--checkpoint_dir", type=str, default="./checkpoints")
    p.add_argument("--checkpoint_path", type=str, default=None)
    p.add_argument("--model_type", type=str, choices=["unet"], default="unet")
    p.add_argument("--base_filters", type=int, default=64)
    p.add_argument("--show_comparison", action="store_true")
    p.add_argument("--show_diff", action="store_true")
    p.add_argument("--cpu", action="store_true")
    p.add_argument("--use_amp", action="store_true")
    p.add_argument("--precision", type=str, default="bf16", choices=["bf16", "fp32"])
    p.add_argument("--batch_size", type=int, default=64, help="slices per forward pass when --input is a directory")
    return p.parse_args(argv)


def enhance(model, inputs_u8, targets_u8=None):
    """uint8 N x H x W (device) -> (uint8 N x 2H x 2W enhanced images, metrics dict or None).  ``targets_u8``: N x 2H x 2W."""
    import torch
    from .utils.evaluation import calculate_metrics, match_histograms
    from .utils.preprocess import preprocess_images
    with torch.no_grad():
        x = preprocess_images(inputs_u8)                                           # scripts/infer.py:97-130
        out = torch.clamp(model(x), 0, 1)                                          # :268-276
        adjusted, metrics = out[:, 0], None
        if targets_u8 is not None:
            tgt = preprocess_images(targets_u8)                                    # :243-256 (same clip / normalise as the input)
            if tgt.shape[-2:] != out.shape[-2:]:
                tgt = torch.nn.functional.interpolate(tgt, size=out.shape[-2:], mode="bicubic", align_corners=False)   # :321-325
            adjusted = match_histograms(out[:, 0], tgt[:, 0])                      # :285-309
            metrics = calculate_metrics(tgt[:, 0], out[:, 0])                      # on the raw model output, as :327 does
        return (adjusted * 255).to(torch.uint8), metrics                           # :336 (truncation)


def main(argv=None) -> int:
    args = parse_args(argv)
    import numpy as np
    import torch
    from PIL import Image
    from .drivers import find_best_checkpoint, load_model, log_message
    if args.cpu or not torch.cuda.is_available():
        raise SystemExit("mri_superresolution_b200.infer: this path runs on CUDA devices only (there is no CPU fallback)")
    device = torch.device("cuda")
    if args.checkpoint_path and os.path.exists(args.checkpoint_path):
        ckpt = args.checkpoint_path
    else:
        ckpt = find_best_checkpoint(args.checkpoint_dir, args.model_type)
    model = load_model(args.model_type, ckpt, device, base_filters=args.base_filters, precision=args.precision)
    if args.show_comparison or args.show_diff:
        log_message("matplotlib views are outside the B200 path (no display stack): skipped")
    if os.path.isdir(args.input):
        names = sorted(f for f in os.listdir(args.input) if f.lower().endswith(".png"))
        os.makedirs(args.output, exist_ok=True)
        pairs = [(os.path.join(args.input, f), os.path.join(args.output, f),
                  os.path.join(args.target, f) if args.target and os.path.exists(os.path.join(args.target, f)) else None) for f in names]
    else:
        pairs = [(args.input, args.output, args.target if args.target and os.path.exists(args.target) else None)]
    load = lambda p: np.asarray(Image.open(p).convert("L"))
    for s in range(0, len(pairs), args.batch_size):
        chunk = pairs[s:s + args.batch_size]
        imgs = [load(p[0]) for p in chunk]
        groups = {}
        for i, im in enumerate(imgs):                                              # one forward pass per distinct slice size
            groups.setdefault((im.shape, chunk[i][2] is not None), []).append(i)
        for (shape, has_t), ids in groups.items():
            if shape[0] % 8 or shape[1] % 8:
                raise SystemExit(f"{chunk[ids[0]][0]}: {shape[0]}x{shape[1]} is not divisible by 8 (three 2x poolings); pad or crop first")
            x = torch.from_numpy(np.stack([imgs[i] for i in ids])).to(device)
            t = None
            if has_t:
                tg = [load(chunk[i][2]) for i in ids]
                if len({a.shape for a in tg}) == 1:
                    t = torch.from_numpy(np.stack(tg)).to(device)
            out, metrics = enhance(model, x, t)
            out = out.cpu().numpy()
            for k, i in enumerate(ids):
                Image.fromarray(out[k]).save(chunk[i][1])
                msg = {"input": chunk[i][0], "output": chunk[i][1]}
                if metrics is not None:
                    msg.update({name: float(v[k]) for name, v in metrics.items()})
                log_message(msg, "inference")
    log_message("Inference completed successfully!")
    return 0


if __name__ == "__main__":
    sys.exit(main())
