"""Command line of the training driver: the argument surface of the reference's scripts/train.py:486-548 (same flags, same
defaults) over the B200 path.

    python -m mri_superresolution_b200.train --full_res_dir HR --low_res_dir LR [--batch_size 64 --augmentation ...]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 -m mri_superresolution_b200.train ...

Under torchrun (WORLD_SIZE > 1) every rank opens the same PNG directories, makes the same seeded train / validation split
(scripts/train.py:205-213) and reads its own shard of every global batch (data.UInt8BatchLoader); gradients are exchanged by
parallel.GradientBuckets inside the recorded step; rank 0 logs and writes checkpoints.  ``--batch_size`` is per GPU.
Flags of the reference that have no meaning here are accepted and answered: ``--cpu`` raises (no CPU fallback), ``--use_amp``
selects what the path always does (bf16 operands, fp32 accumulation; ``--precision fp32`` is the parity mode),
``--use_tensorboard`` / ``--log_dir`` are noted as out of scope.
"""
from __future__ import annotations

import argparse
import os
import random
import sys


def parse_args(argv=None):
    p = argparse.ArgumentParser(description="Train MRI quality enhancement model (B200 path)")
    p.add_argument("--full_res_dir", type=str, required=True, help="Directory containing high-quality MRI slices")
    p.add_argument("--low_res_dir", type=str, required=True, help="Directory containing low-quality MRI slices")
    p.add_argument("--model_type", type=str, choices=["unet"], default="unet")
    p.add_argument("--base_filters", type=int, default=32)
    p.add_argument("--batch_size", type=int, default=8)
    p.add_argument("--epochs", type=int, default=100)
    p.add_argument("--learning_rate", type=float, default=1e-4)
    p.add_argument("--weight_decay", type=float, default=1e-5)
    p.add_argument("--ssim_weight", type=float, default=0.3)
    p.add_argument("--perceptual_weight", type=float, default=0.0)
    p.add_argument("--vgg_layer_idx", type=int, default=35)
    p.add_argument("--perceptual_loss_type", type=str, default="l1", choices=["l1", "l2", "mse"])
    p.add_argument("--initial_alpha", type=float, default=0.0)
    p.add_argument("--validation_split", type=float, default=0.2)
    p.add_argument("--patience", type=int, default=10)
    p.add_argument("--num_workers", type=int, default=min(16, os.cpu_count() or 1))
    p.add_argument("--seed", type=int, default=random.randint(1, 10000))
    p.add_argument("--augmentation", action="store_true")
    p.add_argument("--use_tensorboard", action="store_true")
    p.add_argument("--use_amp", action="store_true")
    p.add_argument("--cpu", action="store_true")
    p.add_argument("--checkpoint_dir", type=str, default="./checkpoints")
    p.add_argument("--log_dir", type=str, default="./logs")
    # additions of this path
    p.add_argument("--precision", type=str, default="bf16", choices=["bf16", "fp32"], help="fp32 = CUDA-core parity mode")
    p.add_argument("--no_graph", action="store_true", help="launch the step eagerly instead of replaying one CUDA graph")
    return p.parse_args(argv)


def main(argv=None) -> int:
    args = parse_args(argv)
    import torch
    import torch.distributed as dist
    from .data import PairedPNGDataset, UInt8BatchLoader, split_indices
    from .drivers import TrainConfig, log_message, train

    if args.cpu or not torch.cuda.is_available():
        raise SystemExit("mri_superresolution_b200.train: this path runs on CUDA devices only (there is no CPU fallback)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    random.seed(args.seed)                                                   # scripts/train.py:145-150
    torch.manual_seed(args.seed)
    if args.use_tensorboard and rank == 0:
        log_message("TensorBoard logging is outside the B200 path (SURVEY.md section 2): the JSON-lines stream carries the same numbers")

    dataset = PairedPNGDataset(args.full_res_dir, args.low_res_dir)
    if len(dataset) == 0:
        raise SystemExit(f"no PNG pairs under {args.full_res_dir} / {args.low_res_dir}")
    tr_idx, va_idx = split_indices(len(dataset), args.validation_split, args.seed)
    if not va_idx:
        raise SystemExit("validation_split leaves no validation data (the reference divides by its length, scripts/train.py:390)")
    tr = UInt8BatchLoader(dataset, tr_idx, args.batch_size, shuffle=True, seed=args.seed, num_workers=args.num_workers, rank=rank, world=world)
    va = UInt8BatchLoader(dataset, va_idx, args.batch_size, shuffle=False, num_workers=args.num_workers, rank=rank, world=world)
    if len(tr) == 0 or len(va) == 0:
        raise SystemExit(f"not enough pairs for batch_size {args.batch_size} x {world} ranks")
    cfg = TrainConfig(epochs=args.epochs, learning_rate=args.learning_rate, weight_decay=args.weight_decay, ssim_weight=args.ssim_weight,
                      perceptual_weight=args.perceptual_weight, vgg_layer_idx=args.vgg_layer_idx,
                      perceptual_loss_type=args.perceptual_loss_type, base_filters=args.base_filters, initial_alpha=args.initial_alpha,
                      patience=args.patience, checkpoint_dir=args.checkpoint_dir, model_type=args.model_type, seed=args.seed,
                      use_graph=not args.no_graph, precision=args.precision, batch_size=args.batch_size, augmentation=args.augmentation)
    try:
        train(cfg, lambda: tr, lambda: va, device=device)
    finally:
        if world > 1:
            torch.cuda.synchronize()
            dist.barrier()
            dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
