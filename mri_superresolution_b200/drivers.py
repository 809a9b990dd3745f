"""The callers of the hot path, restated around the B200 modules: the epoch loop of
scripts/train.py:278-477 and the model-loading / forward half of scripts/infer.py:41-95,268-276.

What is kept from the reference, because other tools read it:
  * the JSON-lines progress schema on stdout: ``params`` (:254-269), ``batch_update`` (:326-333), ``epoch_summary``
    (:431-441), plain ``info`` messages (:54-68) - floats rounded to 6 places, ``type`` key last-writer-wins;
  * checkpoint files ``best_model_unet.pth`` / ``final_model_unet.pth`` with keys ``epoch, model_state_dict,
    optimizer_state_dict, scheduler_state_dict (best only), val_loss, val_ssim`` (:405-418, :467-477), the optimizer state in
    the stock ``torch.optim.Adam`` layout (optim.FusedAdam.state_dict);
  * validation every epoch, ``ReduceLROnPlateau(mode='min', factor=0.5, patience=patience // 2)`` stepped on the validation
    loss (:189,:397), best-model tracking and early stopping after ``patience`` epochs without improvement (:400-424,:462).
What changes, because the reference's loop is host-bound at B200 speeds (5-6 device->host syncs per step):
  * the step is one CUDA-graph launch (utils/graph.GraphedTrainStep) and batches arrive through DevicePrefetcher;
  * loss and SSIM-metric sums are accumulated on the device and read once per epoch (plus one scalar for a ``batch_update``
    line); the reference's second SSIM forward per step (:322-323) is replaced by the mean SSIM the loss kernel already
    produced for the same (output, target) pair (``CombinedLoss.last_components``) when ``ssim_weight > 0``.
No file-format or image IO lives here: batches are any iterable of ``(low_res, high_res)`` CPU or CUDA tensor pairs.
"""
from __future__ import annotations

import json
import os
import time
from typing import Callable, Dict, Iterable, Optional

import torch
import torch.distributed as dist

from .models.unet_model import UNetSuperRes
from .optim import FusedAdam
from .utils.losses import SSIM, CombinedLoss

__all__ = ["log_message", "TrainConfig", "train", "load_model", "find_best_checkpoint", "super_resolve"]


def log_message(message, message_type: str = "info", stream=None) -> dict:
    """One JSON line in the reference's schema (scripts/train.py:54-68); returns the dict that was printed."""
    if isinstance(message, dict):
        out = {k: (round(v, 6) if isinstance(v, float) else v) for k, v in message.items()}
        out["type"] = message_type
    else:
        out = {"type": message_type, "message": str(message)}
    print(json.dumps(out), file=stream, flush=True)
    return out


class TrainConfig:
    """The subset of scripts/train.py's arguments (:486-545) that the loop itself consumes, same names and defaults."""

    def __init__(self, epochs: int = 200, learning_rate: float = 1e-4, weight_decay: float = 1e-5, ssim_weight: float = 0.5,
                 perceptual_weight: float = 0.0, vgg_layer_idx: int = 35, perceptual_loss_type: str = "l1",
                 base_filters: int = 32, initial_alpha: float = 0.0, patience: int = 20, checkpoint_dir: str = "./checkpoints",
                 model_type: str = "unet", seed: Optional[int] = None, use_graph: bool = True, precision: str = "bf16",
                 vgg_weights=None, batch_size: Optional[int] = None, augmentation: bool = False, augmentation_params=None):
        if model_type != "unet":
            raise ValueError(f"Unknown model type: {model_type}")             # scripts/train.py:175
        self.__dict__.update(locals())
        del self.__dict__["self"]


def _device_batches(batches: Iterable, device) -> Iterable:
    """CPU batches go through the double-buffered prefetcher (pinned when they are not already); CUDA batches pass through."""
    it = iter(batches)
    try:
        first = next(it)
    except StopIteration:
        return iter(())

    def chain():
        yield first
        yield from it

    if first[0].is_cuda:
        return chain()
    from .utils.prefetch import DevicePrefetcher
    pinned = ((tuple(t if t.is_pinned() else t.pin_memory() for t in b)) for b in chain())
    return DevicePrefetcher(pinned, device)


def _to_float_pairs(low_res, high_res, augmenter):
    """uint8 batches (data.UInt8BatchLoader: low N x H x W, full N x 2H x 2W gray levels) become the float32 N x 1 x . x . pair
    of ``MRISuperResDataset.__getitem__`` ON THE DEVICE: ``augment_pair`` + ``ToTensor`` (utils/dataset.py:122-136) through
    utils/preprocess.py; float batches pass through."""
    if low_res.dtype != torch.uint8:
        return low_res, high_res
    from .utils.preprocess import to_tensor
    if augmenter is not None:
        return augmenter(high_res, low_res)
    return to_tensor(low_res), to_tensor(high_res)


def train(cfg: TrainConfig, train_batches: Callable[[], Iterable], val_batches: Callable[[], Iterable], device=None,
          model: Optional[UNetSuperRes] = None, log_stream=None) -> Dict:
    """Epoch loop of scripts/train.py:278-477.  ``train_batches()`` / ``val_batches()`` return a fresh iterable of
    ``(low_res N x 1 x H x W, high_res N x 1 x 2H x 2W)`` pairs per epoch (a DataLoader, a list, a generator), float32 in [0, 1] -- or uint8 gray levels
    (low N x H x W, full N x 2H x 2W, e.g. data.UInt8BatchLoader): then ToTensor and, with ``cfg.augmentation``, the reference's
    augment_pair run on the device (utils/preprocess.py).
    Returns ``{'final_path', 'best_path', 'history'}``; ``history`` holds the printed epoch summaries.

    Data parallel (new; the reference is single-device): call under ``torchrun`` after ``dist.init_process_group("nccl")``;
    every rank passes its own shard of the batches (equal counts and shapes).  Gradients are exchanged by
    ``parallel.GradientBuckets`` inside the step, epoch sums are averaged over ranks, so all ranks take the same scheduler and
    early-stopping decisions; rank 0 alone logs and writes checkpoints."""
    device = torch.device(device if device is not None else "cuda")
    if device.type != "cuda":
        raise RuntimeError("the B200 training path needs a CUDA device (there is no CPU fallback)")
    if cfg.seed is not None:
        torch.manual_seed(cfg.seed)                                            # :145
    if model is None:
        model = UNetSuperRes(1, 1, cfg.base_filters, cfg.initial_alpha, precision=cfg.precision)
    model = model.to(device)
    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank() if world > 1 else 0
    buckets = None
    if world > 1:
        from .parallel import GradientBuckets
        buckets = GradientBuckets(model)
    if rank != 0:
        log_stream = open(os.devnull, "w")                                     # one JSON-lines stream per job
    optimizer = FusedAdam(model, lr=cfg.learning_rate, weight_decay=cfg.weight_decay, capturable=cfg.use_graph)     # :186
    scheduler = torch.optim.lr_scheduler.ReduceLROnPlateau(optimizer, mode="min", factor=0.5,
                                                           patience=max(cfg.patience // 2, 0))               # :189
    criterion = CombinedLoss(ssim_weight=cfg.ssim_weight, perceptual_weight=cfg.perceptual_weight,
                             vgg_layer_idx=cfg.vgg_layer_idx, perceptual_loss_type=cfg.perceptual_loss_type,
                             window_size=11, sigma=1.5, val_range=1.0, device=device, vgg_weights=cfg.vgg_weights,
                             precision=cfg.precision)                                                      # :236-245
    ssim_metric = SSIM(window_size=11, sigma=1.5, val_range=1.0, device=device)                                      # :248
    os.makedirs(cfg.checkpoint_dir, exist_ok=True)
    log_message({"model_type": cfg.model_type, "batch_size": cfg.batch_size, "epochs": cfg.epochs,
                 "learning_rate": cfg.learning_rate, "weight_decay": cfg.weight_decay, "ssim_weight": cfg.ssim_weight,
                 "perceptual_weight": cfg.perceptual_weight, "initial_alpha": cfg.initial_alpha, "patience": cfg.patience,
                 "seed": cfg.seed}, "params", log_stream)

    best_val_loss, patience_counter = float("inf"), 0
    best_path = os.path.join(cfg.checkpoint_dir, f"best_model_{cfg.model_type}.pth")
    final_path = os.path.join(cfg.checkpoint_dir, f"final_model_{cfg.model_type}.pth")
    graphed, graphed_shape = None, None
    history = []
    sums = torch.zeros(2, dtype=torch.float32, device=device)                  # running (loss, ssim) of the epoch, on the device
    epoch, val_loss, val_ssim = 0, float("inf"), 0.0

    def metric_of(output, high_res):
        # the loss kernel's own mean SSIM of this pair (unclamped, what ssim_metric would recompute), else a forward-only pass
        comp = criterion.last_components
        if cfg.ssim_weight > 0 and comp is not None:
            return comp[1]
        return ssim_metric(output, high_res)

    augmenter = None
    if getattr(cfg, "augmentation", False):
        from .utils.preprocess import PairAugmenter
        augmenter = PairAugmenter(getattr(cfg, "augmentation_params", None))      # draws from `random` like the reference (:142-166)

    for epoch in range(cfg.epochs):
        t0 = time.time()
        model.train()                                                          # :285
        sums.zero_()
        nb = 0
        batches = train_batches()
        if not hasattr(batches, "__len__"):
            batches = list(batches)
        total = len(batches)
        every = max(10, total // 10)
        for batch_idx, (low_res, high_res) in enumerate(_device_batches(batches, device)):
            low_res, high_res = _to_float_pairs(low_res, high_res, augmenter)
            shape = (tuple(low_res.shape), tuple(high_res.shape))
            if cfg.use_graph and graphed is None:
                from .utils.graph import GraphedTrainStep
                graphed = GraphedTrainStep(model, criterion, optimizer, low_res, high_res, record_metric=True, buckets=buckets)
                graphed_shape = shape
            if cfg.use_graph and shape == graphed_shape:
                loss = graphed(low_res, high_res)
                sums += graphed.loss_and_metric
            else:                                                              # eager, or a ragged last batch of the epoch
                optimizer.zero_grad(set_to_none=True)                          # :301
                output = model(low_res)
                loss = criterion(output, high_res)
                loss.backward()
                optimizer.step(grad_scale=buckets.allreduce() if buckets is not None else 1.0)
                sums[0] += loss.detach()
                with torch.no_grad():
                    sums[1] += metric_of(output.detach(), high_res)
            nb += 1
            if batch_idx % every == 0:                                         # :326
                log_message({"epoch": epoch, "batch": batch_idx, "total_batches": total, "loss": float(loss.item())},
                            "batch_update", log_stream)
        if world > 1:
            dist.all_reduce(sums)                                              # equal batch counts per rank: mean of rank means
        train_loss, train_ssim = (sums / (max(nb, 1) * world)).tolist()        # one read-back per epoch

        model.eval()                                                           # :351
        sums.zero_()
        nv = 0
        with torch.no_grad():
            for low_res, high_res in _device_batches(val_batches(), device):
                low_res, high_res = _to_float_pairs(low_res, high_res, None)       # validation is never augmented
                output = model(low_res)
                sums[0] += criterion(output, high_res)
                sums[1] += metric_of(output, high_res)
                nv += 1
        if nv == 0:
            raise ValueError("validation set is empty (scripts/train.py:390 divides by its length)")
        if world > 1:
            dist.all_reduce(sums)
        val_loss, val_ssim = (sums / (nv * world)).tolist()

        prev_lr = optimizer.param_groups[0]["lr"]
        scheduler.step(val_loss)                                               # :397
        cur_lr = optimizer.param_groups[0]["lr"]
        if cur_lr != prev_lr:
            log_message(f"Learning rate adjusted from {prev_lr:.2e} to {cur_lr:.2e}", stream=log_stream)
        if val_loss < best_val_loss:                                           # :405-420
            best_val_loss, patience_counter = val_loss, 0
            if rank == 0:
                torch.save({"epoch": epoch, "model_state_dict": model.state_dict(), "optimizer_state_dict": optimizer.state_dict(),
                            "scheduler_state_dict": scheduler.state_dict(), "val_loss": val_loss, "val_ssim": val_ssim}, best_path)
            log_message(f"Saved best model with validation loss: {val_loss:.6f}", stream=log_stream)
        else:
            patience_counter += 1
        history.append(log_message({"epoch": epoch, "total_epochs": cfg.epochs, "train_loss": train_loss, "val_loss": val_loss,
                                    "train_ssim": train_ssim, "val_ssim": val_ssim, "elapsed": time.time() - t0, "lr": cur_lr},
                                   "epoch_summary", log_stream))
        if patience_counter >= cfg.patience:                                   # :462
            log_message(f"Early stopping triggered after {epoch + 1} epochs", stream=log_stream)
            break

    if rank == 0:
        torch.save({"epoch": epoch, "model_state_dict": model.state_dict(), "optimizer_state_dict": optimizer.state_dict(),
                    "val_loss": val_loss, "val_ssim": val_ssim}, final_path)       # :467-477
    if graphed is not None:
        graphed.close()                    # a recorded step that holds NCCL work must go before the caller destroys the group
    if buckets is not None:
        buckets.close()
    log_message(f"Training completed. Final model saved to {final_path}", stream=log_stream)
    return {"final_path": final_path, "best_path": best_path if os.path.exists(best_path) else None, "history": history,
            "model": model, "optimizer": optimizer}


# ---- inference half (scripts/infer.py) ------------------------------------------------------------------------------
def find_best_checkpoint(checkpoint_dir: str, model_type: str = "unet") -> str:
    """best_model_<type>.pth, else final_model_<type>.pth, else any *.pth naming the type (scripts/infer.py:74-95)."""
    for name in (f"best_model_{model_type}.pth", f"final_model_{model_type}.pth"):
        path = os.path.join(checkpoint_dir, name)
        if os.path.exists(path):
            return path
    for name in sorted(os.listdir(checkpoint_dir)):
        if name.endswith(".pth") and model_type in name:
            return os.path.join(checkpoint_dir, name)
    raise FileNotFoundError(f"No checkpoint found for {model_type} model in {checkpoint_dir}")


def load_model(model_type: str, checkpoint_path: str, device, base_filters: int = 64, precision: str = "bf16") -> UNetSuperRes:
    """scripts/infer.py:41-68: build the model, accept a training checkpoint dict or a bare state_dict (strict), eval()."""
    if model_type != "unet":
        raise ValueError(f"Unknown model type: {model_type}")
    model = UNetSuperRes(in_channels=1, out_channels=1, base_filters=base_filters, precision=precision).to(device)
    ckpt = torch.load(checkpoint_path, map_location=device, weights_only=False)
    state = ckpt["model_state_dict"] if isinstance(ckpt, dict) and "model_state_dict" in ckpt else ckpt
    model.load_state_dict(state)
    model.eval()
    return model


@torch.no_grad()
def super_resolve(model: UNetSuperRes, slices: torch.Tensor, batch: int = 64, rank: int = 0, world: int = 1,
                  use_graph: bool = True) -> torch.Tensor:
    """Whole-stack inference (BASELINE.json configs 4-5; the reference loops one PNG at a time, scripts/infer.py:268-276):
    ``slices`` is ``S x 1 x H x W`` (a volume with the slice axis first).  Rank ``rank`` of ``world`` takes the contiguous
    block ``[S*rank/world, S*(rank+1)/world)`` - slices are independent, so there is no exchange - and runs it
    in chunks of ``batch`` through ``model(x).clamp(0, 1)``; returns that block's output ``S_r x 1 x 2H x 2W`` on the device."""
    if slices.dim() != 4 or slices.shape[1] != 1:
        raise ValueError(f"expected S x 1 x H x W, got {tuple(slices.shape)}")
    dev = model.flat_parameters().device
    s = slices.shape[0]
    lo, hi = (s * rank) // world, (s * (rank + 1)) // world
    mine = slices[lo:hi]
    out = torch.empty((hi - lo, 1, 2 * slices.shape[2], 2 * slices.shape[3]), dtype=torch.float32, device=dev)
    model.eval()
    graphed = None
    for i in range(0, hi - lo, batch):
        x = mine[i:i + batch].to(dev, non_blocking=True).float()
        if use_graph and x.shape[0] == batch and hi - lo >= 3 * batch:        # a graph pays off from the third chunk on
            if graphed is None:
                from .utils.graph import GraphedForward
                graphed = GraphedForward(model, x)
            out[i:i + batch].copy_(graphed(x))
        else:
            out[i:i + x.shape[0]].copy_(torch.clamp(model(x), 0, 1))
    return out
