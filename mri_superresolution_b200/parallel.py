"""Data-parallel gradient exchange: bucketed sum-allreduce of the flat gradient arena.

The reference is single-device (no DDP/NCCL anywhere); this is the new
data-parallel layer of BASELINE.json config 3.  One process per GPU; gradients of all 64 parameter
tensors live in ONE contiguous fp32 arena, so a bucket is a slice of it: no packing or unpacking
copies.  Buckets are reduced on a side stream so the exchange overlaps whatever the compute stream does
next; 1/world_size is folded into the fused Adam step (``FusedAdam.step(grad_scale=1/world)``).
GroupNorm is per-sample and every loss term is a batch mean, so mean-of-shard-gradients equals the
full-batch gradient exactly.
"""
from __future__ import annotations

from typing import List, Tuple

import torch
import torch.distributed as dist


def bucket_bounds(total: int, n_buckets: int) -> List[Tuple[int, int]]:
    """Contiguous [start, end) element ranges, last bucket first (backward produces the head's gradients first)."""
    n_buckets = max(1, min(n_buckets, total))
    step = -(-total // n_buckets)
    bounds = [(s, min(total, s + step)) for s in range(0, total, step)]
    return bounds[::-1]


class GradientBuckets:
    def __init__(self, model, group=None, n_buckets: int = 3):
        self.model = model
        self.group = group
        self.n_buckets = n_buckets
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self._stream = None

    def allreduce(self) -> float:
        """Sum-allreduce the gradient arena in buckets; returns the grad_scale (1/world) for FusedAdam."""
        if self.world == 1:
            return 1.0
        flat = self.model.flat_grads()
        if flat is None:
            raise RuntimeError("gradients are not arena-backed; call backward() through UNetSuperRes first")
        bounds = bucket_bounds(flat.numel(), self.n_buckets)
        if flat.is_cuda:
            if self._stream is None:
                self._stream = torch.cuda.Stream(device=flat.device)
            self._stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._stream):
                for s, e in bounds:
                    dist.all_reduce(flat[s:e], op=dist.ReduceOp.SUM, group=self.group)
            torch.cuda.current_stream().wait_stream(self._stream)
        else:                                   # gloo / CPU tensors (host-logic tests)
            for s, e in bounds:
                dist.all_reduce(flat[s:e], op=dist.ReduceOp.SUM, group=self.group)
        return 1.0 / self.world
