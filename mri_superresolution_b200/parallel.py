"""Data-parallel gradient exchange: bucketed sum-allreduce of the flat gradient arena, overlapped with backward.

The reference is single-device (no DDP/NCCL anywhere); this is the new
data-parallel layer of BASELINE.json config 3.  One process per GPU; gradients of all 64 parameter
tensors live in ONE contiguous fp32 arena, so a bucket is a slice of it: no packing or unpacking
copies.  The backward pass finishes the conv blocks head -> decoder -> encoder, i.e. from the END of
the arena towards its start, and the C ABI (``mrisr_unet_backward_marked``, include/mrisr.h) records
one CUDA event per bucket on the compute stream the moment every gradient of that bucket is final.
A side stream waits on the event of bucket k and all-reduces it over NCCL while the compute stream
carries on with the encoder half of the backward pass; only the small last bucket (the first two
encoder levels, 0.26 MB at f = 32) is exchanged after backward ends.  1/world_size is folded into the
fused Adam step (``FusedAdam.step(grad_scale=1/world)``).  GroupNorm is per-sample and every loss term
is a batch mean, so mean-of-shard-gradients equals the full-batch gradient exactly.
"""
from __future__ import annotations

import ctypes
from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist

# conv blocks (csrc/unet.cu conv_table index) whose completion closes a bucket: 9 = up1.conv.double_conv.0 (everything from
# the decoder and the head is final: 34% of the arena), 4 = down2.double_conv.0 (down2, down3, up1.up: 62%), 0 = the rest.
SCHEDULE_CUTS = (9, 4, 0)


def bucket_bounds(total: int, n_buckets: int) -> List[Tuple[int, int]]:
    """Contiguous [start, end) element ranges, last bucket first (backward produces the head's gradients first)."""
    n_buckets = max(1, min(n_buckets, total))
    step = -(-total // n_buckets)
    bounds = [(s, min(total, s + step)) for s in range(0, total, step)]
    return bounds[::-1]


def schedule_bounds(total: int, cut_offsets: Sequence[int]) -> List[Tuple[int, int]]:
    """[start, end) ranges from descending arena offsets at which the backward schedule closes a bucket; the last cut is 0."""
    cuts = sorted({int(c) for c in cut_offsets if 0 <= int(c) < total} | {0}, reverse=True)
    bounds, end = [], total
    for c in cuts:
        if c < end:
            bounds.append((c, end))
            end = c
    return bounds


class GradientBuckets:
    def __init__(self, model, group=None, n_buckets: int = 3, overlap: bool = True):
        self.model = model
        self.group = group
        self.n_buckets = n_buckets
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self._stream = None
        self._events: List[int] = []
        self._bounds = None
        flat = model.flat_parameters() if hasattr(model, "flat_parameters") else None
        self.overlap = bool(overlap and self.world > 1 and flat is not None and flat.is_cuda)
        if self.overlap:
            self._arm()

    # ---- overlap with backward: events recorded inside mrisr_unet_backward_marked ----
    def _arm(self) -> None:
        from . import _native
        lib = _native.lib()
        f = self.model.base_filters
        total = int(lib.mrisr_unet_param_elems(f))
        cuts = [int(lib.mrisr_unet_grads_final_from(f, l)) for l in SCHEDULE_CUTS]
        self._bounds = schedule_bounds(total, cuts)
        self._events = []
        for _ in self._bounds:
            ev = lib.mrisr_event_create()
            if not ev:
                _native.check(1, "mrisr_event_create")
            self._events.append(ev)
        self.model.set_gradient_marks([s for s, _ in self._bounds], self._events)

    def close(self) -> None:
        if self._events:
            from . import _native
            self.model.set_gradient_marks(None, None)
            for ev in self._events:
                _native.lib().mrisr_event_destroy(ctypes.c_void_p(ev))
            self._events = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def allreduce(self) -> float:
        """Sum-allreduce the gradient arena in buckets; returns the grad_scale (1/world) for FusedAdam."""
        if self.world == 1:
            return 1.0
        flat = self.model.flat_grads()
        if flat is None:
            raise RuntimeError("gradients are not arena-backed; call backward() through UNetSuperRes first")
        if flat.is_cuda:
            if self._stream is None:
                self._stream = torch.cuda.Stream(device=flat.device)
            marked = self.overlap and self.model.gradient_marks_recorded()
            bounds = self._bounds if marked else bucket_bounds(flat.numel(), self.n_buckets)
            if not marked:
                self._stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._stream):
                for k, (s, e) in enumerate(bounds):
                    if marked:
                        from . import _native
                        _native.check(_native.lib().mrisr_stream_wait_event(
                            ctypes.c_void_p(self._stream.cuda_stream), ctypes.c_void_p(self._events[k])),
                            "mrisr_stream_wait_event")
                    dist.all_reduce(flat[s:e], op=dist.ReduceOp.SUM, group=self.group)
            torch.cuda.current_stream().wait_stream(self._stream)
        else:                                   # gloo / CPU tensors (host-logic tests)
            for s, e in bucket_bounds(flat.numel(), self.n_buckets):
                dist.all_reduce(flat[s:e], op=dist.ReduceOp.SUM, group=self.group)
        return 1.0 / self.world
