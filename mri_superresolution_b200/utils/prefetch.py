"""Double-buffered host -> device staging of training batches.

The reference's training loop moves every batch with ``.to(device)`` at the top of the step (scripts/train.py:296-297, the
DataLoader is built with ``pin_memory=True``, :193-200), so the 84 MB copy of a 64-slice batch and the step serialise.
``DevicePrefetcher`` wraps any iterable of tuples of (pinned) CPU tensors and yields tuples of CUDA tensors: the copy of
batch i+1 runs on a side stream while step i computes, with two device-side slots and events in both directions (a slot
is overwritten only after the work queued on the consumer's stream when it moved on has finished).

Every batch is still copied host -> device exactly once; only the placement in time changes.
"""
from __future__ import annotations

from typing import Iterable, Iterator, List, Optional, Sequence, Tuple

import torch


class DevicePrefetcher:
    def __init__(self, source: Iterable[Sequence[torch.Tensor]], device, slots: int = 2):
        if slots < 2:
            raise ValueError("DevicePrefetcher needs at least two slots")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("DevicePrefetcher stages batches for the CUDA kernels; device must be a CUDA device")
        self._src: Iterator = iter(source)
        self._n = slots
        self._stream = torch.cuda.Stream(self.device)
        self._bufs: List[Optional[Tuple[torch.Tensor, ...]]] = [None] * slots
        self._ready: List[Optional[torch.cuda.Event]] = [None] * slots
        self._free: List[Optional[torch.cuda.Event]] = [None] * slots
        self._head = 0          # batches staged so far
        self._tail = 0          # batches handed out so far
        self._prev: Optional[int] = None
        self._exhausted = False
        self.bytes_copied = 0

    def _fill(self) -> None:
        in_use = self._head - self._tail + (0 if self._prev is None else 1)     # staged + the slot the consumer holds
        if self._exhausted or in_use >= self._n:
            return
        try:
            batch = next(self._src)
        except StopIteration:
            self._exhausted = True
            return
        i = self._head % self._n
        bufs = self._bufs[i]
        if bufs is None or len(bufs) != len(batch) or any(b.shape != t.shape or b.dtype != t.dtype for b, t in zip(bufs, batch)):
            # allocated on the consumer's stream (never freed while the prefetcher lives), filled on the side stream
            bufs = tuple(torch.empty(t.shape, dtype=t.dtype, device=self.device) for t in batch)
            self._bufs[i] = bufs
            self._stream.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self._stream):
            if self._free[i] is not None:
                self._stream.wait_event(self._free[i])
            for dst, src in zip(bufs, batch):
                dst.copy_(src, non_blocking=True)
                self.bytes_copied += src.numel() * src.element_size()
            ev = torch.cuda.Event()
            ev.record(self._stream)
            self._ready[i] = ev
        self._head += 1

    def __iter__(self):
        return self

    def __next__(self) -> Tuple[torch.Tensor, ...]:
        if self._tail == self._head:
            self._fill()
            if self._tail == self._head:
                raise StopIteration
        i = self._tail % self._n
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self._ready[i])
        if self._prev is not None:
            # everything queued on the consumer's stream so far (the step that used the previous slot) guards that slot
            ev = torch.cuda.Event()
            ev.record(cur)
            self._free[self._prev] = ev
        self._prev = i
        self._tail += 1
        self._fill()
        return self._bufs[i]
