"""The whole training step (scripts/train.py:296-323: zero_grad, forward, loss, backward, optimizer step) as ONE CUDA-graph
launch.

A step of the B200 path is ~190 kernel launches of 5-700 us each.  Run eagerly with the loss read back every step (as the
reference's loop does, scripts/train.py:321), the host has to enqueue all of them again after every synchronisation and the
GPU idles behind it at the start of each step.  ``GraphedTrainStep`` records the step once (torch.cuda.graph over the
drop-in modules - every C-ABI entry point launches on the stream it is given, allocates nothing and bakes no host state
into a launch, see include/mrisr.h) and replays it: one launch per step, inputs copied into the graph's static buffers.

Requirements: ``FusedAdam(..., capturable=True)`` (device-side step count) and a fixed batch shape.  Data parallel: pass the
``parallel.GradientBuckets`` of the run; its side-stream NCCL all-reduces are forked from / joined to the recorded stream
by the same gradient-readiness events as in the eager step, so the overlap with backward is part of the graph.
"""
from __future__ import annotations

import torch


def _pin_all(*roots):
    """Pins the workspaces of every module under ``roots`` that caches C-ABI workspaces (UNetSuperRes, VGGFeatureExtractor)."""
    pins, seen = [], set()
    for root in roots:
        mods = root.modules() if hasattr(root, "modules") else [root]
        for m in mods:
            if id(m) not in seen and hasattr(m, "pin_workspaces"):
                seen.add(id(m))
                pins.append((m, m.pin_workspaces()))
    return pins


def _unpin_all(pins) -> None:
    for m, token in pins or ():
        m.unpin_workspaces(token)


class GraphedTrainStep:
    def __init__(self, model, criterion, optimizer, x_example: torch.Tensor, t_example: torch.Tensor, warmup: int = 2,
                 buckets=None, record_metric: bool = False):
        """``record_metric``: also keep ``loss_and_metric`` = device tensor [loss, mean SSIM of (output, target)] of the last
        step, the SSIM being the (unclamped) value the loss kernel computed (``criterion.last_components[1]``) - the
        reference runs a second SSIM forward for it every step (scripts/train.py:322-323)."""
        if not getattr(optimizer, "capturable", False):
            raise ValueError("GraphedTrainStep needs FusedAdam(..., capturable=True): the step count must live on the device")
        if not x_example.is_cuda or not t_example.is_cuda:
            raise RuntimeError("GraphedTrainStep records CUDA work: example tensors must be CUDA tensors (no CPU fallback)")
        self.model, self.criterion, self.optimizer, self.buckets = model, criterion, optimizer, buckets
        self.x = x_example.detach().clone()
        self.t = t_example.detach().clone()
        self.loss = None
        self.loss_and_metric = None
        self._record_metric = bool(record_metric)
        dev = self.x.device
        # The warm-up steps (lazy one-time setup inside the library and the allocator) must not count as training:
        # snapshot parameters and optimizer state, restore them after the capture (which itself executes nothing).
        flat = model.flat_parameters()
        optimizer.step_count = int(optimizer.step_count)
        snap = (flat.clone(), optimizer.exp_avg.clone(), optimizer.exp_avg_sq.clone(), optimizer.step_count)
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._eager()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        # the recorded launches hold raw pointers into the workspaces the warm-up steps allocated: pin them so that no other
        # shape (ragged last batch, validation under no_grad) can evict them while the graph lives
        self._pins = _pin_all(model, criterion)
        self.graph = torch.cuda.CUDAGraph()
        optimizer.zero_grad(set_to_none=True)
        with torch.cuda.graph(self.graph):
            self.loss = self._eager()
        with torch.no_grad():
            flat.copy_(snap[0])
            optimizer.exp_avg.copy_(snap[1])
            optimizer.exp_avg_sq.copy_(snap[2])
            optimizer.step_count = snap[3]
            optimizer._device_state(dev)[0] = snap[3]
        torch.cuda.synchronize(dev)

    def _eager(self):
        self.optimizer.zero_grad(set_to_none=True)
        out = self.model(self.x)
        loss = self.criterion(out, self.t)
        loss.backward()
        scale = self.buckets.allreduce() if self.buckets is not None else 1.0
        self.optimizer.step(grad_scale=scale)
        if self._record_metric:
            comp = getattr(self.criterion, "last_components", None)
            ssim = comp[1] if comp is not None else torch.zeros((), device=loss.device)
            self.loss_and_metric = torch.stack([loss.detach(), ssim])
        return loss.detach()

    def close(self) -> None:
        """Destroys the recorded graph.  A graph that captured NCCL all-reduces keeps the communicator busy:
        ``dist.destroy_process_group()`` waits for every such graph to be destroyed first, so data-parallel callers must
        close the step before tearing the group down (bench.py, drivers.train, tools/check_graph_dp.py)."""
        g, self.graph = getattr(self, "graph", None), None
        if g is not None:
            torch.cuda.synchronize(self.x.device)
            g.reset()
        _unpin_all(getattr(self, "_pins", None))
        self._pins = None
        self.loss = None
        self.loss_and_metric = None

    def __del__(self):
        try:
            _unpin_all(getattr(self, "_pins", None))
        except Exception:
            pass

    def __call__(self, x: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
        """One training step on (x, t); returns the loss as a 0-dim device tensor that the NEXT call overwrites."""
        if self.graph is None:
            raise RuntimeError("GraphedTrainStep has been closed")
        if x.shape != self.x.shape or t.shape != self.t.shape:
            raise ValueError(f"GraphedTrainStep was recorded for {tuple(self.x.shape)} / {tuple(self.t.shape)}, "
                             f"got {tuple(x.shape)} / {tuple(t.shape)}")
        self.x.copy_(x, non_blocking=True)
        self.t.copy_(t, non_blocking=True)
        self.optimizer.sync_lr()                     # a scheduler may have changed param_groups[0]['lr'] since the last step
        self.graph.replay()
        self.optimizer.step_count += 1
        return self.loss


class GraphedForward:
    """The inference forward (scripts/infer.py:268-276: eval, no_grad, model(x), clamp(0, 1)) as one CUDA-graph launch for a
    fixed input shape.  Returns the graph's static output tensor (overwritten by the next call)."""

    def __init__(self, model, x_example: torch.Tensor, warmup: int = 2):
        if not x_example.is_cuda:
            raise RuntimeError("GraphedForward records CUDA work: the example tensor must be a CUDA tensor (no CPU fallback)")
        self.model = model
        self.x = x_example.detach().clone()
        dev = self.x.device
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._eager()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self._pins = _pin_all(model)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.y = self._eager()
        torch.cuda.synchronize(dev)

    def _eager(self):
        with torch.no_grad():
            return torch.clamp(self.model(self.x), 0, 1)

    def close(self) -> None:
        g, self.graph = getattr(self, "graph", None), None
        if g is not None:
            torch.cuda.synchronize(self.x.device)
            g.reset()
        _unpin_all(getattr(self, "_pins", None))
        self._pins = None
        self.y = None

    def __del__(self):
        try:
            _unpin_all(getattr(self, "_pins", None))
        except Exception:
            pass

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        if self.graph is None:
            raise RuntimeError("GraphedForward has been closed")
        if x.shape != self.x.shape:
            raise ValueError(f"GraphedForward was recorded for {tuple(self.x.shape)}, got {tuple(x.shape)}")
        self.x.copy_(x, non_blocking=True)
        self.graph.replay()
        return self.y
