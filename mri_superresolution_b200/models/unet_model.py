"""Drop-in ``UNetSuperRes`` backed by the sm_100a kernels of libmrisr_b200.so.

Mirrors the public surface of the reference module (models/unet_model.py:116-211): constructor
signature, attributes ``in_channels / out_channels / base_filters``, the ``alpha`` parameter, and --
the checkpoint contract -- exactly the reference's 64 ``state_dict`` keys, shapes and registration
order (scripts/train.py:411-418 saves them, scripts/infer.py:56-65 loads them strictly), so
reference checkpoints load here and checkpoints written here load in the reference.

The sub-module tree exists only to carry those names; all arithmetic is done by
``mrisr_unet_forward`` / ``mrisr_unet_backward`` on a flat fp32 parameter arena that the
``nn.Parameter``s are views of.  There is no PyTorch/CPU fallback.
"""
from __future__ import annotations

import math
import os
from typing import Dict, List, Tuple

import torch
import torch.nn as nn

from .. import _native
from ..utils.workspace import WorkspaceCache

__all__ = ["UNetSuperRes", "param_specs"]


def param_specs(base_filters: int) -> List[Tuple[str, Tuple[int, ...]]]:
    """(state_dict key, shape) in the reference's registration order, models/unet_model.py:129-175."""
    f = base_filters
    out: List[Tuple[str, Tuple[int, ...]]] = [("alpha", ())]

    def dc(prefix: str, cin: int, cout: int) -> None:
        for idx, (ci, co) in (("0", (cin, cout)), ("3", (cout, cout))):
            out.append((f"{prefix}.double_conv.{idx}.weight", (co, ci, 3, 3)))
            n = str(int(idx) + 1)
            out.append((f"{prefix}.double_conv.{n}.weight", (co,)))
            out.append((f"{prefix}.double_conv.{n}.bias", (co,)))

    dc("inc", 1, f)
    dc("down1.maxpool_conv.1", f, 2 * f)
    dc("down2.maxpool_conv.1", 2 * f, 4 * f)
    dc("down3.maxpool_conv.1", 4 * f, 8 * f)
    for name, cup, cskip, cout in (("up1", 8 * f, 4 * f, 4 * f), ("up2", 4 * f, 2 * f, 2 * f), ("up3", 2 * f, f, f)):
        out.append((f"{name}.up.1.weight", (cup // 2, cup, 1, 1)))
        out.append((f"{name}.up.2.weight", (cup // 2,)))
        out.append((f"{name}.up.2.bias", (cup // 2,)))
        dc(f"{name}.conv", cskip + cup // 2, cout)
    out += [("final_up_bilinear.1.weight", (f // 2, f, 3, 3)),
            ("final_up_bilinear.2.weight", (f // 2,)), ("final_up_bilinear.2.bias", (f // 2,)),
            ("final_up_pixelshuffle.conv.weight", (2 * f, f, 3, 3)), ("final_up_pixelshuffle.conv.bias", (2 * f,)),
            ("final_up_pixelshuffle.norm.weight", (f // 2,)), ("final_up_pixelshuffle.norm.bias", (f // 2,)),
            ("final_conv.0.weight", (f // 2, f // 2, 3, 3)),
            ("final_conv.1.weight", (f // 2,)), ("final_conv.1.bias", (f // 2,)),
            ("final_conv.3.weight", (1, f // 2, 1, 1)), ("final_conv.3.bias", (1,))]
    return out


class _Node(nn.Module):
    """Name-carrying container (no forward of its own)."""


def _precision_code(precision: str) -> int:
    if precision == "bf16":
        return _native.BF16
    if precision == "fp32":
        return _native.FP32
    raise ValueError(f"precision must be 'bf16' or 'fp32', got {precision!r}")


class _UNetFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, module, *params):
        y = module._run_forward(x, train=True)
        ctx.module = module
        ctx.ticket = module._ticket
        ctx.x_shape = tuple(x.shape)
        ctx.save_for_backward(y)
        ctx.mark_non_differentiable()
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        grads = ctx.module._run_backward(dy, y, ctx.ticket, ctx.x_shape)
        return (None, None) + tuple(grads)


class UNetSuperRes(nn.Module):
    """2x super-resolution U-Net; same constructor as the reference (models/unet_model.py:129) plus
    ``precision`` ('bf16': tcgen05 tensor-core path, 'fp32': CUDA-core parity path)."""

    def __init__(self, in_channels: int = 1, out_channels: int = 1, base_filters: int = 32,
                 initial_alpha: float = 0.0, precision: str | None = None):
        super().__init__()
        if in_channels != 1 or out_channels != 1:
            raise ValueError("the B200 path implements the reference's single-channel configuration "
                             "(in_channels=1, out_channels=1; scripts/train.py:168-173)")
        if base_filters < 16 or base_filters % 16:
            raise ValueError("base_filters must be a positive multiple of 16 (GroupNorm(8) on base_filters/2 channels)")
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.base_filters = base_filters
        self.precision = precision or os.environ.get("MRISR_PRECISION", "bf16")
        _precision_code(self.precision)
        self._specs = param_specs(base_filters)
        for name, shape in self._specs:
            parts = name.split(".")
            node = self
            for part in parts[:-1]:
                if part not in node._modules:
                    node.add_module(part, _Node())
                node = node._modules[part]
            node.register_parameter(parts[-1], nn.Parameter(torch.empty(shape, dtype=torch.float32)))
        self._flat = None
        self._grad_arenas: List[torch.Tensor] = []
        self._workspaces = WorkspaceCache()
        self._ticket = 0
        self._last_train_key = None
        self._marks = None                 # (int64 offsets array, void* events array, n) set by parallel.GradientBuckets
        self._marks_arena_ptr = 0
        self._initialize_weights(initial_alpha)
        self._flatten()

    # ---- initialisation: same distributions as models/unet_model.py:177-187 --------------------------
    def _initialize_weights(self, initial_alpha: float) -> None:
        with torch.no_grad():
            for name, p in self.named_parameters():
                if name == "alpha":
                    p.fill_(initial_alpha / 100.0)                      # :162-163
                elif p.dim() == 4:
                    nn.init.kaiming_normal_(p, mode="fan_out", nonlinearity="leaky_relu")
                elif name.endswith("conv.bias") or name == "final_conv.3.bias":
                    p.zero_()
                elif name.endswith("weight"):
                    p.fill_(1.0)
                else:
                    p.zero_()

    # ---- flat parameter arena ------------------------------------------------------------------------
    def _param_list(self) -> List[nn.Parameter]:
        return [self.get_parameter(n) for n, _ in self._specs]

    def _offsets(self) -> List[int]:
        offs, o = [], 0
        for _, shape in self._specs:
            offs.append(o)
            o += int(math.prod(shape)) if shape else 1
        self._total = o
        return offs

    def _is_flat(self) -> bool:
        if self._flat is None:
            return False
        base = self._flat.data_ptr()
        for p, o in zip(self._param_list(), self._offs):
            if p.data_ptr() != base + 4 * o or p.dtype != torch.float32:
                return False
        return True

    def _flatten(self) -> None:
        """(Re)build the arena and make every parameter a view of it (values preserved)."""
        params = self._param_list()
        self._offs = self._offsets()
        dev = params[0].device
        flat = torch.empty(self._total, dtype=torch.float32, device=dev)
        for p, o in zip(params, self._offs):
            n = p.numel()
            flat[o:o + n].copy_(p.detach().reshape(-1).to(torch.float32))
            p.data = flat[o:o + n].view(p.shape)
        self._flat = flat
        self._grad_arenas = []
        self._workspaces.clear()

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        self._flatten()
        return out

    def flat_parameters(self) -> torch.Tensor:
        """The fp32 arena all parameters are views of (registration order)."""
        if not self._is_flat():
            self._flatten()
        return self._flat

    def flat_grads(self):
        """The arena holding the gradients of the last backward, or None if ``.grad`` is not arena-backed."""
        for arena in self._grad_arenas:
            base = arena.data_ptr()
            if all(p.grad is not None and p.grad.data_ptr() == base + 4 * o
                   for p, o in zip(self._param_list(), self._offs)):
                return arena
        return None

    # ---- kernels -------------------------------------------------------------------------------------
    def _workspace(self, n: int, h: int, w: int, train: bool) -> torch.Tensor:
        prec = _precision_code(self.precision)
        key = (n, h, w, prec, train, self._flat.device)

        def alloc():
            nbytes = _native.lib().mrisr_unet_workspace_bytes(n, h, w, self.base_filters, prec, int(train))
            if nbytes == 0:
                _native.check(-2, "mrisr_unet_workspace_bytes")
            return torch.zeros(nbytes, dtype=torch.uint8, device=self._flat.device)   # zero borders: one-time invariant

        # other un-pinned shapes of the same kind are dropped beyond the two most recent (a workspace is GBs); entries a
        # recorded CUDA graph points into are pinned and never dropped (utils/workspace.py)
        return self._workspaces.get(key, bool(train), alloc)

    def pin_workspaces(self):
        """Keeps every workspace allocated so far alive until ``unpin_workspaces(token)`` (utils/graph.py: a recorded step
        holds raw pointers into them)."""
        return self._workspaces.pin_all()

    def unpin_workspaces(self, token) -> None:
        self._workspaces.unpin(token)

    def _run_forward(self, x: torch.Tensor, train: bool) -> torch.Tensor:
        if x.dim() != 4 or x.shape[1] != 1:
            raise ValueError(f"expected input of shape N x 1 x H x W, got {tuple(x.shape)}")
        if not x.is_cuda:
            raise RuntimeError("UNetSuperRes (B200) needs CUDA tensors: there is no CPU fallback")
        params = self.flat_parameters()
        if params.device != x.device:
            raise RuntimeError(f"input on {x.device} but parameters on {params.device}")
        n, _, h, w = x.shape
        x32 = x.detach().to(torch.float32).contiguous()
        y = torch.empty((n, 1, 2 * h, 2 * w), dtype=torch.float32, device=x.device)
        ws = self._workspace(n, h, w, train)
        rc = _native.lib().mrisr_unet_forward(_native.ptr(params), _native.ptr(x32), _native.ptr(y), _native.ptr(ws),
                                              ws.numel(), n, h, w, self.base_filters,
                                              _precision_code(self.precision), int(train), _native.stream_ptr())
        _native.check(rc, "mrisr_unet_forward")
        if train:
            self._ticket += 1
            self._last_train_key = (n, h, w)
        return y

    def _pick_grad_arena(self) -> torch.Tensor:
        """An arena no live ``.grad`` aliases (so accumulation across backward calls stays correct)."""
        live = {p.grad.untyped_storage().data_ptr() for p in self._param_list() if p.grad is not None}
        for arena in self._grad_arenas:
            if arena.untyped_storage().data_ptr() not in live:
                return arena
        arena = torch.empty(self._total, dtype=torch.float32, device=self._flat.device)
        self._grad_arenas.append(arena)
        return arena

    def _run_backward(self, dy: torch.Tensor, y: torch.Tensor, ticket: int, x_shape):
        if ticket != self._ticket:
            raise RuntimeError("UNetSuperRes.backward: a newer forward pass has overwritten the saved activations "
                               "(one outstanding training forward per module)")
        n, _, h, w = x_shape
        ws = self._workspace(n, h, w, True)
        arena = self._pick_grad_arena()
        dy32 = dy.to(torch.float32).contiguous()
        if self._marks is not None:
            offs, evs, nm = self._marks
            rc = _native.lib().mrisr_unet_backward_marked(_native.ptr(self._flat), _native.ptr(dy32), _native.ptr(y),
                                                          _native.ptr(arena), None, _native.ptr(ws), ws.numel(), n, h, w,
                                                          self.base_filters, _precision_code(self.precision),
                                                          _native.stream_ptr(), nm, offs, evs)
            _native.check(rc, "mrisr_unet_backward_marked")
            self._marks_arena_ptr = arena.data_ptr()
        else:
            rc = _native.lib().mrisr_unet_backward(_native.ptr(self._flat), _native.ptr(dy32), _native.ptr(y),
                                                   _native.ptr(arena), None, _native.ptr(ws), ws.numel(), n, h, w,
                                                   self.base_filters, _precision_code(self.precision),
                                                   _native.stream_ptr())
            _native.check(rc, "mrisr_unet_backward")
        return [arena[o:o + p.numel()].view(p.shape) for p, o in zip(self._param_list(), self._offs)]

    # ---- gradient-readiness marks (data parallel: parallel.GradientBuckets) -------------------------
    def set_gradient_marks(self, offsets, events) -> None:
        """Ask every following backward pass to record ``events[k]`` (library-owned CUDA events) on the compute stream as
        soon as all gradients at arena offset >= ``offsets[k]`` are final (include/mrisr.h, mrisr_unet_backward_marked)."""
        if not offsets:
            self._marks = None
            self._marks_arena_ptr = 0
            return
        if len(offsets) != len(events) or len(offsets) > 32:
            raise ValueError("gradient marks: offsets and events must have the same length (<= 32)")
        import ctypes
        n = len(offsets)
        self._marks = ((ctypes.c_int64 * n)(*[int(o) for o in offsets]), (ctypes.c_void_p * n)(*events), n)

    def gradient_marks_recorded(self) -> bool:
        """True when ``.grad`` is exactly the arena the last marked backward wrote (no accumulation on top of it)."""
        flat = self.flat_grads()
        return self._marks is not None and flat is not None and flat.data_ptr() == self._marks_arena_ptr

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        needs_grad = torch.is_grad_enabled() and any(p.requires_grad for p in self._param_list())
        if needs_grad:
            self.flat_parameters()
            return _UNetFunction.apply(x, self, *self._param_list())
        return self._run_forward(x, train=False)
