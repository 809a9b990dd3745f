"""Fused Adam over the flat parameter arena of UNetSuperRes (replaces ``optim.Adam(model.parameters(),
lr, weight_decay)`` + ``optimizer.step()`` of scripts/train.py:186,318 with one kernel launch)."""
from __future__ import annotations

import torch

from . import _native


class FusedAdam:
    """torch.optim.Adam semantics (L2 weight decay added to the gradient, bias correction), one launch.

    ``grad_scale`` multiplies the gradient first: 1/world_size after a sum-allreduce, 1/loss_scale under AMP.
    """

    def __init__(self, model, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, capturable: bool = False):
        self.model = model
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        flat = model.flat_parameters()
        self.exp_avg = torch.zeros_like(flat)
        self.exp_avg_sq = torch.zeros_like(flat)
        self.step_count = 0
        # capturable: the step number lives on the device (mrisr_adam_step_dev), so step() can be recorded into a CUDA
        # graph (utils/graph.py) without freezing the bias correction; same arithmetic as the host-stepped form
        self.capturable = bool(capturable)
        self._dev_state = None

    def _device_state(self, device) -> torch.Tensor:
        """16 bytes: int64 step count + two floats of scratch; re-synchronised from ``step_count`` when (re)created."""
        if self._dev_state is None or self._dev_state.device != device:
            st = torch.zeros(2, dtype=torch.int64, device=device)
            st[0] = self.step_count
            self._dev_state = st
        return self._dev_state

    def zero_grad(self, set_to_none: bool = True):
        for p in self.model.parameters():
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()

    @torch.no_grad()
    def step(self, grad_scale: float = 1.0):
        flat = self.model.flat_parameters()
        if self.exp_avg.device != flat.device:
            self.exp_avg = self.exp_avg.to(flat.device)
            self.exp_avg_sq = self.exp_avg_sq.to(flat.device)
        grads = self.model.flat_grads()
        if grads is None:                      # .grad tensors are not arena views (e.g. user-assigned): gather them
            grads = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1)
                               for p in self.model._param_list()])
        if self.capturable:
            st = self._device_state(flat.device)      # created from the count BEFORE this step; the kernel advances it
            self.step_count += 1
            rc = _native.lib().mrisr_adam_step_dev(_native.ptr(flat), _native.ptr(grads), _native.ptr(self.exp_avg),
                                                   _native.ptr(self.exp_avg_sq), flat.numel(), self.lr, self.betas[0],
                                                   self.betas[1], self.eps, self.weight_decay, _native.ptr(st),
                                                   float(grad_scale), _native.stream_ptr())
            _native.check(rc, "mrisr_adam_step_dev")
            return
        self.step_count += 1
        rc = _native.lib().mrisr_adam_step(_native.ptr(flat), _native.ptr(grads), _native.ptr(self.exp_avg),
                                           _native.ptr(self.exp_avg_sq), flat.numel(), self.lr, self.betas[0],
                                           self.betas[1], self.eps, self.weight_decay, self.step_count,
                                           float(grad_scale), _native.stream_ptr())
        _native.check(rc, "mrisr_adam_step")

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq,
                "lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": self.weight_decay}

    def load_state_dict(self, sd):
        self.step_count = int(sd["step"])
        if self._dev_state is not None:
            self._dev_state[0] = self.step_count
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.lr, self.betas, self.eps, self.weight_decay = sd["lr"], tuple(sd["betas"]), sd["eps"], sd["weight_decay"]
