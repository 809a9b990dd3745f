"""Fused Adam over the flat parameter arena of UNetSuperRes (replaces ``optim.Adam(model.parameters(),
lr, weight_decay)`` + ``optimizer.step()`` of scripts/train.py:186,318 with one kernel launch).

``FusedAdam`` is a ``torch.optim.Optimizer``: ``param_groups[0]['lr']`` is what ``ReduceLROnPlateau``
(scripts/train.py:189,397) adjusts, and ``state_dict()`` / ``load_state_dict()`` use the stock ``torch.optim.Adam``
layout (per-parameter ``step`` / ``exp_avg`` / ``exp_avg_sq``, SURVEY.md 8(b) "Optimizer state"), so the
``optimizer_state_dict`` of a reference checkpoint (scripts/train.py:411-418) loads here and ours loads there.
Internally both moments are ONE flat fp32 arena each; the per-parameter state tensors are views of them.
"""
from __future__ import annotations

import torch

from . import _native


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam semantics (L2 weight decay added to the gradient, bias correction), one launch.

    ``grad_scale`` multiplies the gradient first: 1/world_size after a sum-allreduce, 1/loss_scale under AMP.
    ``capturable=True`` keeps the step count and the learning rate on the device (``mrisr_adam_step_dev``), so ``step()``
    can be recorded into a CUDA graph (utils/graph.py) without freezing either; same arithmetic as the host-stepped form.
    """

    def __init__(self, model, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, capturable: bool = False):
        self.model = model
        flat = model.flat_parameters()
        defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay, amsgrad=False, maximize=False,
                        foreach=None, capturable=False, differentiable=False, fused=None)
        super().__init__(model._param_list(), defaults)       # arena order == registration order == state_dict order
        self.exp_avg = torch.zeros_like(flat)
        self.exp_avg_sq = torch.zeros_like(flat)
        self.step_count = 0
        self.capturable = bool(capturable)
        self._dev_state = None
        self._dev_lr = None

    # hyper-parameters live in param_groups[0] (one group: the whole arena)
    lr = property(lambda self: self.param_groups[0]["lr"], lambda self, v: self.param_groups[0].__setitem__("lr", v))
    betas = property(lambda self: tuple(self.param_groups[0]["betas"]))
    eps = property(lambda self: self.param_groups[0]["eps"])
    weight_decay = property(lambda self: self.param_groups[0]["weight_decay"])

    def _device_state(self, device) -> torch.Tensor:
        """32 bytes: int64 step count, two floats of scratch, float lr; re-synchronised from the host values when (re)created."""
        if self._dev_state is None or self._dev_state.device != device:
            st = torch.zeros(4, dtype=torch.int64, device=device)
            st[0] = self.step_count
            self._dev_state = st
            self._dev_lr = None
        self.sync_lr()
        return self._dev_state

    def sync_lr(self) -> None:
        """Push ``param_groups[0]['lr']`` to the device state if it changed (called by step() and by GraphedTrainStep)."""
        if self._dev_state is not None and self._dev_lr != self.lr:
            self._dev_state.view(torch.float32)[4] = float(self.lr)
            self._dev_lr = self.lr

    def zero_grad(self, set_to_none: bool = True):
        for p in self.model.parameters():
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()

    @torch.no_grad()
    def step(self, closure=None, grad_scale: float = 1.0):
        if closure is not None:
            raise ValueError("FusedAdam.step does not take a closure")
        flat = self.model.flat_parameters()
        if self.exp_avg.device != flat.device:
            self.exp_avg = self.exp_avg.to(flat.device)
            self.exp_avg_sq = self.exp_avg_sq.to(flat.device)
        grads = self.model.flat_grads()
        if grads is None:                      # .grad tensors are not arena views (e.g. user-assigned): gather them
            grads = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1)
                               for p in self.model._param_list()])
        b1, b2 = self.betas
        if self.capturable:
            st = self._device_state(flat.device)      # created from the count BEFORE this step; the kernel advances it
            self.step_count += 1
            rc = _native.lib().mrisr_adam_step_dev(_native.ptr(flat), _native.ptr(grads), _native.ptr(self.exp_avg),
                                                   _native.ptr(self.exp_avg_sq), flat.numel(), b1, b2, self.eps,
                                                   self.weight_decay, _native.ptr(st), float(grad_scale),
                                                   _native.stream_ptr())
            _native.check(rc, "mrisr_adam_step_dev")
            return
        self.step_count += 1
        rc = _native.lib().mrisr_adam_step(_native.ptr(flat), _native.ptr(grads), _native.ptr(self.exp_avg),
                                           _native.ptr(self.exp_avg_sq), flat.numel(), self.lr, b1, b2, self.eps,
                                           self.weight_decay, self.step_count, float(grad_scale), _native.stream_ptr())
        _native.check(rc, "mrisr_adam_step")

    # ---- stock torch.optim.Adam checkpoint layout -------------------------------------------------------
    def state_dict(self):
        """``{'state': {i: {'step', 'exp_avg', 'exp_avg_sq'}}, 'param_groups': [...]}`` as torch.optim.Adam writes it
        (empty ``state`` before the first step, as in torch)."""
        params = self.model._param_list()
        offs = self.model._offs
        state = {}
        if self.step_count > 0:
            for i, (p, o) in enumerate(zip(params, offs)):
                n = p.numel()
                state[i] = {"step": torch.tensor(float(self.step_count)),
                            "exp_avg": self.exp_avg[o:o + n].view(p.shape).clone(),
                            "exp_avg_sq": self.exp_avg_sq[o:o + n].view(p.shape).clone()}
        group = {k: v for k, v in self.param_groups[0].items() if k != "params"}
        group["params"] = list(range(len(params)))
        return {"state": state, "param_groups": [group]}

    @torch.no_grad()
    def load_state_dict(self, sd):
        params = self.model._param_list()
        offs = self.model._offs
        groups = sd["param_groups"]
        if len(groups) != 1 or len(groups[0]["params"]) != len(params):
            raise ValueError("FusedAdam.load_state_dict: expected one parameter group covering all "
                             f"{len(params)} parameters (torch.optim.Adam(model.parameters()))")
        for k, v in groups[0].items():
            if k != "params":
                self.param_groups[0][k] = tuple(v) if k == "betas" else v
        state = sd["state"]
        steps = set()
        self.exp_avg.zero_()
        self.exp_avg_sq.zero_()
        for i, (p, o) in enumerate(zip(params, offs)):
            s = state.get(i, state.get(str(i)))
            if s is None:
                continue
            n = p.numel()
            self.exp_avg[o:o + n].copy_(s["exp_avg"].reshape(-1))
            self.exp_avg_sq[o:o + n].copy_(s["exp_avg_sq"].reshape(-1))
            steps.add(int(float(s["step"])))
        if len(steps) > 1:
            raise ValueError(f"FusedAdam.load_state_dict: parameters disagree on the step count ({sorted(steps)})")
        self.step_count = steps.pop() if steps else 0
        if self._dev_state is not None:
            self._dev_state[0] = self.step_count
            self._dev_lr = None
            self.sync_lr()
