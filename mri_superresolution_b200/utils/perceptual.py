"""VGG perceptual term of CombinedLoss on the tcgen05 convolution kernels.

Mirrors the reference's ``VGGFeatureExtractor`` (utils/losses.py:83-118) and ``PerceptualLoss`` (:120-151): same
constructor keywords, ``feature_extractor.features`` (an ``nn.Sequential`` of the torchvision VGG19 layers up to
``feature_layer_idx``, frozen), ``mean`` / ``std`` buffers, ``criterion`` and the ``ValueError`` for an unknown loss
type (:133).  The computation itself -- both feature passes, the feature L1 / MSE and the data gradient of the
generated branch -- is one C-ABI call (``mrisr_vgg_perceptual_fwd_bwd``): bf16 operands, fp32 accumulation, the ReLU
and its backward fused into the convolution epilogues.  ``precision='fp32'`` (or MRISR_PRECISION=fp32) runs the same
term with fp32 activations and weights on CUDA cores (``mrisr_vgg_perceptual_fwd_bwd_fp32``): the parity mode the 1e-4
bound of the north star is tested in, sized for test images.  No CPU fallback.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn as nn

from .. import _native
from .workspace import WorkspaceCache

VGG_MEAN = [0.485, 0.456, 0.406]        # utils/losses.py:7
VGG_STD = [0.229, 0.224, 0.225]         # utils/losses.py:8
# torchvision vgg19 "E" configuration (features indices 0..36)
VGG19_CFG = [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512, "M"]


def _vgg19_layers(feature_layer_idx: int) -> List:
    """Layer kinds of vgg19.features[:feature_layer_idx + 1] as ('conv', cout) / ('relu',) / ('pool',)."""
    layers = []
    for item in VGG19_CFG:
        if item == "M":
            layers.append(("pool",))
        else:
            layers.append(("conv", item))
            layers.append(("relu",))
    return layers[: feature_layer_idx + 1]


class VGGFeatureExtractor(nn.Module):
    """Frozen VGG19 prefix.  ``weights``: optional sequence of (weight, bias) pairs for the convolutions (tests, or a
    locally stored checkpoint); None loads torchvision's ImageNet weights exactly as the reference does
    (utils/losses.py:90), which needs the weight file in the torch hub cache."""

    def __init__(self, feature_layer_idx: int = 35, use_maxpool: bool = False,
                 weights: Optional[Sequence[Tuple[torch.Tensor, torch.Tensor]]] = None, precision: Optional[str] = None):
        super().__init__()
        self.precision = precision or os.environ.get("MRISR_PRECISION", "bf16")
        if self.precision not in ("bf16", "fp32"):
            raise ValueError(f"precision must be 'bf16' or 'fp32', got {self.precision!r}")
        kinds = _vgg19_layers(feature_layer_idx)
        if not kinds or kinds[-1][0] != "relu":
            raise ValueError("feature_layer_idx must select a ReLU layer of vgg19.features (e.g. 35 = relu5_4); the fused "
                             "kernels end the feature stack on conv + ReLU")
        if weights is None:
            import torchvision.models as models
            vgg = models.vgg19(weights=models.VGG19_Weights.IMAGENET1K_V1).features
            self.features = nn.Sequential(*list(vgg.children())[: feature_layer_idx + 1])
        else:
            mods, cin, it = [], 3, iter(weights)
            for k in kinds:
                if k[0] == "conv":
                    conv = nn.Conv2d(cin, k[1], 3, padding=1)
                    w, b = next(it)
                    with torch.no_grad():
                        conv.weight.copy_(w)
                        conv.bias.copy_(b)
                    mods.append(conv)
                    cin = k[1]
                elif k[0] == "relu":
                    mods.append(nn.ReLU(inplace=True))
                else:
                    mods.append(nn.MaxPool2d(2, 2))
            self.features = nn.Sequential(*mods)
        for p in self.features.parameters():                      # utils/losses.py:98-99
            p.requires_grad = False
        self.register_buffer("mean", torch.tensor(VGG_MEAN).view(1, 3, 1, 1))
        self.register_buffer("std", torch.tensor(VGG_STD).view(1, 3, 1, 1))
        self.cfg = [k[1] if k[0] == "conv" else -1 for k in kinds if k[0] != "relu"]
        self._packed = None            # (device, flat fp32 weights, packed bf16 operands)
        self._ws = WorkspaceCache()

    # ---- native plumbing ------------------------------------------------------------------
    def _cfg_array(self):
        arr = (ctypes.c_int * len(self.cfg))(*self.cfg)
        return arr, len(self.cfg)

    def _operands(self, device):
        """Flat fp32 weight arena + packed bf16 copies on `device` (built once: the weights are frozen)."""
        if self._packed is not None and self._packed[0] == device:
            return self._packed[1], self._packed[2]
        L = _native.lib()
        arr, n = self._cfg_array()
        flat = torch.cat([t.detach().reshape(-1).to(torch.float32) for m in self.features if isinstance(m, nn.Conv2d)
                          for t in (m.weight, m.bias)]).to(device).contiguous()
        assert flat.numel() == L.mrisr_vgg_weight_elems(arr, n)
        packed = torch.empty(L.mrisr_vgg_packed_bytes(arr, n), dtype=torch.uint8, device=device)
        rc = L.mrisr_vgg_pack_weights(_native.ptr(flat), arr, n, _native.ptr(packed), packed.numel(), _native.stream_ptr())
        _native.check(rc, "mrisr_vgg_pack_weights")
        self._packed = (device, flat, packed)
        return flat, packed

    def _workspace(self, n, h, w, grad, device):
        key = (n, h, w, bool(grad), device, self.precision)

        def alloc():
            arr, nc = self._cfg_array()
            size_fn = _native.lib().mrisr_vgg_workspace_bytes_fp32 if self.precision == "fp32" else _native.lib().mrisr_vgg_workspace_bytes
            nbytes = size_fn(n, h, w, arr, nc, 1 if grad else 0)
            if nbytes == 0:
                raise ValueError(_native.lib().mrisr_last_error().decode())
            return torch.zeros(nbytes, dtype=torch.uint8, device=device)     # zero borders are an invariant of the kernels

        # grad / no-grad workspaces coexist (training step vs validation pass); entries a recorded CUDA graph points into
        # are pinned and never dropped (utils/workspace.py)
        return self._ws.get(key, bool(grad), alloc)

    def pin_workspaces(self):
        return self._ws.pin_all()

    def unpin_workspaces(self, token) -> None:
        self._ws.unpin(token)

    @staticmethod
    def _gray(x: torch.Tensor, name: str) -> torch.Tensor:
        if not x.is_cuda:
            raise RuntimeError(f"{name}: the B200 perceptual kernels need CUDA tensors (no CPU fallback)")
        if x.dim() != 4 or x.shape[1] != 1:
            raise ValueError(f"{name}: expected N x 1 x H x W (the path trains on gray slices), got {tuple(x.shape)}")
        return x.detach().to(torch.float32).contiguous()

    def forward(self, x):
        """N x 1 x H x W in [0,1] -> N x C x H/16 x W/16 features (fp32 copy of the bf16 activations)."""
        img = self._gray(x, "x")
        n, _, h, w = img.shape
        flat, packed = self._operands(img.device)
        ws = self._workspace(n, h, w, False, img.device)
        arr, nc = self._cfg_array()
        pools = sum(1 for c in self.cfg if c < 0)
        cout = [c for c in self.cfg if c > 0][-1]
        feat = torch.empty(n, cout, h >> pools, w >> pools, dtype=torch.float32, device=img.device)
        if self.precision == "fp32":
            rc = _native.lib().mrisr_vgg_features_fp32(_native.ptr(flat), _native.ptr(img), _native.ptr(feat), _native.ptr(ws),
                                                       ws.numel(), n, h, w, arr, nc, _native.stream_ptr())
        else:
            rc = _native.lib().mrisr_vgg_features(_native.ptr(packed), _native.ptr(flat), _native.ptr(img), _native.ptr(feat),
                                                  _native.ptr(ws), ws.numel(), n, h, w, arr, nc, _native.stream_ptr())
        _native.check(rc, "mrisr_vgg_features")
        return feat


class _PerceptualFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, generated, target, extractor, loss_type):
        pred = VGGFeatureExtractor._gray(generated, "generated")
        tgt = VGGFeatureExtractor._gray(target, "target")
        if pred.shape != tgt.shape:
            raise ValueError(f"PerceptualLoss: shape mismatch {tuple(generated.shape)} vs {tuple(target.shape)}")
        n, _, h, w = pred.shape
        need_grad = generated.requires_grad
        flat, packed = extractor._operands(pred.device)
        ws = extractor._workspace(n, h, w, need_grad, pred.device)
        arr, nc = extractor._cfg_array()
        out1 = torch.empty(1, dtype=torch.float32, device=pred.device)
        dpred = torch.zeros_like(pred) if need_grad else None
        if extractor.precision == "fp32":
            rc = _native.lib().mrisr_vgg_perceptual_fwd_bwd_fp32(
                _native.ptr(flat), _native.ptr(pred), _native.ptr(tgt), _native.ptr(dpred), _native.ptr(out1),
                _native.ptr(ws), ws.numel(), n, h, w, arr, nc, int(loss_type), 1.0, _native.stream_ptr())
        else:
            rc = _native.lib().mrisr_vgg_perceptual_fwd_bwd(
                _native.ptr(packed), _native.ptr(flat), _native.ptr(pred), _native.ptr(tgt), _native.ptr(dpred), _native.ptr(out1),
                _native.ptr(ws), ws.numel(), n, h, w, arr, nc, int(loss_type), 1.0, _native.stream_ptr())
        _native.check(rc, "mrisr_vgg_perceptual_fwd_bwd")
        ctx.shape, ctx.dtype = tuple(generated.shape), generated.dtype
        if need_grad:
            ctx.save_for_backward(dpred)
        return out1[0].clone()

    @staticmethod
    def backward(ctx, grad_out):
        (dpred,) = ctx.saved_tensors
        return (dpred * grad_out).view(ctx.shape).to(ctx.dtype), None, None, None


class PerceptualLoss(nn.Module):
    """criterion(VGG(generated), no_grad VGG(target)) (utils/losses.py:120-151)."""

    def __init__(self, feature_layer_idx=35, loss_type="l1", vgg_weights=None, precision=None):
        super().__init__()
        self.feature_extractor = VGGFeatureExtractor(feature_layer_idx=feature_layer_idx, weights=vgg_weights, precision=precision)
        if loss_type == "l1":
            self.criterion = nn.L1Loss()
            self._loss_code = 1
        elif loss_type in ("l2", "mse"):
            self.criterion = nn.MSELoss()
            self._loss_code = 2
        else:
            raise ValueError(f"Unsupported loss type for PerceptualLoss: {loss_type}")
        self.feature_extractor.eval()

    def forward(self, generated, target):
        return _PerceptualFunction.apply(generated, target, self.feature_extractor, self._loss_code)
