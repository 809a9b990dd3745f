"""Train-step / inference-step oracle: the loop bodies of the reference drivers restated over the
functional oracle (TEST INFRASTRUCTURE, see oracle/__init__.py).

train step  = scripts/train.py:296-323 (zero_grad, forward, CombinedLoss, backward,
              Adam(lr, weight_decay) step, loss.item(), SSIM metric)
infer step  = scripts/infer.py:268-276 (no_grad forward + clamp(0,1))
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch

from .losses_oracle import combined_loss, ssim
from .unet_oracle import unet_forward


class OracleTrainer:
    def __init__(self, params: Dict[str, torch.Tensor], lr: float = 1e-4, weight_decay: float = 1e-5,
                 ssim_weight: float = 0.4, perceptual_weight: float = 0.0, vgg_params=None):
        self.params = {k: v.clone().requires_grad_(True) for k, v in params.items()}
        # scripts/train.py:186  optim.Adam(model.parameters(), lr=..., weight_decay=...)
        self.opt = torch.optim.Adam(list(self.params.values()), lr=lr, weight_decay=weight_decay)
        self.ssim_weight = ssim_weight
        self.perceptual_weight = perceptual_weight
        self.vgg_params = vgg_params

    def step(self, lr_img: torch.Tensor, hr_img: torch.Tensor) -> Tuple[float, float]:
        self.opt.zero_grad(set_to_none=True)                                  # :301
        out = unet_forward(self.params, lr_img)                               # :313
        loss = combined_loss(out, hr_img, self.ssim_weight, self.perceptual_weight, self.vgg_params)  # :314
        loss.backward()                                                       # :317
        self.opt.step()                                                       # :318
        with torch.no_grad():
            s = ssim(out, hr_img)                                             # :322-323
        return float(loss.item()), float(s.item())


@torch.no_grad()
def infer_step(params: Dict[str, torch.Tensor], lr_img: torch.Tensor) -> torch.Tensor:
    return torch.clamp(unet_forward(params, lr_img), 0, 1)                    # scripts/infer.py:275-276
