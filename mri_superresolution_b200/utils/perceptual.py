"""VGG19[:36] perceptual term of CombinedLoss (utils/losses.py:83-151 in the reference)."""
import torch.nn as nn


class PerceptualLoss(nn.Module):
    def __init__(self, feature_layer_idx=35, loss_type="l1"):
        super().__init__()
        if loss_type not in ("l1", "l2", "mse"):
            raise ValueError(f"Unsupported loss type for PerceptualLoss: {loss_type}")
        raise NotImplementedError("the VGG19 perceptual branch is not built yet on the B200 kernels "
                                  "(SURVEY.md 8(a) rows L6/L7); use perceptual_weight=0")
