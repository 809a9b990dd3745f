"""B200-native (sm_100a) implementation of the rdd0582/mri_superresolution U-Net super-resolution hot path.

Public surface mirrors the reference's module paths:
    mri_superresolution_b200.models.unet_model.UNetSuperRes
    mri_superresolution_b200.utils.losses.{ssim, SSIM, CombinedLoss}
plus ``optim.FusedAdam`` and ``parallel.GradientBuckets`` for the data-parallel training step.
"""
from . import _native  # noqa: F401

__version__ = "0.1.0"
