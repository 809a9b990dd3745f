"""`utils.losses` of the reference (utils/losses.py) served by the B200 modules."""
from mri_superresolution_b200.utils.losses import SSIM, CombinedLoss, create_window, gaussian_window, ssim  # noqa: F401
from mri_superresolution_b200.utils.perceptual import PerceptualLoss, VGGFeatureExtractor  # noqa: F401
