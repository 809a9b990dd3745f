"""`models.unet_model` of the reference (models/unet_model.py:116-211) served by the B200 module."""
from mri_superresolution_b200.models.unet_model import UNetSuperRes  # noqa: F401
