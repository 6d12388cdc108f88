"""Backend ops — mirror of PyraPose/backend/common.py (decode math, shift) on CUDA tensors."""
from .common import bbox_transform_inv, box3D_transform_inv, box3D_denormalization, shift  # noqa: F401
