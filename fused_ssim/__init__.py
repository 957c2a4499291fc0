"""Drop-in package name of the reference's fused-ssim submodule
(submodules/fused-ssim/fused_ssim/__init__.py): train.py:34-38 does `from fused_ssim import fused_ssim`
and, when that import succeeds, computes the D-SSIM term with it (train.py:157-158) instead of the
conv2d-based utils/loss_utils.py:56-86.  Implemented in gaussian_splatting_b200.ssim over libgsrast.so.
"""
from gaussian_splatting_b200.ssim import FusedSSIMMap, allowed_padding, fused_ssim  # noqa: F401
