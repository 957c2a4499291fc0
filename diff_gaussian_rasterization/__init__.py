"""Drop-in package name of the reference's rasterizer
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py): the
reference's gaussian_renderer/__init__.py:14 does
`from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer`.
Everything is implemented in gaussian_splatting_b200.rasterizer over libgsrast.so.
`SparseGaussianAdam` is deliberately absent, as in this fork (train.py:41-45 probes for it).
"""
from gaussian_splatting_b200.rasterizer import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizer,
    _RasterizeGaussians,
    cpu_deep_copy_tuple,
    rasterize_gaussians,
)
