"""Drop-in package name of the reference's rasterizer
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py): the
reference's gaussian_renderer/__init__.py:14 does
`from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer`.
Everything is implemented in gaussian_splatting_b200.rasterizer over libgsrast.so.
`SparseGaussianAdam` (probed by train.py:41-45, render.py:23-27, scene/gaussian_model.py:25-28; absent in
the fork) is provided by gaussian_splatting_b200.optim, which switches the fork's own accelerated branches on:
render() then passes `dc=` / `shs=` separately and `--optimizer_type sparse_adam` uses the fused Adam kernel.
"""
from gaussian_splatting_b200.optim import SparseGaussianAdam  # noqa: F401
from gaussian_splatting_b200.rasterizer import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizer,
    _RasterizeGaussians,
    cpu_deep_copy_tuple,
    rasterize_gaussians,
)
