"""B200-native differentiable Gaussian rasterizer (drop-in for diff_gaussian_rasterization)."""
__version__ = "0.1.0"
