"""Drop-in replacement for the reference package `diff_gaussian_rasterization`
(thirdparty/diff-gaussian-rasterization-w-pose/diff_gaussian_rasterization/__init__.py):
same three public names, same argument order, outputs and gradients, backed by the
B200-native library in wildgs_slam_blackwell_b200 (hand-written sm_100a CUDA behind a C ABI).

    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
"""
from wildgs_slam_blackwell_b200.rasterizer import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizer,
    rasterize_gaussians,
    _RasterizeGaussians,
    cpu_deep_copy_tuple,
    grad_sink,
    tile_rows,
)

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]
