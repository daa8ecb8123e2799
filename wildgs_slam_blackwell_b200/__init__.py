"""B200-native differentiable 3D-Gaussian rasterizer with camera-pose gradients: the one hot path
of yuhang1008/WildGS-SLAM-Blackwell (thirdparty/diff-gaussian-rasterization-w-pose), rebuilt as
hand-written sm_100a CUDA behind a C ABI (include/wgs_rasterizer.h) with the reference's Python API."""
from .rasterizer import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizer,
    rasterize_gaussians,
)

__version__ = "0.1.0"
