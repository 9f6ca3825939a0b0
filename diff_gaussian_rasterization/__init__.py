"""Un-suffixed alias named by BASELINE.json's north_star
(`diff_gaussian_rasterization.GaussianRasterizer / GaussianRasterizationSettings`)."""
from diff_gaussian_rasterization_32 import *  # noqa: F401,F403
from diff_gaussian_rasterization_32 import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizer_32,
    GaussianRasterizer,
    rasterize_gaussians,
    _C,
)
