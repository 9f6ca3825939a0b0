"""Drop-in for the reference package `diff_gaussian_rasterization_32`
(/root/reference/submodules/diff-gaussian-rasterization-32/
diff_gaussian_rasterization_32/__init__.py), backed by the B200-native
library.  `from diff_gaussian_rasterization_32 import
GaussianRasterizationSettings, GaussianRasterizer_32` -- the import GUAVA makes at
models/UbodyAvatar/gaussian_render.py:4 -- works unchanged.
"""
from guava_renderer_b200.rasterizer import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizer_32,
    GaussianRasterizer,
    _RasterizeGaussians,
    rasterize_gaussians,
    _C,
)
import guava_renderer_b200 as _impl  # noqa: F401  (loads the CUDA library or raises)


def cpu_deep_copy_tuple(input_tuple):
    import torch
    copied_tensors = [item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple]
    return tuple(copied_tensors)
