"""guava_renderer_b200 -- B200-native (sm_100a) differentiable 32-channel
Gaussian-splatting rasterizer, drop-in for GUAVA's
`submodules/diff-gaussian-rasterization-32`.

Only what the hot path needs lives here:
  csrc/          hand-written CUDA kernels + the C ABI (include/guava_rasterizer.h)
  _lib.py        ctypes binding of the C ABI (no fallback: raises if the .so is missing)
  rasterizer.py  host-side mirror of the reference's `_C` module and Python surface
  renderer.py    batched replacement for GaussianRenderer's per-view loop
  deform.py      fused deformation -> rasterizer inputs (the step before the path)
  graphed.py     CUDA-graph capture of a fixed-shape render (+ loss + backward) step
  synthetic.py   seeded synthetic clouds / cameras of the benchmark shapes
"""
from ._lib import GuavaLibraryError, launch_count, set_option, get_option, version, buffer_layout  # noqa: F401
from .rasterizer import (  # noqa: F401
    NUM_CHANNELS,
    GaussianRasterizationSettings,
    GaussianRasterizer_32,
    GaussianRasterizer,
    BatchRasterizationSettings,
    GaussianBatchRasterizer,
    rasterize_gaussians,
    rasterize_gaussians_batch,
    _C,
)
from .renderer import render_views  # noqa: F401
from .graphed import GraphedRenderStep  # noqa: F401
from .deform import deform_gaussians  # noqa: F401

# importing the package must fail loudly when the CUDA library is absent
from . import _lib as _lib_mod
_lib_mod.load()
