"""Drop-in at the reference's import path:
    from submodules.diff_gaussian_rasterization.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
(gaussian_renderer/__init__.py:5 of sharonal10/rain).  Everything is implemented in rain_b200."""
from rain_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, _RasterizeGaussians,  # noqa: F401
                                  cpu_deep_copy_tuple, rasterize_gaussians)
from rain_b200 import _C  # noqa: F401
