"""rain_b200 — sm_100a replacement for the CUDA extensions of sharonal10/rain (diff_gaussian_rasterization + simple-knn).

Only the hot path lives here: ``csrc/`` (hand-written CUDA kernels + the C ABI of include/rain_rast.h, built into
``librainrast.so``) and the host-side mirror of the reference's Python interface.  Importing this package does not load
the CUDA library; the first op call does, and raises if it has not been built (there is no CPU fallback).
"""
from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians  # noqa: F401
from ._C import distCUDA2, mark_visible  # noqa: F401

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "distCUDA2", "mark_visible"]
