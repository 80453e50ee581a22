"""B200-native differentiable Gaussian-splatting rasterizer.

Drop-in for the one hot path of ChikaYan/mvs_gaussian_splatting: the rasterizer its
`gaussian_renderer.render()` imports and calls (/root/reference/gaussian_renderer/__init__.py:15,
:42-57, :257-265).  Python host code over a C-ABI shared library of hand-written sm_100a CUDA
(include/gsrast_b200.h); no Triton, no multi-backend dispatch, no CPU fallback.
"""
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians,
                         rasterize_forward, rasterize_backward, mark_visible, RasterContext,
                         set_capacity_hint, last_forward_stats, GradientArena, blend_pair_counts,
                         activate, rasterize_forward_raw, rasterize_backward_raw)
from . import _native

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians",
           "rasterize_forward", "rasterize_backward", "mark_visible", "RasterContext",
           "set_capacity_hint", "last_forward_stats", "GradientArena", "blend_pair_counts",
           "activate", "rasterize_forward_raw", "rasterize_backward_raw", "build_native"]


def build_native(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into lib/libgsrast_b200.so (in-tree)."""
    return _native.build(verbose=verbose)
