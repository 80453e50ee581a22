"""Import shim: lets the reference's unmodified `gaussian_renderer/__init__.py:15`
(`from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer`)
resolve to the B200-native rasterizer when this repository is on sys.path."""
from mvs_gaussian_splatting_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                                                    rasterize_gaussians, _RasterizeGaussians)

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]
