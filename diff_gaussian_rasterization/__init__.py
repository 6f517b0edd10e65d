"""Drop-in alias: `from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer`
(what gaussian_renderer/__init__.py:15 of the reference does) resolves to the B200-native implementation when this
repository's root is on sys.path ahead of the reference extension."""
from langsplat_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians,  # noqa: F401
                                       cpu_deep_copy_tuple, _RasterizeGaussians)
