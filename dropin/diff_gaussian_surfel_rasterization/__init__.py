"""Top-level alias so `from diff_gaussian_surfel_rasterization import ...` (GR:96-98) resolves to pings_b200."""
from pings_b200.diff_gaussian_surfel_rasterization import *  # noqa: F401,F403
from pings_b200.diff_gaussian_surfel_rasterization import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                                                           rasterize_gaussians, _RasterizeGaussians)
