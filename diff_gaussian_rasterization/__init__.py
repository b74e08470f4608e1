"""Drop-in for the reference package of the same name
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py): MonoGS's
gaussian_splatting/gaussian_renderer/__init__.py:15-18 imports these two names and nothing else."""
from monogs_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                                    rasterize_gaussians, _RasterizeGaussians)
