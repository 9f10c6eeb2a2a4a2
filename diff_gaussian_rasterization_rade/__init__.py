"""Drop-in for the reference package `diff_gaussian_rasterization_rade`
(submodules/RaDe-GS/submodules/diff-gaussian-rasterization; imported by infer_batch.py:23 and
RaDe-GS gaussian_renderer/__init__.py:14).  Everything is served by igs_b200."""
from igs_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, _RasterizeGaussians,  # noqa: F401
                                 cpu_deep_copy_tuple, rasterize_gaussians)
from igs_b200 import _C  # noqa: F401
