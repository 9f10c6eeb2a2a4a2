"""Drop-in for the reference package `diff_gaussian_rasterization_rade_clamp`
(submodules/RaDe-GS/submodules/diff-gaussian-rasterization-clamp; imported by igs/models/gs.py:39):
identical to the non-clamp package except that five gradients are clamped to +-15."""
from igs_b200.rasterizer import GaussianRasterizationSettings, cpu_deep_copy_tuple  # noqa: F401
from igs_b200.rasterizer import GaussianRasterizerClamp as GaussianRasterizer  # noqa: F401
from igs_b200.rasterizer import _RasterizeGaussiansClamp as _RasterizeGaussians  # noqa: F401
from igs_b200 import _C  # noqa: F401


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)
