"""igs_b200 — B200-native differentiable Gaussian-splat rasterizer for IGS's refine loop.

Only what the hot path needs lives here: `csrc/` (hand-written sm_100a kernels + C-ABI),
`rasterizer.py` (the reference's Python API), `multiview.py` (batched V-view call),
`refine.py` (the refine-iteration driver used by bench.py).
"""
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, GaussianRasterizerClamp,  # noqa: F401
                         rasterize_gaussians)
from .multiview import MultiViewSettings, rasterize_views  # noqa: F401

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "GaussianRasterizerClamp", "rasterize_gaussians",
           "MultiViewSettings", "rasterize_views"]
