"""Batched multi-view rasterization: V cameras over one replicated set of Gaussians in ONE native
call (one preprocess launch, one sort, one blend grid of V*tiles CTAs), with per-Gaussian gradients
summed over the V views in backward.  This is the schedule BASELINE.json's refine iteration uses
(20 views per Adam step); V == 1 reproduces `GaussianRasterizer` exactly.
"""
from __future__ import annotations

from typing import NamedTuple, Optional

import torch

from . import _C
from .rasterizer import _empty, _opt


class MultiViewSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    kernel_size: float
    bg: torch.Tensor            # [3]
    scale_modifier: float
    viewmatrices: torch.Tensor  # [V,4,4] (reference row-vector convention, one per view)
    projmatrices: torch.Tensor  # [V,4,4]
    sh_degree: int
    campos: torch.Tensor        # [V,3]
    require_depth: bool = True
    require_coord: bool = True
    debug: bool = False
    tile_rows: Optional[torch.Tensor] = None   # [V,2] int32 CUDA: tile-row window rendered per view (None: whole image)
    loss_rows: Optional[list] = None            # flat [2V] pixel rows whose loss terms this call owns (refine loop)


class _RasterizeViews(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, sh, opacities, scales, rotations, s: MultiViewSettings):
        e = _empty()
        out = _C.rasterize_views(s.bg, means3D, e, opacities, scales, rotations, s.scale_modifier, e,
                                 s.viewmatrices, s.projmatrices, s.tanfovx, s.tanfovy, s.kernel_size,
                                 s.image_height, s.image_width, sh, s.sh_degree, s.campos, False,
                                 s.require_coord, s.require_depth, s.debug,
                                 s.tile_rows)
        num_rendered, color, coord, mcoord, alpha, normal, depth, mdepth, radii, geom, binning, img = out
        ctx.s = s
        ctx.num_rendered = num_rendered
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(radii)
        ctx.save_for_backward(means3D, scales, rotations, normal, radii, sh, geom, binning, img, alpha)
        return color, radii, coord, mcoord, depth, mdepth, alpha, normal

    @staticmethod
    def backward(ctx, g_color, g_radii, g_coord, g_mcoord, g_depth, g_mdepth, g_alpha, g_normal):
        s = ctx.s
        means3D, scales, rotations, normal, radii, sh, geom, binning, img, alpha = ctx.saved_tensors
        e = _empty()
        (g_means2D, _g_colors, g_opac, g_means3D, _g_cov, g_sh, g_scales, g_rot) = _C.rasterize_views_backward(
            s.bg, means3D, radii, e, scales, rotations, s.scale_modifier, e, s.viewmatrices, s.projmatrices,
            s.tanfovx, s.tanfovy, s.kernel_size, _opt(g_color), _opt(g_coord), _opt(g_mcoord), _opt(g_depth),
            _opt(g_mdepth), _opt(g_alpha), _opt(g_normal), normal, sh, s.sh_degree, s.campos, geom,
            ctx.num_rendered, binning, img, alpha, s.require_coord, s.require_depth, s.debug)
        ctx.grad_means2D = g_means2D  # [V,P,3] screen-space gradients (densification statistics)
        return g_means3D, g_sh, g_opac, g_scales, g_rot, None


def rasterize_views(means3D, shs, opacities, scales, rotations, settings: MultiViewSettings):
    """Returns (color [V,3,H,W], radii [V,P], coord, mcoord, depth, mdepth, alpha, normal)."""
    return _RasterizeViews.apply(means3D, shs, opacities, scales, rotations, settings)
