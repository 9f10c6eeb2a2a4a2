"""Python surface of the B200-native splat rasterizer.

Mirrors the reference package `diff_gaussian_rasterization_rade` (and its `_clamp` twin) name for
name: `GaussianRasterizationSettings` (reference __init__.py:177-192), `GaussianRasterizer`
(:194-312), `rasterize_gaussians` (:21-42) and `_RasterizeGaussians` (:44-175).  Differences, all
invisible to a caller that only uses results:
  * the native extension is `igs_b200._C` (hand-written sm_100a kernels behind a C-ABI); it is
    imported eagerly and there is NO CPU fallback — import fails loudly if it is not built;
  * `backward` receives `None` for output maps the loss never touched (set_materialize_grads(False))
    and passes that on as "cotangent identically zero", which lets the kernels skip work that the
    reference multiplies by zero;
  * the reference's seven `assert not torch.isnan(..).any()` host syncs (:156-162) run only when
    `settings.debug` is set or IGS_RASTER_NAN_CHECK=1;
  * `integrate()` (GOF mesh extraction, never reached from IGS) raises NotImplementedError.
The `_clamp` variant clamps five gradients to +-15 after the native call (reference clamp
__init__.py:157-161); it is the same class with `clamp_grads=True`.
"""
from __future__ import annotations

import os
from typing import NamedTuple, Optional

import torch
import torch.nn as nn

try:
    from . import _C  # noqa: F401  (built by igs_b200.build / __graft_entry__.build)
except ImportError as exc:  # pragma: no cover - exercised only on a broken install
    raise ImportError(
        "igs_b200._C (the CUDA extension) is not built; run `python -m igs_b200.build`. "
        "There is deliberately no CPU or PyTorch fallback."
    ) from exc

_NAN_CHECK = os.environ.get("IGS_RASTER_NAN_CHECK", "0") == "1"


def cpu_deep_copy_tuple(input_tuple):
    copied_tensors = [item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple]
    return tuple(copied_tensors)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    kernel_size: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    require_depth: bool
    require_coord: bool
    debug: bool


def _empty():
    return torch.Tensor([])


def _opt(t):
    return t if t is not None else _empty()


class _RasterizeGaussians(torch.autograd.Function):
    clamp_grads = False

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings):
        args = (
            raster_settings.bg,
            means3D,
            colors_precomp,
            opacities,
            scales,
            rotations,
            raster_settings.scale_modifier,
            cov3Ds_precomp,
            raster_settings.viewmatrix,
            raster_settings.projmatrix,
            raster_settings.tanfovx,
            raster_settings.tanfovy,
            raster_settings.kernel_size,
            raster_settings.image_height,
            raster_settings.image_width,
            sh,
            raster_settings.sh_degree,
            raster_settings.campos,
            raster_settings.prefiltered,
            raster_settings.require_coord,
            raster_settings.require_depth,
            raster_settings.debug,
        )
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                out = _C.rasterize_gaussians(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            out = _C.rasterize_gaussians(*args)
        num_rendered, color, coord, mcoord, alpha, normal, depth, mdepth, radii, geomBuffer, binningBuffer, imgBuffer = out

        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(radii)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, normal, radii, sh,
                              geomBuffer, binningBuffer, imgBuffer, alpha)
        return color, radii, coord, mcoord, depth, mdepth, alpha, normal

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_coord, grad_mcoord, grad_depth, grad_mdepth, grad_alpha, grad_normal):
        num_rendered = ctx.num_rendered
        raster_settings = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, normal, radii, sh,
         geomBuffer, binningBuffer, imgBuffer, alpha) = ctx.saved_tensors

        args = (
            raster_settings.bg,
            means3D,
            radii,
            colors_precomp,
            scales,
            rotations,
            raster_settings.scale_modifier,
            cov3Ds_precomp,
            raster_settings.viewmatrix,
            raster_settings.projmatrix,
            raster_settings.tanfovx,
            raster_settings.tanfovy,
            raster_settings.kernel_size,
            _opt(grad_color),
            _opt(grad_coord),
            _opt(grad_mcoord),
            _opt(grad_depth),
            _opt(grad_mdepth),
            _opt(grad_alpha),
            _opt(grad_normal),
            normal,
            sh,
            raster_settings.sh_degree,
            raster_settings.campos,
            geomBuffer,
            num_rendered,
            binningBuffer,
            imgBuffer,
            alpha,
            raster_settings.require_coord,
            raster_settings.require_depth,
            raster_settings.debug,
        )
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                grads_native = _C.rasterize_gaussians_backward(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            grads_native = _C.rasterize_gaussians_backward(*args)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations) = grads_native

        if ctx._forward_cls.clamp_grads:
            grad_means3D = torch.clamp(grad_means3D, -15, 15)
            grad_sh = torch.clamp(grad_sh, -15, 15)
            grad_opacities = torch.clamp(grad_opacities, -15, 15)
            grad_scales = torch.clamp(grad_scales, -15, 15)
            grad_rotations = torch.clamp(grad_rotations, -15, 15)

        if raster_settings.debug or _NAN_CHECK:
            assert not torch.isnan(grad_means3D).any()
            assert not torch.isnan(grad_means2D).any()
            assert not torch.isnan(grad_sh).any()
            assert not torch.isnan(grad_colors_precomp).any()
            assert not torch.isnan(grad_opacities).any()
            assert not torch.isnan(grad_scales).any()
            assert not torch.isnan(grad_rotations).any()

        def _match(g, ref):
            # absent inputs are CPU empties in the reference API; hand back None for them
            return g if (ref is not None and ref.numel() != 0) else None

        grads = (
            grad_means3D,
            grad_means2D,
            _match(grad_sh, sh),
            _match(grad_colors_precomp, colors_precomp),
            grad_opacities,
            _match(grad_scales, scales),
            _match(grad_rotations, rotations),
            _match(grad_cov3Ds_precomp, cov3Ds_precomp),
            None,
        )
        return grads


class _RasterizeGaussiansClamp(_RasterizeGaussians):
    clamp_grads = True


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, clamp: bool = False):
    fn = _RasterizeGaussiansClamp if clamp else _RasterizeGaussians
    return fn.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    clamp_grads = False

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            raster_settings = self.raster_settings
            visible = _C.mark_visible(positions, raster_settings.viewmatrix, raster_settings.projmatrix)
        return visible

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        raster_settings = self.raster_settings

        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')

        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        if shs is None:
            shs = _empty()
        if colors_precomp is None:
            colors_precomp = _empty()
        if scales is None:
            scales = _empty()
        if rotations is None:
            rotations = _empty()
        if cov3D_precomp is None:
            cov3D_precomp = _empty()

        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   raster_settings, clamp=self.clamp_grads)

    def integrate(self, *args, **kwargs):
        raise NotImplementedError(
            "integrate() (GOF point integration for mesh extraction, reference __init__.py:245-312) is outside "
            "the refine-loop hot path this library implements")


class GaussianRasterizerClamp(GaussianRasterizer):
    clamp_grads = True
