"""RaDe-GS frame-0 training step on the B200-native rasterizer.

Reference: submodules/RaDe-GS/train.py:109-161 (one iteration: SH-degree ramp, render, L1 + D-SSIM,
depth-normal consistency regulariser, backward), submodules/RaDe-GS/gaussian_renderer/__init__.py:22-110
(`render`: 3D-filtered scales / opacities into the rasterizer), submodules/RaDe-GS/scene/gaussian_model.py:118-160,
181-233 (`get_scaling_with_3D_filter`, `get_opacity_with_3D_filter`, `compute_3D_filter`),
submodules/RaDe-GS/utils/graphics_utils.py:97-126 (`depths_double_to_points`, `point_double_to_normal`).

What is native here: the rasterizer with its non-zero `kernel_size` and all seven cotangents (forward + the 25-value
backward blend), the fused L1 + D-SSIM loss kernel (the same form as the refine loss with lambda_l1 = 1 - lambda_dssim)
and the depth-normal consistency loss (`_C.depth_normal_loss`, csrc/geo_loss.cu: value and the three gradients in one
launch instead of ~20 torch kernels and their autograd nodes).  The 3D-filter algebra is a handful of [P]-sized torch
ops, as in the reference.  `*_torch` functions restate the reference's torch code with a device argument (the
reference hard-codes `.cuda()`); they are what the tests compare the kernels with.
"""
from __future__ import annotations

import math
from typing import NamedTuple, Optional

import torch
import torch.nn.functional as F

from . import _C
from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer
from .refine import refine_loss_sum


# ------------------------------------------------------------------------------------------------
# 3D filter (Mip-Splatting style), scene/gaussian_model.py:181-233 and :118-160
# ------------------------------------------------------------------------------------------------
@torch.no_grad()
def compute_3D_filter(xyz: torch.Tensor, viewmatrices: torch.Tensor, width: int, height: int, tanfovx: float,
                      tanfovy: float) -> torch.Tensor:
    """filter_3D [P,1] = (min camera-space depth over the cameras that see the point) / focal * sqrt(0.2).

    viewmatrices [V,4,4] are world_view_transforms (transposed w2c): xyz_cam = xyz @ vm[:3,:3] + vm[3,:3], which is
    the reference's `xyz @ R + T` with its transposed-R storage.  All cameras share one resolution / FoV here."""
    P = xyz.shape[0]
    distance = torch.full((P,), 100000.0, device=xyz.device)
    valid_points = torch.zeros(P, device=xyz.device, dtype=torch.bool)
    focal_x = width / (2.0 * tanfovx)
    focal_y = height / (2.0 * tanfovy)
    for vm in viewmatrices:
        cam = xyz @ vm[:3, :3] + vm[3, :3][None, :]
        valid_depth = cam[:, 2] > 0.2
        z = torch.clamp(cam[:, 2], min=0.001)
        x = cam[:, 0] / z * focal_x + width / 2.0
        y = cam[:, 1] / z * focal_y + height / 2.0
        in_screen = (x >= -0.15 * width) & (x <= width * 1.15) & (y >= -0.15 * height) & (y <= 1.15 * height)
        valid = valid_depth & in_screen
        distance[valid] = torch.min(distance[valid], z[valid])
        valid_points |= valid
    if bool(valid_points.any()):
        distance[~valid_points] = distance[valid_points].max()
    return (distance / focal_x * (0.2 ** 0.5))[..., None]


def scaling_n_opacity_with_3D_filter(log_scale: torch.Tensor, opacity_logit: torch.Tensor, filter_3D: torch.Tensor):
    """get_scaling_n_opacity_with_3D_filter (scene/gaussian_model.py:157-170): differentiable torch ops."""
    scales = torch.exp(log_scale)
    s2 = torch.square(scales)
    det1 = s2.prod(dim=1)
    s2f = s2 + torch.square(filter_3D)
    det2 = s2f.prod(dim=1)
    coef = torch.sqrt(det1 / det2)
    return torch.sqrt(s2f), torch.sigmoid(opacity_logit) * coef[..., None]


# ------------------------------------------------------------------------------------------------
# depth -> normal (utils/graphics_utils.py:97-126), torch restatement with a device argument
# ------------------------------------------------------------------------------------------------
def depths_double_to_points_torch(width, height, tanfovx, tanfovy, depthmap1, depthmap2):
    dev = depthmap1.device
    fx = width / (2 * tanfovx)
    fy = height / (2 * tanfovy)
    intrins_inv = torch.tensor([[1 / fx, 0., -width / (2 * fx)], [0., 1 / fy, -height / (2 * fy)], [0., 0., 1.0]]).float().to(dev)
    grid_x, grid_y = torch.meshgrid(torch.arange(width) + 0.5, torch.arange(height) + 0.5, indexing='xy')
    points = torch.stack([grid_x, grid_y, torch.ones_like(grid_x)], dim=0).reshape(3, -1).float().to(dev)
    rays_d = intrins_inv @ points
    points1 = depthmap1.reshape(1, -1) * rays_d
    points2 = depthmap2.reshape(1, -1) * rays_d
    return points1.reshape(3, height, width), points2.reshape(3, height, width)


def point_double_to_normal_torch(points1, points2):
    points = torch.stack([points1, points2], dim=0)
    output = torch.zeros_like(points)
    dx = points[..., 2:, 1:-1] - points[..., :-2, 1:-1]
    dy = points[..., 1:-1, 2:] - points[..., 1:-1, :-2]
    output[..., 1:-1, 1:-1] = F.normalize(torch.cross(dx, dy, dim=1), dim=1)
    return output


def depth_normal_loss_torch(normal, map_expected, map_median, width, height, tanfovx, tanfovy, use_coord=False,
                            depth_ratio: float = 0.6):
    """train.py:143-157 for one view: normal [3,H,W], maps [1,H,W] depths or [3,H,W] coords."""
    if use_coord:
        dn = point_double_to_normal_torch(map_expected, map_median)
    else:
        dn = point_double_to_normal_torch(*depths_double_to_points_torch(width, height, tanfovx, tanfovy, map_expected, map_median))
    err = 1 - (normal.unsqueeze(0) * dn).sum(dim=1)
    return (1 - depth_ratio) * err[0].mean() + depth_ratio * err[1].mean()


class _DepthNormalLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, m0, m1, normal, use_coord, tanfovx, tanfovy, depth_ratio):
        loss, d0, d1, dn = _C.depth_normal_loss(m0, m1, normal, bool(use_coord), float(tanfovx), float(tanfovy),
                                                float(depth_ratio), 1.0)
        ctx.save_for_backward(d0.view_as(m0), d1.view_as(m1), dn)
        return loss

    @staticmethod
    def backward(ctx, g):
        d0, d1, dn = ctx.saved_tensors
        return d0 * g, d1 * g, dn * g, None, None, None, None


def depth_normal_loss(normal, map_expected, map_median, tanfovx, tanfovy, use_coord=False, depth_ratio: float = 0.6):
    """Fused kernel: normal [V,3,H,W] (or [3,H,W]); maps [V,1,H,W] / [1,H,W] depths, or [V,3,H,W] / [3,H,W] coords.
    Returns the SUM over views of the per-view losses."""
    if normal.dim() == 3:
        normal, map_expected, map_median = normal.unsqueeze(0), map_expected.unsqueeze(0), map_median.unsqueeze(0)
    return _DepthNormalLoss.apply(map_expected.contiguous(), map_median.contiguous(), normal.contiguous(), use_coord,
                                  tanfovx, tanfovy, depth_ratio)


# ------------------------------------------------------------------------------------------------
# render + one training iteration
# ------------------------------------------------------------------------------------------------
class Frame0Options(NamedTuple):
    """OptimizationParams / ModelParams fields the step reads (arguments/__init__.py:79-101)."""
    lambda_dssim: float = 0.2
    lambda_depth_normal: float = 0.05
    regularization_from_iter: int = 10_000
    kernel_size: float = 0.0
    use_coord_map: bool = False
    max_sh_degree: int = 3


def active_sh_degree(iteration: int, max_sh_degree: int = 3) -> int:
    """`oneupSHdegree` every 1000 iterations starting from degree 0 (train.py:112-114)."""
    return min(max_sh_degree, iteration // 1000)


def render(cam, params, filter_3D, bg, kernel_size, sh_degree, require_coord=True, require_depth=True, module=None):
    """gaussian_renderer.render for one camera.  cam: dict(viewmatrix, projmatrix, campos, tanfovx, tanfovy, W, H);
    params: dict(xyz, shs [P,16,3], scaling (log), rotation (raw), opacity (logit)).  `module` is the rasterizer
    package to render with (default: this library; the tests pass the reference build)."""
    m = module
    Settings = m.GaussianRasterizationSettings if m is not None else GaussianRasterizationSettings
    Rast = m.GaussianRasterizer if m is not None else GaussianRasterizer
    settings = Settings(image_height=int(cam["H"]), image_width=int(cam["W"]), tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                        kernel_size=kernel_size, bg=bg, scale_modifier=1.0, viewmatrix=cam["viewmatrix"],
                        projmatrix=cam["projmatrix"], sh_degree=sh_degree, campos=cam["campos"], prefiltered=False,
                        require_coord=require_coord, require_depth=require_depth, debug=False)
    means2D = torch.zeros_like(params["xyz"], requires_grad=True)
    scales, opacity = scaling_n_opacity_with_3D_filter(params["scaling"], params["opacity"], filter_3D)
    img, radii, ecoord, mcoord, edepth, mdepth, alpha, normal = Rast(raster_settings=settings)(
        means3D=params["xyz"], means2D=means2D, shs=params["shs"], colors_precomp=None, opacities=opacity, scales=scales,
        rotations=F.normalize(params["rotation"]), cov3D_precomp=None)
    return {"render": img, "mask": alpha, "expected_coord": ecoord, "median_coord": mcoord, "expected_depth": edepth,
            "median_depth": mdepth, "viewspace_points": means2D, "visibility_filter": radii > 0, "radii": radii,
            "normal": normal}


class Frame0Losses(NamedTuple):
    loss: torch.Tensor
    rgb_loss: torch.Tensor
    depth_normal_loss: torch.Tensor
    render_pkg: dict


def training_loss(cam, params, filter_3D, bg, gt_image, iteration: int, opt: Frame0Options = Frame0Options(),
                  impl: str = "fused", module=None) -> Frame0Losses:
    """Forward part of train.py:109-159 for one iteration; call `.loss.backward()` on the result.

    impl="fused": L1 + D-SSIM and the depth-normal regulariser through the hand-written kernels; impl="torch": the
    reference's torch formulation (restated) — the two agree to fp32 rounding (tests/test_frame0_gpu.py)."""
    reg = iteration >= opt.regularization_from_iter
    require_depth = (not opt.use_coord_map) and reg
    require_coord = opt.use_coord_map and reg
    pkg = render(cam, params, filter_3D, bg, opt.kernel_size, active_sh_degree(iteration, opt.max_sh_degree),
                 require_coord=require_coord, require_depth=require_depth, module=module)
    img = pkg["render"]
    rgb = refine_loss_sum(img.unsqueeze(0), gt_image.unsqueeze(0), 1.0 - opt.lambda_dssim, "fused" if impl == "fused" else "torch")
    if reg:
        if opt.use_coord_map:
            m0, m1 = pkg["expected_coord"], pkg["median_coord"]
        else:
            m0, m1 = pkg["expected_depth"], pkg["median_depth"]
        if impl == "fused":
            dn = depth_normal_loss(pkg["normal"], m0, m1, cam["tanfovx"], cam["tanfovy"], opt.use_coord_map)
        else:
            dn = depth_normal_loss_torch(pkg["normal"], m0, m1, int(cam["W"]), int(cam["H"]), cam["tanfovx"], cam["tanfovy"],
                                         opt.use_coord_map)
        lam = opt.lambda_depth_normal
    else:
        dn = torch.zeros((), device=img.device)
        lam = 0.0
    return Frame0Losses(rgb + dn * lam, rgb, dn, pkg)
