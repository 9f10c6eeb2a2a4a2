"""Batched multi-view forward of the streamed Gaussians: what AGM-Net's renderer asks of the
rasterizer, on the batched native call.

Reference: igs/models/gs.py:90-113 (`Camera`, `Camera.from_c2w`), :566-713 (`GS3DRenderer.forward_single_view`: the
`_rade_clamp` rasterizer with kernel_size 0, require_coord = require_depth = True; optional second render of the 2-D
flow of the masked Gaussians through `colors_precomp`), :824-856 (`forward_single_batch`: a Python loop over the Nv
cameras, outputs stacked).  Here the Nv colour renders are ONE `rasterize_views` call (one preprocess, one sort, one
blend grid over Nv x tiles) with gradients summed over views in backward; the flow render keeps one call per view
because its colours are per-view.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch

from .multiview import MultiViewSettings, rasterize_views
from .rasterizer import GaussianRasterizationSettings, GaussianRasterizerClamp
from .scenes import get_projection_matrix


class Camera:
    """igs/models/gs.py:90-113: world_view_transform = w2c^T, znear 0.01, zfar 100."""

    def __init__(self, w2c: torch.Tensor, FoVx: float, FoVy: float, resolution):
        self.FoVx, self.FoVy = float(FoVx), float(FoVy)
        self.height, self.width = int(resolution[0]), int(resolution[1])
        self.world_view_transform = w2c.transpose(0, 1)
        self.zfar, self.znear = 100.0, 0.01
        self.projection_matrix = get_projection_matrix(self.znear, self.zfar, self.FoVx, self.FoVy).transpose(0, 1).to(w2c.device)
        self.full_proj_transform = self.world_view_transform.unsqueeze(0).bmm(self.projection_matrix.unsqueeze(0)).squeeze(0)
        self.camera_center = self.world_view_transform.inverse()[3, :3]

    @staticmethod
    def from_c2w(c2w: torch.Tensor, FOV, resolution) -> "Camera":
        return Camera(torch.inverse(c2w), FOV[0], FOV[1], resolution)


def settings_from_c2ws(c2ws: torch.Tensor, FOV, resolution, bg: torch.Tensor, sh_degree: int = 3,
                       scale_modifier: float = 1.0) -> MultiViewSettings:
    """All Nv cameras of a batch as one MultiViewSettings (batched inverse / matmul instead of Nv Camera objects)."""
    w2c = torch.inverse(c2ws)
    wv = w2c.transpose(1, 2).contiguous()
    proj = get_projection_matrix(0.01, 100.0, float(FOV[0]), float(FOV[1])).transpose(0, 1).to(c2ws.device)
    full = wv @ proj
    campos = torch.inverse(wv)[:, 3, :3].contiguous()
    return MultiViewSettings(int(resolution[0]), int(resolution[1]), math.tan(float(FOV[0]) * 0.5), math.tan(float(FOV[1]) * 0.5),
                             0.0, bg, scale_modifier, wv, full.contiguous().float(), sh_degree, campos, True, True, False)


def flow_colors(flow: Dict[str, torch.Tensor], viewmatrix: torch.Tensor, fx: float, fy: float) -> torch.Tensor:
    """The per-Gaussian "colours" of the flow render (gs.py:693-700): the predicted displacement rotated into the camera
    and scaled by focal / z — the reference divides by the WORLD z of the Gaussian, reproduced as is.  [Pm,3], third
    channel zero."""
    means3D = flow["xyz"]
    flow_cam = flow["resi_xyz"] @ viewmatrix[:3, :3]
    f2 = flow["resi_xyz"].new_zeros((means3D.shape[0], 3))
    f2[:, 0] = flow_cam[:, 0] * fx / (means3D[:, 2] + 1e-6)
    f2[:, 1] = flow_cam[:, 1] * fy / (means3D[:, 2] + 1e-6)
    return f2


def forward_single_batch(gs: Dict[str, torch.Tensor], c2ws: torch.Tensor, FOV, resolution, background_color: torch.Tensor,
                         sh_degree: int = 3, scaling_modifier: float = 1.0, flow: Optional[Dict[str, torch.Tensor]] = None,
                         flow_resolution=None, clamp: bool = True) -> Dict[str, torch.Tensor]:
    """gs: ACTIVATED Gaussians dict(xyz, shs [P,16,3], opacity [P,1], scaling [P,3], rotation [P,4]) (the reference's
    `gs.get_*`).  Returns images_pred [Nv,3,H,W], depth_pred [Nv,1,H,W] (+ alpha / normal / radii), differentiable.
    flow: optional dict(xyz, opacity, scaling, rotation of the masked original Gaussians, resi_xyz [Pm,3]) -> flow_pred
    [Nv,2,h,w], flow_mask [Nv,1,h,w] (gs.py:659-713)."""
    st = settings_from_c2ws(c2ws, FOV, resolution, background_color, sh_degree, scaling_modifier)
    if clamp and torch.is_grad_enabled() and any(t.requires_grad for t in gs.values()):
        per_view = []
        for v in range(c2ws.shape[0]):
            rs = GaussianRasterizationSettings(image_height=st.image_height, image_width=st.image_width, tanfovx=st.tanfovx,
                                               tanfovy=st.tanfovy, kernel_size=0.0, bg=background_color,
                                               scale_modifier=scaling_modifier, viewmatrix=st.viewmatrices[v],
                                               projmatrix=st.projmatrices[v], sh_degree=sh_degree, campos=st.campos[v],
                                               prefiltered=False, require_depth=True, require_coord=True, debug=False)
            per_view.append(GaussianRasterizerClamp(raster_settings=rs)(
                means3D=gs["xyz"], means2D=torch.zeros_like(gs["xyz"], requires_grad=True), shs=gs["shs"], colors_precomp=None,
                opacities=gs["opacity"], scales=gs["scaling"], rotations=gs["rotation"], cov3D_precomp=None))
        color, radii, _coord, _mcoord, depth, _mdepth, alpha, normal = (torch.stack(x) for x in zip(*per_view))
    else:   # nothing to clamp without gradients: the batched call is exact
        color, radii, _coord, _mcoord, depth, _mdepth, alpha, normal = rasterize_views(
            gs["xyz"], gs["shs"], gs["opacity"], gs["scaling"], gs["rotation"], st)
    out = {"images_pred": color, "depth_pred": depth, "alpha": alpha, "normal": normal, "radii": radii}
    if flow is not None:
        fh, fw = flow_resolution if flow_resolution is not None else resolution
        fx = fw / (2 * st.tanfovx)
        fy = fh / (2 * st.tanfovy)
        flows, masks = [], []
        for v in range(c2ws.shape[0]):
            rs = GaussianRasterizationSettings(image_height=int(fh), image_width=int(fw), tanfovx=st.tanfovx, tanfovy=st.tanfovy,
                                               kernel_size=0.0, bg=background_color, scale_modifier=scaling_modifier,
                                               viewmatrix=st.viewmatrices[v], projmatrix=st.projmatrices[v], sh_degree=sh_degree,
                                               campos=st.campos[v], prefiltered=False, require_depth=True, require_coord=True,
                                               debug=False)
            means3D = flow["xyz"]
            f2 = flow_colors(flow, st.viewmatrices[v], fx, fy)
            r = GaussianRasterizerClamp(raster_settings=rs)(
                means3D=means3D, means2D=torch.zeros_like(means3D, requires_grad=True), shs=None, colors_precomp=f2.float(),
                opacities=flow["opacity"], scales=flow["scaling"], rotations=flow["rotation"], cov3D_precomp=None)
            flows.append(r[0][:2])
            masks.append(r[6])
        out["flow_pred"] = torch.stack(flows)
        out["flow_mask"] = torch.stack(masks)
    return out


def quaternion_multiply(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """Hamilton product of the NORMALISED quaternions (w, x, y, z) — igs/utils/general_utils.py:176-203."""
    a, b = torch.nn.functional.normalize(a), torch.nn.functional.normalize(b)
    w1, x1, y1, z1 = a.unbind(-1)
    w2, x2, y2, z2 = b.unbind(-1)
    return torch.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                        w1 * y2 + y1 * w2 + z1 * x2 - x1 * z2, w1 * z2 + z1 * w2 + x1 * y2 - y1 * x2], dim=-1)


def deform(gs: Dict[str, torch.Tensor], resi: Dict[str, torch.Tensor], mask: torch.Tensor) -> Dict[str, torch.Tensor]:
    """The streamed Gaussians moved by AGM-Net's predicted residuals (`GaussianModel.deform`, igs/models/gs.py:347-375):
    gs = dict(xyz, opacity, rotation, scaling, shs) PRE-activation; on the masked Gaussians the rotation is composed
    (quaternion product), `shs` [Pm,48] and every other residual are added.  Returns the deformed dict plus `mask` and
    `resi_xyz` (what the flow render needs)."""
    out = {k: gs[k].clone() for k in ("xyz", "opacity", "rotation", "scaling", "shs")}
    for key, r in resi.items():
        if key == "rotation":
            out[key][mask] = quaternion_multiply(out[key][mask], r)
        elif key == "shs":
            out[key][mask] = r.reshape(-1, 16, 3) + out[key][mask]
        else:
            out[key][mask] = r + out[key][mask]
    out["mask"] = mask
    out["resi_xyz"] = resi["xyz"].clone()
    return out


def activated(gs: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """`get_xyz / get_features / get_opacity / get_scaling / get_rotation` of the stream model (gs.py:230-262)."""
    return dict(xyz=gs["xyz"], shs=gs["shs"], opacity=torch.sigmoid(gs["opacity"]), scaling=torch.exp(gs["scaling"]),
                rotation=torch.nn.functional.normalize(gs["rotation"]))


def flow_inputs(deformed: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """The `flow` argument of forward_single_batch for a deformed model: the masked Gaussians AFTER the deformation
    (the reference passes `original_gs = gs`, gs.py:846) and the predicted displacement (gs.py:686-695)."""
    a, m = activated(deformed), deformed["mask"]
    return dict(xyz=a["xyz"][m], opacity=a["opacity"][m], scaling=a["scaling"][m], rotation=a["rotation"][m],
                resi_xyz=deformed["resi_xyz"])
