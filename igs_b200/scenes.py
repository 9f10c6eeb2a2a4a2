"""Deterministic synthetic scenes and cameras shared by tests, smoke() and bench.py.

Camera matrices are built exactly as the reference's `Camera` does (igs/models/gs.py:62-106):
world_view_transform = w2c^T, projection = getProjectionMatrix(0.01, 100, FoVx, FoVy)^T,
full_proj = world_view @ projection, camera_center = inverse(world_view)[3,:3].
Everything is generated on the CPU with seeded generators so that the GPU box, this container and
the CPU oracle see bit-identical inputs.
"""
from __future__ import annotations

import math
from typing import Dict, NamedTuple

import numpy as np
import torch

BBOX = ((-14.0, -3.0, 9.0), (7.0, 8.0, 17.0))   # reference configs/bbox.json, sear_steak
LOOK_AT = (-3.5, 2.5, 13.0)

CONFIGS: Dict[str, dict] = {
    # BASELINE.json configs[0..4]
    "C1": dict(P=10_000, V=1, W=256, H=256, log_scale_mean=math.log(0.03)),
    "C2": dict(P=200_000, V=1, W=512, H=512, log_scale_mean=math.log(0.03)),
    "C3": dict(P=300_000, V=20, W=1352, H=1014, log_scale_mean=math.log(0.03)),
    "C4": dict(P=150_000, V=13, W=1280, H=720, log_scale_mean=math.log(0.03)),
    "C5": dict(P=2_000_000, V=8, W=3840, H=2160, log_scale_mean=math.log(0.015)),
}


class Scene(NamedTuple):
    xyz: torch.Tensor          # [P,3]
    log_scale: torch.Tensor    # [P,3]  (activation: exp)
    rot: torch.Tensor          # [P,4]  (activation: normalize); stored normalised
    opacity_logit: torch.Tensor  # [P,1] (activation: sigmoid)
    sh: torch.Tensor           # [P,16,3]


class Cameras(NamedTuple):
    viewmatrix: torch.Tensor   # [V,4,4]  world_view_transform (transposed w2c)
    projmatrix: torch.Tensor   # [V,4,4]  full_proj_transform
    campos: torch.Tensor       # [V,3]
    tanfovx: float
    tanfovy: float
    W: int
    H: int


def make_gaussians(P: int, seed: int = 0, log_scale_mean: float = math.log(0.03)) -> Scene:
    g = torch.Generator().manual_seed(seed)
    lo = torch.tensor(BBOX[0])
    hi = torch.tensor(BBOX[1])
    xyz = lo + (hi - lo) * torch.rand(P, 3, generator=g)
    log_scale = (log_scale_mean + 0.6 * torch.randn(P, 3, generator=g)).clamp(math.log(1e-3), math.log(0.5))
    rot = torch.randn(P, 4, generator=g)
    rot = rot / rot.norm(dim=1, keepdim=True)
    opacity_logit = 2.0 * torch.randn(P, 1, generator=g)
    sh = torch.empty(P, 16, 3)
    sh[:, 0, :] = torch.randn(P, 3, generator=g)
    sh[:, 1:, :] = 0.1 * torch.randn(P, 15, 3, generator=g)
    return Scene(xyz.float().contiguous(), log_scale.float().contiguous(), rot.float().contiguous(),
                 opacity_logit.float().contiguous(), sh.float().contiguous())


def get_projection_matrix(znear: float, zfar: float, fovX: float, fovY: float) -> torch.Tensor:
    tanHalfFovY = math.tan(fovY / 2)
    tanHalfFovX = math.tan(fovX / 2)
    top = tanHalfFovY * znear
    bottom = -top
    right = tanHalfFovX * znear
    left = -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def _look_at_c2w(eye: np.ndarray, target: np.ndarray, up: np.ndarray) -> np.ndarray:
    # camera looks down +z (the rasterizer culls z_view <= 0.2), x right, y down
    fwd = target - eye
    fwd = fwd / np.linalg.norm(fwd)
    right = np.cross(fwd, up)
    right = right / np.linalg.norm(right)
    down = np.cross(fwd, right)  # right-handed: x right, y down, z forward
    c2w = np.eye(4)
    c2w[:3, 0] = right
    c2w[:3, 1] = down
    c2w[:3, 2] = fwd
    c2w[:3, 3] = eye
    return c2w


def make_cameras(V: int, W: int, H: int, seed: int = 1) -> Cameras:
    rng = np.random.RandomState(seed)
    xs = np.linspace(-3.0, 3.0, V) if V > 1 else np.array([0.0])
    ys = rng.normal(0.0, 0.2, size=V)
    fovx = math.radians(60.0)
    fovy = 2.0 * math.atan(math.tan(fovx / 2) * H / W)
    proj = get_projection_matrix(0.01, 100.0, fovx, fovy).transpose(0, 1)
    views, fulls, centers = [], [], []
    for i in range(V):
        eye = np.array([xs[i], ys[i], 0.0])
        c2w = torch.tensor(_look_at_c2w(eye, np.array(LOOK_AT), np.array([0.0, -1.0, 0.0])), dtype=torch.float32)
        w2c = torch.inverse(c2w)
        wv = w2c.transpose(0, 1).contiguous()
        full = (wv.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
        views.append(wv)
        fulls.append(full.float())
        centers.append(wv.inverse()[3, :3])
    return Cameras(torch.stack(views).contiguous(), torch.stack(fulls).contiguous(), torch.stack(centers).contiguous(),
                   math.tan(fovx * 0.5), math.tan(fovy * 0.5), W, H)


def make_targets(V: int, H: int, W: int, seed: int = 2) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    return torch.rand(V, 3, H, W, generator=g)


def make_cotangents(V: int, H: int, W: int, seed: int = 3):
    """(color, coord, mcoord, depth, mdepth, alpha, normal) cotangents ~ N(0,1)."""
    g = torch.Generator().manual_seed(seed)
    shapes = [3, 3, 3, 1, 1, 1, 3]
    return tuple(torch.randn(V, c, H, W, generator=g) for c in shapes)


def activated(scene: Scene):
    """Reference activations (igs/models/gaussian_model.py:89-127): exp, sigmoid, normalize."""
    scales = torch.exp(scene.log_scale)
    rot = torch.nn.functional.normalize(scene.rot)
    opac = torch.sigmoid(scene.opacity_logit)
    return scales, rot, opac
