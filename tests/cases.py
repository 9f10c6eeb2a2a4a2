"""Small deterministic rasterizer cases shared by the golden generator (tools/make_golden_ref.py, run on
the GPU box against the UNMODIFIED reference CUDA build) and by the parity tests.  The golden files
store the inputs too, so tests never depend on RNG reproducibility across torch versions.

Edge cases follow SURVEY.md §4(3): Gaussians at z ~= 0.2 (near cull, T3), rects on the image border
and ragged image sizes (T7), degenerate / huge covariances (T5, T6), empty tiles (T10), opacities
below 1/255, precomputed colours / covariances (T23), SH degrees 0..3 (T22).
"""
from __future__ import annotations

import math
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

from igs_b200 import scenes  # noqa: E402

GOLDEN_DIR = Path(__file__).resolve().parent / "golden"

MAP_NAMES = ["color", "coord", "mcoord", "depth", "mdepth", "alpha", "normal"]
MAP_CH = [3, 3, 3, 1, 1, 1, 3]
GRAD_NAMES = ["means3D", "means2D", "sh", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp"]


def _base(P, W, H, seed, log_scale_mean, view=0, V=3):
    sc = scenes.make_gaussians(P, seed, log_scale_mean)
    cams = scenes.make_cameras(V, W, H, seed + 1)
    scales, rot, opac = scenes.activated(sc)
    return dict(
        P=P, W=W, H=H, sh_degree=3, kernel_size=0.0, scale_modifier=1.0,
        require_coord=True, require_depth=True,
        tanfovx=float(cams.tanfovx), tanfovy=float(cams.tanfovy),
        bg=np.zeros(3, np.float32),
        means3D=sc.xyz.numpy().copy(), scales=scales.numpy().copy(), rotations=rot.numpy().copy(),
        opacities=opac.numpy().copy(), shs=sc.sh.numpy().copy(),
        viewmatrix=cams.viewmatrix[view].numpy().copy(), projmatrix=cams.projmatrix[view].numpy().copy(),
        campos=cams.campos[view].numpy().copy(),
    )


def case_small():
    """1500 Gaussians, ragged 100x76 image (7x5 tiles, partial edge tiles)."""
    return _base(1500, 100, 76, 10, math.log(0.25))


def case_dense():
    """Few tiles, long per-tile lists (> 256 instances: several staging rounds) with early termination."""
    c = _base(3000, 64, 48, 20, math.log(0.6))
    c["opacities"] = np.clip(c["opacities"] * 1.5, 0, 1).astype(np.float32)
    c["bg"] = np.array([0.3, 0.1, 0.7], np.float32)
    return c


def case_edge():
    """Near-plane, border, degenerate and sub-threshold Gaussians, scale_modifier != 1."""
    c = _base(800, 90, 70, 30, math.log(0.3), view=1)
    rng = np.random.RandomState(5)
    m, s, o = c["means3D"], c["scales"], c["opacities"]
    V = c["viewmatrix"]  # row-vector convention: p_view = [p,1] @ V
    Vinv = np.linalg.inv(V.astype(np.float64))

    def at_view(x, y, z):
        return (np.array([x, y, z, 1.0]) @ Vinv)[:3].astype(np.float32)

    # 0..39: depths straddling the 0.2 near cull
    for i in range(40):
        z = 0.2 + (i - 20) * 1e-3
        m[i] = at_view(rng.uniform(-0.05, 0.05), rng.uniform(-0.05, 0.05), z)
        s[i] = 0.01
    # 40..79: far outside the frustum laterally (no lateral cull in the reference, clamps in cov2D)
    for i in range(40, 80):
        m[i] = at_view(rng.choice([-1, 1]) * rng.uniform(8, 30), rng.uniform(-3, 3), rng.uniform(2, 10))
    # 80..119: huge splats covering the whole image
    s[80:120] = rng.uniform(2.0, 6.0, size=(40, 3)).astype(np.float32)
    # 120..159: needle / pancake (ill-conditioned covariance -> eigen branch forward.cu:142-155)
    s[120:160, 0] = 1e-4
    s[140:160, 1] = 1e-4
    # 160..199: almost-zero scales (det underflow, T5/T6 max(0.1,..))
    s[160:200] = 1e-6
    # 200..239: opacities around the 1/255 threshold and exactly 0 / 1
    o[200:240, 0] = np.linspace(0.0, 2.0 / 255.0, 40, dtype=np.float32)
    o[240, 0] = 1.0
    o[241, 0] = 0.0
    # 242..260: identical depth and position (stable tie-break on Gaussian index, T9)
    m[242:261] = m[242]
    c["scale_modifier"] = 1.3
    c["bg"] = np.array([1.0, 1.0, 1.0], np.float32)
    return c


def case_precomp():
    """colors_precomp + cov3D_precomp instead of SH + scale/rotation (T23)."""
    c = _base(600, 80, 64, 40, math.log(0.25))
    S = c.pop("scales").astype(np.float64)
    q = c.pop("rotations").astype(np.float64)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                  2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                  2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).reshape(-1, 3, 3)
    L = R * S[:, None, :]
    cov = L @ L.transpose(0, 2, 1)
    c["cov3D_precomp"] = np.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]],
                                  -1).astype(np.float32)
    rng = np.random.RandomState(7)
    c["colors_precomp"] = rng.uniform(0, 1, size=(600, 3)).astype(np.float32)
    c.pop("shs")
    return c


def case_sh1():
    """SH degree 1 of 16 stored coefficients, require_coord only."""
    c = _base(700, 72, 56, 50, math.log(0.3), view=2)
    c["sh_degree"] = 1
    c["require_depth"] = False
    return c


def case_sh0_nogeo():
    """SH degree 0, no coord / depth maps requested (the <false,false,false> template)."""
    c = _base(700, 64, 64, 60, math.log(0.3))
    c["sh_degree"] = 0
    c["require_depth"] = False
    c["require_coord"] = False
    return c


def case_depthonly():
    c = _base(500, 48, 80, 70, math.log(0.3))
    c["sh_degree"] = 2
    c["require_coord"] = False
    return c


def case_empty_view():
    """Camera looking away: everything culled, R == 0 (rasterizer_impl.cu:386)."""
    c = _base(300, 64, 48, 80, math.log(0.3))
    V = c["viewmatrix"].copy()
    V[:, 2] *= -1.0  # flip view-space z
    c["viewmatrix"] = V
    return c


def case_kernel3d():
    """kernel_size != 0 (RaDe-GS frame-0 training uses 0.1): the low-pass term of computeCov2D and the opacity
    compensation coef = sqrt(det_0 / det_1), forward and backward (SURVEY T5 / T19 with kernel_size != 0)."""
    c = _base(1200, 96, 80, 90, math.log(0.2), view=1)
    c["kernel_size"] = 0.1
    return c


def case_scalemod():
    """scale_modifier != 1 together with a large kernel_size."""
    c = _base(900, 80, 64, 100, math.log(0.3), view=2)
    c["kernel_size"] = 0.3
    c["scale_modifier"] = 0.7
    return c


CASES = {
    "small": case_small, "dense": case_dense, "edge": case_edge, "precomp": case_precomp, "sh1": case_sh1,
    "sh0_nogeo": case_sh0_nogeo, "depthonly": case_depthonly, "empty_view": case_empty_view,
    "kernel3d": case_kernel3d, "scalemod": case_scalemod,
}


def cotangents(case: dict, seed: int = 3):
    """Seven N(0,1) cotangent maps [C,H,W] in MAP_NAMES order."""
    rng = np.random.RandomState(seed)
    H, W = case["H"], case["W"]
    return [rng.standard_normal((c, H, W)).astype(np.float32) for c in MAP_CH]


def load_golden(name: str):
    f = GOLDEN_DIR / f"ref_{name}.npz"
    if not f.exists():
        return None
    return dict(np.load(f, allow_pickle=False))


def case_from_golden(g: dict) -> dict:
    c = {}
    for k, v in g.items():
        if k.startswith("in_"):
            c[k[3:]] = v.item() if v.ndim == 0 else v
    for k in ("P", "W", "H", "sh_degree"):
        c[k] = int(c[k])
    for k in ("require_coord", "require_depth"):
        c[k] = bool(c[k])
    for k in ("kernel_size", "scale_modifier", "tanfovx", "tanfovy"):
        c[k] = float(c[k])
    return c


def to_torch_args(case: dict, device):
    """Keyword tensors for GaussianRasterizer.forward + the settings kwargs."""
    def t(k):
        return torch.from_numpy(np.ascontiguousarray(case[k])).to(device) if k in case else None
    settings = dict(image_height=case["H"], image_width=case["W"], tanfovx=case["tanfovx"], tanfovy=case["tanfovy"],
                    kernel_size=case["kernel_size"], bg=t("bg"), scale_modifier=case["scale_modifier"],
                    viewmatrix=t("viewmatrix"), projmatrix=t("projmatrix"), sh_degree=case["sh_degree"],
                    campos=t("campos"), prefiltered=False, require_depth=case["require_depth"],
                    require_coord=case["require_coord"], debug=False)
    tensors = dict(means3D=t("means3D"), opacities=t("opacities"), shs=t("shs"), colors_precomp=t("colors_precomp"),
                   scales=t("scales"), rotations=t("rotations"), cov3D_precomp=t("cov3D_precomp"))
    return settings, tensors
