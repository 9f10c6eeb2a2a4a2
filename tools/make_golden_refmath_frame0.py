"""Generates tests/golden/refmath_frame0.npz by IMPORTING the reference's own RaDe-GS helpers in the build container
(CPU): utils/graphics_utils.py depth_double_to_normal / point_double_to_normal (their hard-coded `.cuda()` calls are
made no-ops for the duration of the call) evaluated inside the depth-normal loss expression of train.py:143-157, with
autograd gradients; and scene/gaussian_model.py's 3D-filter algebra (get_scaling_n_opacity_with_3D_filter,
compute_3D_filter) called as unbound functions on a stand-in object.
Run here (needs /root/reference):  python tools/make_golden_refmath_frame0.py
"""
import importlib.util
import math
import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
RADE = Path("/root/reference/submodules/RaDe-GS")
sys.path.insert(0, str(ROOT))


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def main():
    gu = load(RADE / "utils" / "graphics_utils.py", "ref_rade_graphics_utils")
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    out = {}
    try:
        g = torch.Generator().manual_seed(77)
        for name, (H, W, fovx) in dict(a=(23, 37, 1.0), b=(40, 64, 0.7)).items():
            fovy = 2 * math.atan(math.tan(fovx / 2) * H / W)
            view = types.SimpleNamespace(image_width=W, image_height=H, FoVx=fovx, FoVy=fovy)
            yy, xx = torch.meshgrid(torch.arange(H).float(), torch.arange(W).float(), indexing="ij")
            base = 3.0 + 0.02 * xx + 0.03 * yy + 0.3 * torch.sin(xx / 5.0) * torch.cos(yy / 4.0)
            d1 = (base + 0.02 * torch.randn(H, W, generator=g)).unsqueeze(0).requires_grad_(True)
            d2 = (base + 0.05 * torch.randn(H, W, generator=g)).unsqueeze(0).requires_grad_(True)
            nrm = torch.nn.functional.normalize(torch.randn(3, H, W, generator=g), dim=0)
            nrm = (nrm * (0.5 + 0.5 * torch.rand(1, H, W, generator=g))).requires_grad_(True)   # alpha-weighted, like the rasterizer's
            dn = gu.depth_double_to_normal(view, d1, d2)
            err = (1 - (nrm.unsqueeze(0) * dn).sum(dim=1))
            loss = 0.4 * err[0].mean() + 0.6 * err[1].mean()
            loss.backward()
            out.update({f"{name}_depth1": d1.detach().numpy(), f"{name}_depth2": d2.detach().numpy(), f"{name}_normal": nrm.detach().numpy(),
                        f"{name}_loss": np.float32(loss.item()), f"{name}_d_depth1": d1.grad.numpy(), f"{name}_d_depth2": d2.grad.numpy(),
                        f"{name}_d_normal": nrm.grad.numpy(), f"{name}_fov": np.float32([fovx, fovy])})
            # coord variant
            p1 = (torch.stack([xx / W, yy / H, base]) + 0.02 * torch.randn(3, H, W, generator=g)).requires_grad_(True)
            p2 = (torch.stack([xx / W, yy / H, base]) + 0.04 * torch.randn(3, H, W, generator=g)).requires_grad_(True)
            dn = gu.point_double_to_normal(view, p1, p2)
            err = (1 - (nrm.detach().unsqueeze(0) * dn).sum(dim=1))
            loss = 0.4 * err[0].mean() + 0.6 * err[1].mean()
            loss.backward()
            out.update({f"{name}_coord1": p1.detach().numpy(), f"{name}_coord2": p2.detach().numpy(), f"{name}_closs": np.float32(loss.item()),
                        f"{name}_d_coord1": p1.grad.numpy(), f"{name}_d_coord2": p2.grad.numpy()})
    finally:
        torch.Tensor.cuda = orig_cuda

    # 3D filter algebra: the reference methods, unbound, on a stand-in object (the class itself needs simple_knn etc.)
    src = (RADE / "scene" / "gaussian_model.py").read_text()
    import re
    import textwrap
    def method(nm):
        m = re.search(rf"\n(    def {nm}\(self.*?)(?=\n    (?:@|def ))", src, re.S)
        return textwrap.dedent(m.group(1))
    ns = {"torch": torch, "math": math}
    exec(method("get_scaling_n_opacity_with_3D_filter"), ns)
    exec(method("compute_3D_filter").replace("@torch.no_grad()", ""), ns)
    from igs_b200 import scenes
    P = 500
    sc = scenes.make_gaussians(P, 5, -1.8)
    cams = scenes.make_cameras(3, 64, 48, 1)
    obj = types.SimpleNamespace()
    obj.get_xyz = sc.xyz
    obj.get_scaling = torch.exp(sc.log_scale)
    obj.opacity_activation = torch.sigmoid
    obj._opacity = sc.opacity_logit
    fovx = 2 * math.atan(cams.tanfovx)
    fovy = 2 * math.atan(cams.tanfovy)
    camobjs = []
    for v in range(3):
        vm = cams.viewmatrix[v]
        # world_view_transform = getWorld2View2(R, T).T with Rt[:3,:3] = R.T, so R = vm[:3,:3] and T = vm[3,:3]
        camobjs.append(types.SimpleNamespace(image_width=64, image_height=48, FoVx=fovx, FoVy=fovy,
                                             R=vm[:3, :3].numpy(), T=vm[3, :3].numpy()))
    with torch.no_grad():
        ns["compute_3D_filter"](obj, camobjs)
    scales, opac = ns["get_scaling_n_opacity_with_3D_filter"](obj)
    out.update({"f_xyz": sc.xyz.numpy(), "f_log_scale": sc.log_scale.numpy(), "f_opacity_logit": sc.opacity_logit.numpy(),
                "f_viewmatrices": cams.viewmatrix.numpy(), "f_tanfov": np.float32([cams.tanfovx, cams.tanfovy]),
                "f_filter_3D": obj.filter_3D.numpy(), "f_scales": scales.numpy(), "f_opacity": opac.numpy()})
    np.savez_compressed(ROOT / "tests" / "golden" / "refmath_frame0.npz", **out)
    print({k: (v.shape if hasattr(v, "shape") and v.ndim else v) for k, v in out.items()})


if __name__ == "__main__":
    main()
