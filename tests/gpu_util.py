"""Helpers for the GPU parity tests: run a rasterizer package on a tests/cases.py case."""
import numpy as np
import torch

import cases


def run_case(mod, case, dev, cot=None, which="color", native=None):
    """Returns dict(maps=[7 np arrays], radii, grads={name: np}, decoded={...} if native given)."""
    settings_kw, t = cases.to_torch_args(case, dev)
    leaves = {k: v.clone().requires_grad_(True) for k, v in t.items() if v is not None}
    m2 = torch.zeros_like(leaves["means3D"], requires_grad=True)
    rast = mod.GaussianRasterizer(raster_settings=mod.GaussianRasterizationSettings(**settings_kw))
    color, radii, coord, mcoord, depth, mdepth, alpha, normal = rast(means2D=m2, **leaves)
    maps = [color, coord, mcoord, depth, mdepth, alpha, normal]
    out = dict(maps=[m.detach().cpu().numpy() for m in maps], radii=radii.cpu().numpy(), grads={})
    if cot is not None:
        cts = [torch.from_numpy(c).to(dev) for c in cot]
        loss = (maps[0] * cts[0]).sum() if which == "color" else sum((m * c).sum() for m, c in zip(maps, cts))
        loss.backward()
        key = dict(means3D="means3D", sh="shs", colors_precomp="colors_precomp", opacities="opacities", scales="scales",
                   rotations="rotations", cov3D_precomp="cov3D_precomp")
        for g, l in key.items():
            if l in leaves and leaves[l].grad is not None:
                out["grads"][g] = leaves[l].grad.cpu().numpy()
        out["grads"]["means2D"] = m2.grad.cpu().numpy()
    if native is not None:
        e = torch.Tensor([])
        o = lambda x: x if x is not None else e  # noqa: E731
        nat = native.rasterize_gaussians(settings_kw["bg"], t["means3D"], o(t["colors_precomp"]), t["opacities"], o(t["scales"]),
                                         o(t["rotations"]), settings_kw["scale_modifier"], o(t["cov3D_precomp"]),
                                         settings_kw["viewmatrix"], settings_kw["projmatrix"], settings_kw["tanfovx"],
                                         settings_kw["tanfovy"], settings_kw["kernel_size"], case["H"], case["W"], o(t["shs"]),
                                         case["sh_degree"], settings_kw["campos"], False, case["require_coord"],
                                         case["require_depth"], False)
        R = int(nat[0])
        names = ["keys", "point_list", "ranges", "n_contrib", "tiles_touched", "depths", "means2D", "conic_opacity", "rgb",
                 "cov3D", "view_points", "camera_planes", "ray_planes", "ts", "normals"]
        out["R"] = R
        if case["P"] > 0:
            dec = native.decode_buffers(case["P"], 1, R, case["H"], case["W"], nat[9], nat[10], nat[11])
            out["decoded"] = {n: d.cpu().numpy() for n, d in zip(names, dec)}
    return out
