"""Generates tests/golden/ref_renderer.npz and tests/golden/ref_ply_*.{ply,npz} by running the reference's OWN code
(/root/reference/igs/models/gs.py and igs/models/gaussian_model.py, unmodified) in the build container (CPU):

  * `Camera.from_c2w` (gs.py:90-113) on seeded camera-to-world matrices: world_view_transform, projection_matrix,
    full_proj_transform, camera_center;
  * `GaussianModel.deform` (gs.py:347-375) + `GS3DRenderer.forward_single_batch` (gs.py:824-856) ->
    `forward_single_view` (gs.py:566-713) with `render_flow` on.  The CUDA rasterizer module it imports is replaced by
    a RECORDER (tools/ref_import.py): the golden holds exactly what the reference hands to the rasterizer for every
    view — the 15 settings fields and the (activated) tensors, for the colour render and for the flow render.  On the
    GPU the tests feed these recorded arguments to the reference CUDA build (baseline/_ref; the `_clamp` package's
    CUDA is byte-identical) and compare with this repo's batched `renderer.forward_single_batch` on the raw inputs;
  * `.ply` I/O: the stream model's `save_ply` (gs.py:317-342), the refine model's `save_ply(save_type='all')`
    (gaussian_model.py:177-212, normalised rotation) and `load_ply` with the 3D-filter fold (gs.py:400-460,476-490).
    `plyfile` is absent from the image; tools/plyfile_min.py stands in for the few calls made (same PLY container).

Run here (needs /root/reference):  python tools/make_golden_renderer.py
"""
from __future__ import annotations

import math
import os
import sys
import types
from pathlib import Path

os.environ.setdefault("TORCHDYNAMO_DISABLE", "1")   # the reference decorates helpers with @torch.compile; run them eagerly here

import numpy as np  # noqa: E402
import torch  # noqa: E402

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))
import ref_import  # noqa: E402
from igs_b200 import scenes  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"
P, V, W, H, FW, FH = 1500, 3, 96, 72, 48, 36


def renderer_golden(gs_mod):
    sc = scenes.make_gaussians(P, 21, math.log(0.25))
    cams = scenes.make_cameras(V, W, H, 22)
    c2ws = torch.inverse(cams.viewmatrix.transpose(1, 2)).contiguous()
    FOV = (2 * math.atan(cams.tanfovx), 2 * math.atan(cams.tanfovy))
    g = torch.Generator().manual_seed(23)
    mask = torch.sort(torch.randperm(P, generator=g)[:400]).values
    resi = dict(xyz=0.02 * torch.randn(400, 3, generator=g), rotation=torch.nn.functional.normalize(
        torch.tensor([1.0, 0, 0, 0]) + 0.05 * torch.randn(400, 4, generator=g)), shs=0.02 * torch.randn(400, 48, generator=g))
    bg = torch.tensor([1.0, 1.0, 1.0])
    out = dict(in_xyz=sc.xyz, in_opacity=sc.opacity_logit, in_rotation=sc.rot, in_scaling=sc.log_scale, in_shs=sc.sh,
               in_c2ws=c2ws, in_fov=torch.tensor(FOV, dtype=torch.float64), in_mask=mask, in_resi_xyz=resi["xyz"], in_resi_rotation=resi["rotation"],
               in_resi_shs=resi["shs"], in_bg=bg, in_res=torch.tensor([H, W]), in_flow_res=torch.tensor([FH, FW]))
    # ---- Camera.from_c2w ----
    for v in range(V):
        cam = gs_mod.Camera.from_c2w(c2ws[v], FOV, (H, W))
        out[f"cam{v}_world_view"] = cam.world_view_transform
        out[f"cam{v}_proj"] = cam.projection_matrix
        out[f"cam{v}_full"] = cam.full_proj_transform
        out[f"cam{v}_center"] = cam.camera_center
    # ---- forward_single_batch with the recording rasterizer ----
    r = gs_mod.GS3DRenderer.__new__(gs_mod.GS3DRenderer)
    torch.nn.Module.__init__(r)
    r.cfg = types.SimpleNamespace(lbs=False, ret_rgb=True, render_flow=True, flow_height=FH, flow_width=FW, scaling_modifier=1.0,
                                  sh_degree=3, use_rgb=False)
    r.device = torch.device("cpu")
    gauss = gs_mod.GaussianModel(sc.xyz, sc.opacity_logit, sc.rot, sc.log_scale, sc.sh)
    ref_import.RecordingRasterizer.calls.clear()
    res = r.forward_single_batch(resi, gauss, c2ws, FOV, (H, W), bg, mask)
    assert res["images_pred"].shape == (V, 3, H, W) and res["flow_pred"].shape == (V, 2, FH, FW)
    calls = ref_import.RecordingRasterizer.calls
    assert len(calls) == 2 * V
    for v in range(V):
        for kind, (st, kw) in (("rgb", calls[2 * v]), ("flow", calls[2 * v + 1])):
            pre = f"v{v}_{kind}_"
            for k in ("image_height", "image_width", "tanfovx", "tanfovy", "scale_modifier", "sh_degree", "kernel_size"):
                out[pre + k] = torch.tensor(float(st[k]), dtype=torch.float64)
            for k in ("prefiltered", "debug", "require_coord", "require_depth"):
                out[pre + k] = torch.tensor(bool(st[k]))
            for k in ("bg", "viewmatrix", "projmatrix", "campos"):
                out[pre + k] = st[k]
            for k in ("means3D", "opacities", "scales", "rotations", "shs", "colors_precomp"):
                if kw.get(k) is None:
                    continue
                # the Gaussians handed over are the same for every view (only the flow colours are per view): stored once
                first = out.get(f"v0_{kind}_{k}")
                if v > 0 and k != "colors_precomp":
                    assert torch.equal(first, kw[k].detach()), (v, kind, k)
                    continue
                out[pre + k] = kw[k].detach().clone()
            assert kw.get("cov3D_precomp") is None
    d = res["3dgs"]
    for k in ("xyz", "opacity", "rotation", "scaling", "shs"):
        out["deformed_" + k] = getattr(d, k)
    np.savez_compressed(GOLDEN / "ref_renderer.npz", **{k: (v.detach().numpy() if torch.is_tensor(v) else v) for k, v in out.items()})
    print("renderer golden:", len(out), "arrays;", len(calls), "recorded rasterizer calls")


def ply_golden(gs_mod):
    g = torch.Generator().manual_seed(31)
    n = 64
    f = dict(xyz=torch.randn(n, 3, generator=g), opacity=torch.randn(n, 1, generator=g), rotation=torch.randn(n, 4, generator=g),
             scaling=torch.randn(n, 3, generator=g) - 3.0, shs=torch.randn(n, 16, 3, generator=g))
    out = {"in_" + k: v.numpy() for k, v in f.items()}
    # (1) stream model writer (raw fields)
    gs_mod.GaussianModel(f["xyz"], f["opacity"], f["rotation"], f["scaling"], f["shs"]).save_ply(str(GOLDEN / "ref_ply_stream.ply"))
    # (2) refine model writer, save_type='all' (inverse activations of the activations; rotation NORMALISED)
    gm = ref_import.load_gaussian_model()
    with ref_import.cuda_as_cpu():
        m = gm.GaussianModel(3)
        stream = ref_import.AttrDict({k: v.clone() for k, v in f.items()}, mask=None)
        item = ref_import.AttrDict(use_new_shs=False, use_mask=False, no_shs=False, no_opacity=False, no_scaling=False)
        m.load_fromstream(stream, dict(position_lr_init=0.0016, rotation_lr=0.01, feature_lr=0.0025, opacity_lr=0.05, scaling_lr=0.005), item)
        m.save_ply(str(GOLDEN / "ref_ply_refine_all.ply"), save_type="all")
    # (3) a frame-0 style file WITH filter_3D (laid out with the plyfile stand-in) -> the reference loader's result
    import plyfile_min
    names = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] + [f"f_rest_{i}" for i in range(45)] + \
            ["opacity"] + [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)] + ["filter_3D"]
    tab = np.empty(n, dtype=[(k, "f4") for k in names])
    cols = torch.randn(n, len(names), generator=g).numpy().astype(np.float32)
    cols[:, names.index("filter_3D")] = np.abs(cols[:, names.index("filter_3D")]) * 0.02 + 0.001
    for i, k in enumerate(names):
        if k.startswith("scale_"):
            cols[:, i] = cols[:, i] * 0.5 - 3.0
        tab[k] = cols[:, i]
    plyfile_min.PlyData([plyfile_min.PlyElement.describe(tab, "vertex")]).write(str(GOLDEN / "ref_ply_frame0.ply"))
    loaded = gs_mod.load_ply(str(GOLDEN / "ref_ply_frame0.ply"))
    for k in ("xyz", "opacity", "rotation", "scaling", "shs"):
        out["loaded_" + k] = getattr(loaded, k).numpy()
    np.savez_compressed(GOLDEN / "ref_ply.npz", **out)
    print("ply golden written")


if __name__ == "__main__":
    gs_mod = ref_import.load_gs_module()
    renderer_golden(gs_mod)
    ply_golden(gs_mod)
