"""CPU: the host side of the batched renderer entry and the `.ply` I/O against what the REFERENCE'S
OWN code produced (tools/make_golden_renderer.py: unmodified igs/models/gs.py and igs/models/gaussian_model.py run in
the build container; the CUDA rasterizer module replaced by a recorder, so the golden holds exactly what
`GS3DRenderer.forward_single_batch` hands to the rasterizer for every view)."""
import math
from pathlib import Path

import numpy as np
import pytest

torch = pytest.importorskip("torch")

GOLDEN = Path(__file__).resolve().parent / "golden"


def _g():
    return np.load(GOLDEN / "ref_renderer.npz")


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def raw_inputs(g, dev="cpu"):
    gs = {k: _t(g["in_" + k]).to(dev) for k in ("xyz", "opacity", "rotation", "scaling", "shs")}
    resi = {k: _t(g["in_resi_" + k]).to(dev) for k in ("xyz", "rotation", "shs")}
    return gs, resi, _t(g["in_mask"]).to(dev), _t(g["in_c2ws"]).to(dev), tuple(float(x) for x in g["in_fov"])


def test_camera_from_c2w_matches_the_reference_class():
    from igs_b200 import renderer
    g = _g()
    _gs, _resi, _mask, c2ws, FOV = raw_inputs(g)
    H, W = (int(x) for x in g["in_res"])
    st = renderer.settings_from_c2ws(c2ws, FOV, (H, W), torch.ones(3))
    for v in range(c2ws.shape[0]):
        cam = renderer.Camera.from_c2w(c2ws[v], FOV, (H, W))
        for ours, key in ((cam.world_view_transform, "world_view"), (cam.projection_matrix, "proj"),
                          (cam.full_proj_transform, "full"), (cam.camera_center, "center")):
            assert torch.equal(ours, _t(g[f"cam{v}_{key}"])), (v, key)       # same torch ops: bit-identical
        # the batched construction (one inverse / matmul for all views) agrees to rounding
        assert torch.allclose(st.viewmatrices[v], _t(g[f"cam{v}_world_view"]), rtol=0, atol=1e-6)
        assert torch.allclose(st.projmatrices[v], _t(g[f"cam{v}_full"]), rtol=1e-6, atol=1e-6)
        assert torch.allclose(st.campos[v], _t(g[f"cam{v}_center"]), rtol=1e-6, atol=1e-6)


def test_rasterizer_arguments_match_what_the_reference_renderer_passes():
    from igs_b200 import renderer
    g = _g()
    gs, resi, mask, c2ws, FOV = raw_inputs(g)
    H, W = (int(x) for x in g["in_res"])
    FH, FW = (int(x) for x in g["in_flow_res"])
    d = renderer.deform(gs, resi, mask)
    for k in ("xyz", "opacity", "rotation", "scaling", "shs"):
        assert torch.allclose(d[k], _t(g["deformed_" + k]), rtol=1e-6, atol=1e-7), k
    a = renderer.activated(d)
    for ours, key in ((a["xyz"], "means3D"), (a["opacity"], "opacities"), (a["scaling"], "scales"), (a["rotation"], "rotations"),
                      (a["shs"], "shs")):
        assert torch.allclose(ours, _t(g["v0_rgb_" + key]), rtol=1e-6, atol=1e-7), key
    st = renderer.settings_from_c2ws(c2ws, FOV, (H, W), _t(g["in_bg"]))
    fl = renderer.flow_inputs(d)
    for ours, key in ((fl["xyz"], "means3D"), (fl["opacity"], "opacities"), (fl["scaling"], "scales"), (fl["rotation"], "rotations")):
        assert torch.allclose(ours, _t(g["v0_flow_" + key]), rtol=1e-6, atol=1e-7), key
    for v in range(c2ws.shape[0]):
        for kind, (h, w) in (("rgb", (H, W)), ("flow", (FH, FW))):
            pre = f"v{v}_{kind}_"
            assert (int(g[pre + "image_height"]), int(g[pre + "image_width"])) == (h, w)
            assert abs(float(g[pre + "tanfovx"]) - st.tanfovx) <= 1e-12 and abs(float(g[pre + "tanfovy"]) - st.tanfovy) <= 1e-12
            assert float(g[pre + "kernel_size"]) == st.kernel_size == 0.0 and float(g[pre + "scale_modifier"]) == st.scale_modifier
            assert int(g[pre + "sh_degree"]) == st.sh_degree
            assert bool(g[pre + "require_coord"]) and bool(g[pre + "require_depth"]) and not bool(g[pre + "prefiltered"])
            assert torch.allclose(st.viewmatrices[v], _t(g[pre + "viewmatrix"]), rtol=0, atol=1e-6)
            assert torch.allclose(st.projmatrices[v], _t(g[pre + "projmatrix"]), rtol=1e-6, atol=1e-6)
            assert torch.allclose(st.campos[v], _t(g[pre + "campos"]), rtol=1e-6, atol=1e-6)
        fx, fy = FW / (2 * st.tanfovx), FH / (2 * st.tanfovy)
        col = renderer.flow_colors(fl, st.viewmatrices[v], fx, fy)
        ref = _t(g[f"v{v}_flow_colors_precomp"])
        assert torch.allclose(col, ref, rtol=1e-5, atol=1e-6 * float(ref.abs().max())), v


def test_ply_reader_and_writers_match_the_reference(tmp_path):
    from igs_b200 import ply
    g = np.load(GOLDEN / "ref_ply.npz")
    # reader: a frame-0 style file with filter_3D -> what the reference's load_ply returned (3D filter folded in)
    got = ply.load_ply(GOLDEN / "ref_ply_frame0.ply")
    for k in ("xyz", "rotation", "shs"):
        assert torch.equal(got[k], _t(g["loaded_" + k])), k
    for k in ("opacity", "scaling"):
        assert torch.allclose(got[k], _t(g["loaded_" + k]), rtol=1e-6, atol=1e-6), k
    # reader: the file the reference's stream model wrote -> the values it was given
    got = ply.load_ply(GOLDEN / "ref_ply_stream.ply")
    for k in ("xyz", "opacity", "rotation", "scaling", "shs"):
        assert torch.equal(got[k], _t(g["in_" + k])), k
    # writer (default): byte-identical table to the reference's stream-model writer
    f = {k: _t(g["in_" + k]) for k in ("xyz", "opacity", "rotation", "scaling", "shs")}
    ply.save_ply(tmp_path / "a.ply", f["xyz"], f["shs"], f["opacity"], f["scaling"], f["rotation"])
    ta, tb = ply.read_vertex_table(tmp_path / "a.ply"), ply.read_vertex_table(GOLDEN / "ref_ply_stream.ply")
    assert list(ta) == list(tb)
    for k in ta:
        assert np.array_equal(ta[k], tb[k]), k
    # writer (normalize_rotation=True): the refine model's save_ply(save_type='all')
    ply.save_ply(tmp_path / "b.ply", f["xyz"], f["shs"], f["opacity"], f["scaling"], f["rotation"], normalize_rotation=True)
    ta, tb = ply.read_vertex_table(tmp_path / "b.ply"), ply.read_vertex_table(GOLDEN / "ref_ply_refine_all.ply")
    assert list(ta) == list(tb)
    for k in ta:
        if k == "opacity" or k.startswith("scale_"):     # the reference writes inverse_activation(activation(x))
            assert np.allclose(ta[k], tb[k], rtol=1e-5, atol=1e-5), k
        else:
            assert np.array_equal(ta[k], tb[k]), k
    assert abs(float(np.sqrt(sum(tb[f"rot_{i}"] ** 2 for i in range(4))).max()) - 1.0) <= 1e-6
