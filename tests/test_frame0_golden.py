"""CPU: torch restatements of the RaDe-GS frame-0 helpers (igs_b200/frame0.py) against values produced by the
reference's own code (tests/golden/refmath_frame0.npz, generator tools/make_golden_refmath_frame0.py)."""
import math

import numpy as np
import torch

import cases


def _g():
    return dict(np.load(cases.GOLDEN_DIR / "refmath_frame0.npz"))


def test_depth_normal_loss_restatement_matches_reference(native):
    from igs_b200 import frame0
    g = _g()
    for name in ("a", "b"):
        nrm = torch.from_numpy(g[f"{name}_normal"]).requires_grad_(True)
        H, W = nrm.shape[1:]
        tfx, tfy = (math.tan(float(f) / 2) for f in g[f"{name}_fov"])
        d1 = torch.from_numpy(g[f"{name}_depth1"]).requires_grad_(True)
        d2 = torch.from_numpy(g[f"{name}_depth2"]).requires_grad_(True)
        loss = frame0.depth_normal_loss_torch(nrm, d1, d2, W, H, tfx, tfy)
        loss.backward()
        assert abs(float(loss) - float(g[f"{name}_loss"])) <= 1e-6
        for t, k in ((d1, "d_depth1"), (d2, "d_depth2"), (nrm, "d_normal")):
            ref = g[f"{name}_{k}"]
            assert np.abs(t.grad.numpy() - ref).max() <= 1e-5 * np.abs(ref).max() + 1e-9, k
        p1 = torch.from_numpy(g[f"{name}_coord1"]).requires_grad_(True)
        p2 = torch.from_numpy(g[f"{name}_coord2"]).requires_grad_(True)
        loss = frame0.depth_normal_loss_torch(nrm.detach(), p1, p2, W, H, tfx, tfy, use_coord=True)
        loss.backward()
        assert abs(float(loss) - float(g[f"{name}_closs"])) <= 1e-6
        for t, k in ((p1, "d_coord1"), (p2, "d_coord2")):
            ref = g[f"{name}_{k}"]
            assert np.abs(t.grad.numpy() - ref).max() <= 1e-5 * np.abs(ref).max() + 1e-9, k


def test_3d_filter_matches_reference(native):
    from igs_b200 import frame0
    g = _g()
    xyz = torch.from_numpy(g["f_xyz"])
    f3d = frame0.compute_3D_filter(xyz, torch.from_numpy(g["f_viewmatrices"]), 64, 48, float(g["f_tanfov"][0]), float(g["f_tanfov"][1]))
    assert np.abs(f3d.numpy() - g["f_filter_3D"]).max() <= 1e-6 * np.abs(g["f_filter_3D"]).max()
    sc, op = frame0.scaling_n_opacity_with_3D_filter(torch.from_numpy(g["f_log_scale"]), torch.from_numpy(g["f_opacity_logit"]),
                                                     torch.from_numpy(g["f_filter_3D"]))
    assert np.abs(sc.numpy() - g["f_scales"]).max() <= 1e-6 * np.abs(g["f_scales"]).max()
    assert np.abs(op.numpy() - g["f_opacity"]).max() <= 1e-6


def test_sh_degree_ramp():
    from igs_b200 import frame0
    assert [frame0.active_sh_degree(i) for i in (1, 999, 1000, 2500, 3000, 29999)] == [0, 0, 1, 2, 3, 3]
