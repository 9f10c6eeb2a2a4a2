"""GPU: the CUDA path against the CPU oracle on seeded scenes the oracle finishes in seconds (C1: 10k Gaussians,
256x256; plus the multi-view batched call), and the reference CUDA build itself when baseline/_ref travelled."""
import numpy as np
import pytest
import torch

import cases
import oracle_py
from gpu_util import run_case
from tolerances import GRAD_RTOL, MAP_ATOL, grad_err, well_conditioned_rotation_mask

pytestmark = pytest.mark.gpu


def _c1_case(view=0, P=10_000, W=256, H=256, lsm=-2.2):
    c = cases._base(P, W, H, 0, lsm, view=view, V=3)
    return c


@pytest.mark.parametrize("view", [0, 2])
def test_c1_against_oracle(cuda_dev, native, view):
    import diff_gaussian_rasterization_rade as dgr
    case = _c1_case(view)
    cot = cases.cotangents(case)
    out = run_case(dgr, case, cuda_dev, cot=cot, which="all", native=native)
    fw = oracle_py.forward_case(case)
    assert out["R"] == fw["R"]
    np.testing.assert_array_equal(out["radii"], fw["geom"]["radii"])
    d = out["decoded"]
    np.testing.assert_array_equal(d["keys"].astype(np.uint64), fw["keys"])
    np.testing.assert_array_equal(d["point_list"].astype(np.uint32), fw["point_list"])
    np.testing.assert_array_equal(d["ranges"].astype(np.uint32), fw["ranges"])
    np.testing.assert_array_equal(d["n_contrib"].reshape(2, -1).astype(np.uint32), fw["image"]["n_contrib"])
    for n, m in zip(cases.MAP_NAMES, out["maps"]):
        assert np.abs(m.reshape(-1) - fw["image"][n].reshape(-1)).max() <= MAP_ATOL, n
    names = ["dL_color", "dL_coord", "dL_mcoord", "dL_depth", "dL_mdepth", "dL_alpha", "dL_normal"]
    gr = oracle_py.backward(fw, **dict(zip(names, cot)))
    for k in ("means3D", "means2D", "sh", "opacities", "scales", "rotations"):
        got, ref = out["grads"][k], gr[k].reshape(out["grads"][k].shape)
        if k == "rotations":
            m = well_conditioned_rotation_mask(case["scales"])
            got, ref = got[m], ref[m]
        assert grad_err(got, ref) <= GRAD_RTOL, (k, grad_err(got, ref))


def test_multiview_equals_stacked_single_views(cuda_dev, native):
    import igs_b200
    from igs_b200 import scenes
    from igs_b200.multiview import MultiViewSettings, rasterize_views
    P, W, H, V = 20000, 200, 152, 3
    dev = cuda_dev
    sc = scenes.make_gaussians(P, 0, -2.0)
    cams = scenes.make_cameras(V, W, H, 1)
    scales, rot, opac = scenes.activated(sc)
    base = [sc.xyz.to(dev), sc.sh.to(dev), opac.to(dev), scales.to(dev), rot.to(dev)]
    bg = torch.zeros(3, device=dev)
    cot = scenes.make_cotangents(V, H, W, 3)[0].to(dev)
    leaves = [b.clone().requires_grad_(True) for b in base]
    st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, bg, 1.0, cams.viewmatrix.to(dev), cams.projmatrix.to(dev), 3,
                           cams.campos.to(dev))
    outs = rasterize_views(*leaves, st)
    (outs[0] * cot).sum().backward()
    leaves2 = [b.clone().requires_grad_(True) for b in base]
    singles = []
    for v in range(V):
        s1 = igs_b200.GaussianRasterizationSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, bg, 1.0, cams.viewmatrix[v].to(dev),
                                                    cams.projmatrix[v].to(dev), 3, cams.campos[v].to(dev), False, True, True, False)
        m2 = torch.zeros(P, 3, device=dev, requires_grad=True)
        o = igs_b200.GaussianRasterizer(s1)(means3D=leaves2[0], means2D=m2, shs=leaves2[1], opacities=leaves2[2],
                                            scales=leaves2[3], rotations=leaves2[4])
        singles.append(o)
        (o[0] * cot[v]).sum().backward()
    for i in range(8):
        assert torch.equal(outs[i], torch.stack([s[i] for s in singles])), i
    for a, b in zip(leaves, leaves2):
        assert float((a.grad - b.grad).abs().max()) <= 1e-4 * float(b.grad.abs().max())


def test_against_reference_cuda_build_when_present(cuda_dev, native):
    import ref_loader
    ref = ref_loader.load_reference()
    if ref is None:
        pytest.skip("baseline/_ref (reference CUDA build) did not travel to this box")
    import diff_gaussian_rasterization_rade as dgr
    case = _c1_case(1)
    cot = cases.cotangents(case)
    a = run_case(dgr, case, cuda_dev, cot=cot, which="color")
    acc = None
    for _ in range(3):
        b = run_case(ref, case, cuda_dev, cot=cot, which="color")
        acc = b["grads"] if acc is None else {k: acc[k] + b["grads"][k] for k in acc}
    np.testing.assert_array_equal(a["radii"], b["radii"])
    for n, x, y in zip(cases.MAP_NAMES, a["maps"], b["maps"]):
        assert np.abs(x - y).max() <= MAP_ATOL, n
    for k in ("means3D", "means2D", "sh", "opacities", "scales", "rotations"):
        ref_g = acc[k] / 3
        got = a["grads"][k]
        if k == "rotations":
            m = well_conditioned_rotation_mask(case["scales"])
            got, ref_g = got[m], ref_g[m]
        assert grad_err(got, ref_g) <= GRAD_RTOL, k
