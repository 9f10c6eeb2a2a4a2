"""CPU: the oracle (oracle/raster_oracle.c) against golden vectors produced by the UNMODIFIED reference
CUDA rasterizer on a B200 (tests/golden/ref_*.npz, generator tools/make_golden_ref.py).  This is what
pins the oracle: the reference's own tests hold no vectors for this path."""
import numpy as np
import pytest

import cases
import oracle_py
from tolerances import GRAD_RTOL, MAP_ATOL, grad_err, well_conditioned_rotation_mask

NAMES = list(cases.CASES)


@pytest.fixture(scope="module", params=NAMES)
def pair(request):
    g = cases.load_golden(request.param)
    assert g is not None, f"missing golden for {request.param}"
    case = cases.case_from_golden(g)
    return request.param, g, case, oracle_py.forward_case(case)


def test_inputs_match_case_definition(pair):
    name, g, case, _ = pair
    fresh = cases.CASES[name]()
    for k in ("means3D", "opacities", "viewmatrix", "projmatrix"):
        np.testing.assert_array_equal(fresh[k], case[k])


def test_bit_exact_binning(pair):
    _, g, case, fw = pair
    assert fw["R"] == int(g["R"])
    np.testing.assert_array_equal(fw["geom"]["radii"], g["radii"])
    np.testing.assert_array_equal(fw["geom"]["tiles_touched"].astype(np.int64), g["tiles_touched"].astype(np.int64))
    np.testing.assert_array_equal(fw["keys"].astype(np.int64), g["point_list_keys"])
    np.testing.assert_array_equal(fw["point_list"].astype(np.int64), g["point_list"].astype(np.int64))
    np.testing.assert_array_equal(fw["ranges"].astype(np.int64), g["ranges"].astype(np.int64))
    np.testing.assert_array_equal(fw["image"]["n_contrib"], g["n_contrib"].astype(np.uint32))


def test_bit_exact_preprocess_scalars(pair):
    _, g, case, fw = pair
    vis = g["radii"] > 0
    for k in ("depths", "means2D", "conic_opacity"):
        a = fw["geom"][k][vis].view(np.uint32)
        b = g["geom_" + k][vis].view(np.uint32)
        assert (a != b).sum() == 0, k


def test_maps_within_1e4(pair):
    _, g, case, fw = pair
    for n in cases.MAP_NAMES:
        d = np.abs(fw["image"][n].reshape(-1) - g["map_" + n].reshape(-1)).max()
        assert d <= MAP_ATOL, (n, d)


@pytest.mark.parametrize("which", ["color", "all"])
def test_gradients_within_1e3(pair, which):
    name, g, case, fw = pair
    # "edge" holds Gaussians 1e-3 behind the near plane (magnification ~focal/0.2) whose gradients amplify the
    # last-ulp differences between glibc expf / CPU rounding and the reference's GPU arithmetic: the CPU oracle is
    # pinned at 2e-3 there; the CUDA path is held to 1e-3 against the same golden (test_parity_golden_gpu.py).
    rtol = 2e-3 if name == "edge" else GRAD_RTOL
    cot = cases.cotangents(case)
    names = ["dL_color", "dL_coord", "dL_mcoord", "dL_depth", "dL_mdepth", "dL_alpha", "dL_normal"]
    kw = dict(dL_color=cot[0]) if which == "color" else dict(zip(names, cot))
    gr = oracle_py.backward(fw, ref_bug=True, **kw)
    for k in cases.GRAD_NAMES:
        key = f"grad_{which}_{k}"
        if key not in g:
            continue
        ref, got = g[key], gr[k].reshape(g[key].shape)
        if k == "rotations":
            m = well_conditioned_rotation_mask(case["scales"], scale_modifier=case["scale_modifier"])
            ref, got = ref[m], got[m]
        assert grad_err(got, ref) <= rtol, (k, grad_err(got, ref))


def test_reference_bug_T19_switch_is_observable_only_with_kernel_size():
    """With kernel_size == 0 (what IGS passes) the mis-passed conic_opacity only scales a term that cancels;
    the two oracle modes must agree to rounding, which is why goldens cannot tell them apart there."""
    g = cases.load_golden("small")
    case = cases.case_from_golden(g)
    fw = oracle_py.forward_case(case)
    cot = cases.cotangents(case)
    a = oracle_py.backward(fw, dL_color=cot[0], ref_bug=True)
    b = oracle_py.backward(fw, dL_color=cot[0], ref_bug=False)
    for k in ("means3D", "scales", "rotations", "opacities"):
        assert grad_err(a[k], b[k]) < 1e-4, k
