"""GPU: full-size parity (tests/fullsize.py: C2 200k/512^2, a C3 view 300k/1352x1014, and a batched 3-view C3 call
that takes the multi-launch sort path) against the UNMODIFIED reference CUDA build.

  * live, when baseline/_ref travelled to the box: every integer output bit-exact per view, the seven maps
    <= 1e-4 max-abs over every pixel, the six gradients <= 1e-3 max-normalised AND <= 1e-4 relative L2 over every element
    (reference = mean of three runs);
  * always: the same integer outputs against the committed sha256 digests of the reference's, and the float outputs
    against the committed probes (tools/make_golden_fullsize.py -> tests/golden/ref_fullsize_*.npz), so nothing is
    skipped on a box without the reference build.
Reference: DGR/cuda_rasterizer/rasterizer_impl.cu:190-250 (carve), :373-390 (sort + ranges)."""
import numpy as np
import pytest
import torch

import fullsize
from tolerances import GRAD_RL2, GRAD_RTOL, MAP_ATOL

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=list(fullsize.CASES))
def ours(request, cuda_dev, native):
    inp = fullsize.inputs(request.param, cuda_dev)
    out = fullsize.run_ours(inp, native)
    torch.cuda.synchronize()
    return request.param, inp, out


def _masked(name, t, inp):
    if name != "rotations":
        return t
    m = fullsize.rotation_mask(inp)
    return t if isinstance(m, slice) else t[torch.from_numpy(m).to(t.device)]


def test_integer_outputs_match_committed_reference_digests(ours):
    name, inp, out = ours
    f = fullsize.golden_path(name)
    assert f.exists(), f"{f} missing: run tools/make_golden_fullsize.py on a box with baseline/_ref"
    g = np.load(f, allow_pickle=False)
    assert out["R"] == int(g["R"])
    for n in fullsize.INT_NAMES:
        for v, t in enumerate(out["ints"][n]):
            assert fullsize.digest(t) == str(g["sha_" + n][v]), (name, n, v)


def test_float_outputs_match_committed_reference_probes(ours):
    name, inp, out = ours
    g = np.load(fullsize.golden_path(name), allow_pickle=False)
    V = len(inp["views"])
    pix = torch.from_numpy(g["probe_pix"]).to(inp["means3D"].device)
    for n in fullsize.MAP_NAMES:
        m = out["maps"][n]
        got = m.reshape(V, m.shape[1], -1)[:, :, pix].cpu().numpy()
        assert np.abs(got - g["map_" + n]).max() <= MAP_ATOL, (name, n)
    gi = torch.from_numpy(g["probe_gauss"]).to(inp["means3D"].device)
    rm = fullsize.rotation_mask(inp)
    for n in fullsize.GRAD_NAMES:
        t = out["grads"][n]
        got = (t[:, gi] if n == "means2D" else t[gi]).cpu().numpy().astype(np.float64)
        ref = g["grad_" + n].astype(np.float64)
        if n == "rotations" and not isinstance(rm, slice):
            keep = rm[g["probe_gauss"]]
            got, ref = got[keep], ref[keep]
        assert np.abs(got - ref).max() / float(g["gradmax_" + n]) <= GRAD_RTOL, (name, n)
        # probe rows against the probe rows' own norm (a subset: looser than the full-tensor bar of the live test)
        assert np.linalg.norm(got - ref) / (np.linalg.norm(ref) + 1e-30) <= 10 * GRAD_RL2, (name, n)


def test_live_against_reference_build(ours):
    import ref_loader
    ref = ref_loader.load_reference()
    if ref is None:
        pytest.skip("baseline/_ref (reference CUDA build) did not travel to this box; digests / probes were asserted instead")
    name, inp, out = ours
    # The reference accumulates gradients with atomics: on the rotation gradient two single runs of the UNMODIFIED
    # reference differ by up to 1.1e-3 max-normalised at C3 and two 3-run means still by 3e-4 (tools/grad_noise_floor.py,
    # profiles/r02_grad_noise_floor.log), which made the 1e-3 bar a matter of luck with a 3-run mean.  The reference
    # value is therefore the mean of 12 runs (a better estimate of the same quantity; the bar itself is unchanged).
    r = fullsize.run_ref(ref, inp, grad_runs=12)
    assert out["R"] == r["R"]
    for n in fullsize.INT_NAMES:
        for v, (a, b) in enumerate(zip(out["ints"][n], r["ints"][n])):
            assert a.shape == b.shape and bool(torch.equal(a, b)), (name, n, v)
    for n in fullsize.MAP_NAMES:
        d = float((out["maps"][n] - r["maps"][n]).abs().max())
        assert d <= MAP_ATOL, (name, n, d)
    for n in fullsize.GRAD_NAMES:
        a = _masked(n, out["grads"][n], inp).double()
        b = _masked(n, r["grads"][n], inp).double()
        assert bool(torch.isfinite(a).all()), (name, n)
        e_max = float((a - b).abs().max() / (b.abs().max() + 1e-30))
        e_l2 = float((a - b).norm() / (b.norm() + 1e-30))
        print(f"[{name}] grad {n:10s} max-norm {e_max:.3e} rel-l2 {e_l2:.3e}")
        assert e_max <= GRAD_RTOL, (name, n, e_max)
        assert e_l2 <= GRAD_RL2, (name, n, e_l2)
