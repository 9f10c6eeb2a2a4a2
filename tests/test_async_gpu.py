"""GPU: the capacity mode of the forward (no host read of the instance count; reference's blocking cudaMemcpy is
DGR/cuda_rasterizer/rasterizer_impl.cu:350-358) — equal to the exact mode, overflow is flagged and guards the Adam step
on the device, and a whole refine iteration can be captured in a CUDA graph."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(dev, P=20000, V=3, W=200, H=152):
    from igs_b200 import refine, scenes
    from igs_b200.multiview import MultiViewSettings
    sc = scenes.make_gaussians(P, 0, -2.0)
    cams = scenes.make_cameras(V, W, H, 1)
    st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=dev), 1.0,
                           cams.viewmatrix.to(dev).contiguous(), cams.projmatrix.to(dev).contiguous(), 3,
                           cams.campos.to(dev).contiguous())
    gt = torch.rand(V, 3, H, W, generator=torch.Generator().manual_seed(4)).to(dev)

    def fresh():
        return refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, dev)
    return st, gt, fresh


def test_capacity_mode_equals_exact_mode(cuda_dev, native):
    from igs_b200 import refine
    st, gt, fresh = _setup(cuda_dev)
    a, b = fresh(), fresh()
    V = gt.shape[0]
    ra = refine.refine_iteration(a, st, gt, V)
    b.enable_async_binning()
    refine.refine_iteration(b, st, gt, V)            # first iteration: exact (learns the count), identical by construction
    b.flat.copy_(fresh().flat); b.exp_avg.zero_(); b.exp_avg_sq.zero_(); b.step = 0
    assert b.instance_capacity(V) >= ra.num_rendered > 0
    rb = refine.refine_iteration(b, st, gt, V)       # capacity mode
    assert rb.num_rendered == b.instance_capacity(V)  # the capacity is what is handed around as R
    b.poll_status(wait=True)
    assert b.overflows == 0 and b._r_seen[V] == ra.num_rendered
    assert torch.equal(ra.loss, rb.loss) and torch.equal(a.grad, b.grad) and torch.equal(a.flat, b.flat)
    assert torch.equal(ra.radii, rb.radii) and torch.equal(ra.grad_means2D, rb.grad_means2D)


def test_overflow_is_flagged_and_skips_the_adam_step(cuda_dev, native):
    from igs_b200 import refine
    st, gt, fresh = _setup(cuda_dev)
    s = fresh()
    V = gt.shape[0]
    R = refine.refine_iteration(s, st, gt, V, do_step=False).num_rendered
    s.enable_async_binning()
    s._r_seen[V] = R // 3                            # pretend a much smaller scene was seen: capacity < R
    before, m_before, step_before = s.flat.clone(), s.exp_avg.clone(), s.step
    refine.refine_iteration(s, st, gt, V)
    torch.cuda.synchronize()
    assert int(s.status[0]) == R and int(s.status[1]) == 1
    assert torch.equal(s.flat, before) and torch.equal(s.exp_avg, m_before)   # guarded on the device
    assert bool(torch.isfinite(s.grad).all())
    s.poll_status(wait=True)
    assert s.overflows == 1 and s._r_seen[V] == R     # the true count was learnt: the next capacity fits
    assert s.instance_capacity(V) >= R


def test_iteration_in_a_cuda_graph(cuda_dev, native):
    from igs_b200 import refine
    st, gt, fresh = _setup(cuda_dev, V=1)
    eager, graphed = fresh(), fresh()
    R = refine.refine_iteration(fresh(), st, gt, 1).num_rendered
    losses = [refine.refine_iteration(eager, st, gt, 1).loss.clone() for _ in range(4)]
    g = refine.GraphedIteration(graphed, st, gt, 1, capacity=int(R * 1.5))
    got = [g.run(st.viewmatrices, st.projmatrices, st.campos, gt).clone() for _ in range(4)]
    assert g.finish() == 0 and graphed.step == 4
    # the graph computes Adam's bias corrections on the device (double pow), the eager step on the host: last-bit room
    for a, b in zip(losses, got):
        assert abs(float(a) - float(b)) <= 1e-6 * abs(float(a))
    assert float((eager.flat - graphed.flat).abs().max()) <= 1e-6 * float(eager.flat.abs().max())
    assert torch.allclose(eager.exp_avg_sq, graphed.exp_avg_sq, rtol=1e-5, atol=0)
