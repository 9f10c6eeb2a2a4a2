"""GPU: the densify kernels (igs_densify_stats / igs_densify_gather) against their torch restatements, and a refine
loop that densifies in the middle (P changes, Adam state carried, iterations keep running)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_stats_kernel_matches_torch(cuda_dev, native):
    from igs_b200 import densify
    g = torch.Generator().manual_seed(5)
    P, V = 10007, 6
    g2d = torch.randn(V, P, 3, generator=g).to(cuda_dev)
    radii = torch.randint(-2, 9, (V, P), generator=g, dtype=torch.int32).clamp_min(0).to(cuda_dev)
    a, b = densify.DensifyStats(P, cuda_dev), densify.DensifyStats(P, cuda_dev)
    for _ in range(2):
        a.update(g2d, radii)
        b.update_torch(g2d, radii)
    assert torch.allclose(a.grad_accum, b.grad_accum, rtol=1e-6, atol=1e-7)
    assert torch.equal(a.denom, b.denom) and torch.equal(a.max_radii2D, b.max_radii2D)


def test_gather_kernel_matches_torch(cuda_dev, native):
    from igs_b200 import densify, refine
    g = torch.Generator().manual_seed(6)
    P = 20011
    xyz = torch.randn(P, 3, generator=g)
    scaling = torch.randn(P, 3, generator=g) * 0.8 + math.log(0.05)
    state = refine.RefineState(xyz, scaling, torch.randn(P, 4, generator=g), torch.randn(P, 1, generator=g) * 3,
                               torch.randn(P, 16, 3, generator=g), cuda_dev)
    state.exp_avg.copy_(torch.randn(P * 59, generator=g))
    state.exp_avg_sq.copy_(torch.rand(P * 59, generator=g))
    accum = (torch.rand(P, generator=g) * 4e-4).to(cuda_dev)
    denom = torch.randint(0, 4, (P,), generator=g).float().to(cuda_dev)
    gen = torch.Generator(device=cuda_dev).manual_seed(11)
    plan = densify.plan_densify(None, state.param("scaling"), state.param("opacity"), accum, denom, 0.00015, 0.005, 3.0,
                                None, 150000, 0.01, generator=gen)
    assert plan.n_clone > 0 and plan.n_split > 0 and plan.n_pruned > 0
    P_new = int(plan.src.numel())
    ref = densify.apply_plan_torch(state.flat, state.exp_avg, state.exp_avg_sq, P, plan)
    got = native.densify_gather(state.flat, state.exp_avg, state.exp_avg_sq, P, plan.src, plan.kind, plan.samples,
                                plan.split_div, densify._offsets(P), densify._offsets(P_new))
    assert torch.allclose(got[0], ref[0], rtol=2e-6, atol=2e-6)
    moved = (got[0] != ref[0])
    child = (plan.kind == 2)
    # only the split children's xyz / scaling may differ (rounding of the rotation product / log-exp)
    off = dict(zip(densify._ROLE_ORDER, densify._offsets(P_new)))
    mask = torch.zeros(P_new * 59, dtype=torch.bool, device=cuda_dev)
    mask[off["xyz"]:off["xyz"] + 3 * P_new] = child.repeat_interleave(3)
    mask[off["scaling"]:off["scaling"] + 3 * P_new] = child.repeat_interleave(3)
    assert not bool((moved & ~mask).any())
    assert torch.equal(got[1], ref[1]) and torch.equal(got[2], ref[2])


def test_refine_loop_with_densification(cuda_dev, native):
    from igs_b200 import densify, refine, scenes
    from igs_b200.multiview import MultiViewSettings
    P, V, W, H = 8000, 4, 160, 120
    sc = scenes.make_gaussians(P, 0, -1.8)
    cams = scenes.make_cameras(V, W, H, 1)
    st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=cuda_dev), 1.0,
                           cams.viewmatrix.to(cuda_dev).contiguous(), cams.projmatrix.to(cuda_dev).contiguous(), 3,
                           cams.campos.to(cuda_dev).contiguous())
    state = refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, cuda_dev)
    with torch.no_grad():
        target = refine.forward_views(state, st)[0].clone()
    gen = torch.Generator().manual_seed(3)
    state.param("shs").add_(0.3 * torch.randn(P, 16, 3, generator=gen).to(cuda_dev))
    stats = densify.DensifyStats(P, cuda_dev)
    losses, sizes = [], []
    dgen = torch.Generator(device=cuda_dev).manual_seed(1)
    for it in range(1, 31):
        r = refine.refine_iteration(state, st, target, V)
        losses.append(float(r.loss))
        stats.update(r.grad_means2D, r.radii)
        if it % 10 == 0:   # infer_batch.py:315-321 with densification_interval 10
            step_before = state.step
            plan = densify.densify_and_prune(state, stats, 1e-6, 0.005, 3.0, None, max_num=9000, percent_dense=0.01,
                                             generator=dgen)
            assert state.P == int(plan.src.numel()) and state.flat.numel() == state.P * 59 and state.step == step_before
            assert state.P <= 9000 + 2 * plan.n_split and stats.denom.numel() == state.P
            assert float(stats.denom.abs().max()) == 0.0
        sizes.append(state.P)
    assert len(set(sizes)) > 1, sizes                      # the number of Gaussians did change
    assert all(math.isfinite(x) for x in losses) and losses[-1] < losses[0]
    assert bool(torch.isfinite(state.flat).all())


@pytest.mark.parametrize("case", ["refine", "bounded", "cloned"])
def test_kernels_match_the_reference_class(cuda_dev, native, case):
    """igs_densify_stats + igs_densify_gather against what the reference's own GaussianModel produced
    (tests/golden/ref_densify_*.npz, tools/make_golden_densify.py).  The split samples are planned on the CPU so that
    the random stream is the reference's (global CPU generator)."""
    import numpy as np
    from igs_b200 import densify
    from test_densify_host import GOLDEN, _check_against_golden, _state_from_golden
    g = np.load(GOLDEN / f"ref_densify_{case}.npz")
    P = int(g["cfg_P"])
    stats = densify.DensifyStats(P, cuda_dev)
    for i in range(3):
        radii = torch.where(torch.from_numpy(g[f"stat{i}_vis"]), torch.from_numpy(g[f"stat{i}_radii"]), torch.zeros(P)).to(torch.int32)
        stats.update(torch.from_numpy(g[f"stat{i}_grad"]).unsqueeze(0).to(cuda_dev), radii.unsqueeze(0).to(cuda_dev))
    assert torch.allclose(stats.grad_accum.cpu(), torch.from_numpy(g["grad_accum"]).reshape(-1), rtol=1e-6, atol=0)
    assert torch.equal(stats.denom.cpu(), torch.from_numpy(g["denom"]).reshape(-1))
    assert torch.equal(stats.max_radii2D.cpu(), torch.from_numpy(g["max_radii2D"]).reshape(-1))
    st_cpu = _state_from_golden(g)
    plan = densify.plan_densify(None, st_cpu.param("scaling"), st_cpu.param("opacity"), torch.from_numpy(g["grad_accum"]).reshape(-1),
                                torch.from_numpy(g["denom"]).reshape(-1), float(g["cfg_thr"]), 0.005, float(g["cfg_extent"]), None,
                                int(g["cfg_max_num"]), float(g["cfg_percent_dense"]), True,
                                generator=torch.Generator().manual_seed(int(g["cfg_rng_seed"])))
    st = _state_from_golden(g, cuda_dev)
    P_new = int(plan.src.numel())
    flat, m, v = native.densify_gather(st.flat, st.exp_avg, st.exp_avg_sq, P, plan.src.to(cuda_dev), plan.kind.to(cuda_dev),
                                       plan.samples.to(cuda_dev), float(plan.split_div), densify._offsets(P), densify._offsets(P_new))
    st.resize(P_new, flat, m, v)
    _check_against_golden(st, g)


def test_batched_step_accumulates_the_same_statistics_as_single_view_steps(cuda_dev, native):
    """The reference feeds add_densification_stats the gradient of ONE view's undivided loss (infer_batch.py:302-313).
    A V-view step divides its loss by V; with grad_scale = V the accumulated statistics equal those of V single-view
    steps at the same parameters, so densify_grad_threshold keeps its meaning under the batched schedule."""
    from igs_b200 import densify, refine, scenes
    from igs_b200.multiview import MultiViewSettings
    from igs_b200.stream import select_views
    P, V, W, H = 6000, 3, 160, 120
    sc = scenes.make_gaussians(P, 0, -1.8)
    cams = scenes.make_cameras(V, W, H, 1)
    st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=cuda_dev), 1.0,
                           cams.viewmatrix.to(cuda_dev).contiguous(), cams.projmatrix.to(cuda_dev).contiguous(), 3,
                           cams.campos.to(cuda_dev).contiguous())
    gt = torch.rand(V, 3, H, W, generator=torch.Generator().manual_seed(9)).to(cuda_dev)
    state = refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, cuda_dev)
    a = densify.DensifyStats(P, cuda_dev)
    r = refine.refine_iteration(state, st, gt, V, do_step=False)
    a.update(r.grad_means2D, r.radii, grad_scale=float(V))
    b = densify.DensifyStats(P, cuda_dev)
    for v in range(V):
        r1 = refine.refine_iteration(state, select_views(st, [v]), gt[v:v + 1], 1, do_step=False)
        b.update(r1.grad_means2D, r1.radii)
    assert torch.equal(a.denom, b.denom) and torch.equal(a.max_radii2D, b.max_radii2D)
    assert float(a.denom.max()) == V
    assert torch.allclose(a.grad_accum, b.grad_accum, rtol=1e-4, atol=1e-6 * float(b.grad_accum.max()))
