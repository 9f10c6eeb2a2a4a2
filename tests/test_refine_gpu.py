"""GPU: the pieces of the refine iteration around the rasterizer — fused L1+SSIM loss kernel vs the reference's
loss_utils golden and vs torch at image size, fused Adam vs torch.optim.Adam(eps=1e-15), and the iteration itself
(loss decreases; sharded-by-view gradient sum equals the single batch)."""
import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu


def test_fused_loss_matches_reference_golden(cuda_dev, native):
    g = dict(np.load(cases.GOLDEN_DIR / "refmath_loss.npz"))
    for name in ("a", "b"):
        pred = torch.from_numpy(g[f"{name}_pred"]).unsqueeze(0).to(cuda_dev)
        gt = torch.from_numpy(g[f"{name}_gt"]).unsqueeze(0).to(cuda_dev)
        loss, dpred = native.refine_loss(pred, gt, 0.8)
        assert abs(float(loss) - float(g[f"{name}_loss"])) <= 2e-6
        ref = g[f"{name}_dpred"]
        assert np.abs(dpred[0].cpu().numpy() - ref).max() <= 1e-3 * np.abs(ref).max()


def test_fused_loss_matches_torch_at_image_size(cuda_dev, native):
    from igs_b200 import refine
    gen = torch.Generator().manual_seed(5)
    V, H, W = 2, 1014, 1352
    pred = torch.rand(V, 3, H, W, generator=gen).to(cuda_dev).requires_grad_(True)
    gt = torch.rand(V, 3, H, W, generator=gen).to(cuda_dev)
    ref = refine.refine_loss_torch(pred, gt, 0.8).sum()
    ref.backward()
    loss, dpred = native.refine_loss(pred.detach(), gt, 0.8)
    assert abs(float(loss) - float(ref)) <= 1e-5 * abs(float(ref))
    assert float((dpred - pred.grad).abs().max()) <= 1e-3 * float(pred.grad.abs().max())
    # through the autograd wrapper, scaled
    p2 = pred.detach().clone().requires_grad_(True)
    (refine.refine_loss_sum(p2, gt, 0.8, "fused") / 7.0).backward()
    assert float((p2.grad - pred.grad / 7.0).abs().max()) <= 1e-3 * float(pred.grad.abs().max()) / 7.0


@pytest.mark.parametrize("H,W", [(70, 93), (33, 40), (129, 258)])
def test_fused_loss_odd_sizes_match_torch(cuda_dev, native, H, W):
    """Widths that are not a multiple of 4 take the scalar staging path; ragged right/bottom tiles either way."""
    from igs_b200 import refine
    gen = torch.Generator().manual_seed(H * 1000 + W)
    pred = torch.rand(2, 3, H, W, generator=gen).to(cuda_dev).requires_grad_(True)
    gt = torch.rand(2, 3, H, W, generator=gen).to(cuda_dev)
    ref = refine.refine_loss_torch(pred, gt, 0.8).sum()
    ref.backward()
    loss, dpred = native.refine_loss(pred.detach(), gt, 0.8, 0.5)
    assert abs(float(loss) - float(ref)) <= 1e-5 * abs(float(ref))
    assert float((dpred - 0.5 * pred.grad).abs().max()) <= 1e-3 * float(pred.grad.abs().max()) * 0.5


def test_fused_adam_matches_torch_adam(cuda_dev, native):
    gen = torch.Generator().manual_seed(9)
    n = 100_003
    p0 = torch.randn(n, generator=gen).to(cuda_dev)
    a = p0.clone()
    m = torch.zeros_like(a)
    v = torch.zeros_like(a)
    b = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([b], lr=0.0016, eps=1e-15)
    for step in range(1, 6):
        gr = torch.randn(n, generator=gen).to(cuda_dev) * (10.0 ** (step - 3))
        native.adam_step(a, gr, m, v, 0.0016, 0.9, 0.999, 1e-15, step)
        b.grad = gr.clone()
        opt.step()
        assert float((a - b.detach()).abs().max()) <= 1e-6


def _state_and_settings(dev, P, V, W, H, views=None):
    from igs_b200 import refine, scenes
    from igs_b200.multiview import MultiViewSettings
    sc = scenes.make_gaussians(P, 0, -1.8)
    cams = scenes.make_cameras(V, W, H, 1)
    views = list(range(V)) if views is None else views
    st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=dev), 1.0,
                           cams.viewmatrix[views].to(dev).contiguous(), cams.projmatrix[views].to(dev).contiguous(), 3,
                           cams.campos[views].to(dev).contiguous())
    state = refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, dev)
    return state, st


@pytest.mark.parametrize("loss_impl", ["torch", "fused"])
def test_refine_iterations_reduce_the_loss(cuda_dev, native, loss_impl):
    from igs_b200 import refine
    P, V, W, H = 20000, 4, 160, 120
    state, st = _state_and_settings(cuda_dev, P, V, W, H)
    with torch.no_grad():
        target = refine.forward_views(state, st)[0].clone()
    # perturb the colours and positions, then refine back towards the target renders
    gen = torch.Generator().manual_seed(3)
    state.param("shs").add_(0.3 * torch.randn(P, 16, 3, generator=gen).to(cuda_dev))
    losses = []
    for _ in range(25):
        r = refine.refine_iteration(state, st, target, V, loss_impl=loss_impl)
        losses.append(float(r.loss))
    assert losses[-1] < 0.7 * losses[0], losses
    assert r.radii.shape == (V, P) and r.num_rendered > 0


def test_sharded_gradients_equal_single_batch(cuda_dev, native):
    """What N ranks + all-reduce compute (each rank its views, loss scaled by 1/total_views) summed == one rank
    rendering all views.  The collective itself is covered by tests/test_dist_gloo.py."""
    from igs_b200 import refine
    P, V, W, H = 20000, 5, 160, 120
    state, st_all = _state_and_settings(cuda_dev, P, V, W, H)
    gen = torch.Generator().manual_seed(4)
    gt = torch.rand(V, 3, H, W, generator=gen).to(cuda_dev)
    r_all = refine.refine_iteration(state, st_all, gt, V, do_step=False)
    g_all = state.grad.clone()
    total = torch.zeros_like(g_all)
    loss = 0.0
    for rank in range(2):
        views = refine.shard_views(V, rank, 2)
        s2, st = _state_and_settings(cuda_dev, P, V, W, H, views)
        r = refine.refine_iteration(s2, st, gt[views], V, do_step=False)
        total += s2.grad
        loss += float(r.loss)
    assert abs(loss - float(r_all.loss)) <= 1e-5 * abs(loss)
    assert float((total - g_all).abs().max()) <= 1e-4 * float(g_all.abs().max())


def test_native_step_matches_autograd_step(cuda_dev, native):
    """_C.refine_grad (no autograd graph, gradients written in place into the packed record) against the same
    iteration through torch autograd over rasterize_views: bit-identical when both use the activation kernels,
    and within the gradient parity bar when autograd uses torch's exp / normalize / sigmoid (their one-ulp
    differences in the activated parameters move the ill-conditioned position gradients by a few 1e-4)."""
    import tolerances
    from igs_b200 import refine
    from igs_b200.multiview import rasterize_views
    P, V, W, H = 20001, 3, 176, 120   # odd P: the record's slices are not 16-byte aligned
    state, st = _state_and_settings(cuda_dev, P, V, W, H)
    gen = torch.Generator().manual_seed(11)
    gt = torch.rand(V, 3, H, W, generator=gen).to(cuda_dev)
    # total_views = 8: the 1/total_views scaling of dL/dcolor is exact in every path
    ra = refine.refine_iteration(state, st, gt, 8, do_step=False, impl="autograd")
    g_auto = state.grad.clone()
    state.grad.fill_(float("nan"))   # every element must be overwritten
    rn = refine.refine_iteration(state, st, gt, 8, do_step=False, impl="native")
    g_nat = state.grad.clone()
    assert rn.num_rendered == ra.num_rendered and torch.equal(rn.radii, ra.radii)
    assert abs(float(rn.loss) - float(ra.loss)) <= 1e-6 * abs(float(ra.loss))
    assert bool(torch.isfinite(g_nat).all())
    for name, _ in refine.GROUPS:
        a, b, _k = state.slices[name]
        scale = float(g_auto[a:b].abs().max())
        assert float((g_nat[a:b] - g_auto[a:b]).abs().max()) <= tolerances.GRAD_RTOL * scale, name
    # same activated inputs -> same kernels -> identical bits
    sc, rot, op = native.activate_forward(state.param("scaling"), state.param("rotation"), state.param("opacity"))
    leaves = [t.detach().clone().requires_grad_(True) for t in (state.param("xyz"), state.param("shs"), op, sc, rot)]
    color = rasterize_views(*leaves, st)[0]
    loss = refine.refine_loss_sum(color, gt, 0.8, "fused") / 8.0
    g_xyz, g_sh, g_op, g_sc, g_rot = torch.autograd.grad(loss, leaves)
    d_ls, d_rr, d_lg = native.activate_backward(sc, state.param("rotation"), op, g_sc, g_rot, g_op)
    for name, ref in (("xyz", g_xyz), ("shs", g_sh), ("scaling", d_ls), ("rotation", d_rr), ("opacity", d_lg)):
        a, b, _k = state.slices[name]
        assert torch.equal(g_nat[a:b], ref.reshape(-1)), name
    assert torch.equal(rn.grad_means2D, color.grad_fn.grad_means2D)
    assert float(rn.loss) == float(loss)


def test_activation_kernels_match_torch(cuda_dev, native):
    import torch.nn.functional as F
    gen = torch.Generator().manual_seed(12)
    P = 4099
    ls = (torch.randn(P, 3, generator=gen) - 2).to(cuda_dev).requires_grad_(True)
    rr = torch.randn(P, 4, generator=gen).to(cuda_dev)
    rr[:5] *= 1e-3
    rr.requires_grad_(True)
    lg = (3 * torch.randn(P, 1, generator=gen)).to(cuda_dev).requires_grad_(True)
    sc, rot, op = native.activate_forward(ls.detach(), rr.detach(), lg.detach())
    tsc, trot, top = torch.exp(ls), F.normalize(rr), torch.sigmoid(lg)
    assert float((sc - tsc).abs().max()) <= 2e-7 * float(tsc.abs().max())
    assert float((rot - trot).abs().max()) <= 3e-7
    assert float((op - top).abs().max()) <= 2e-7
    gs, gr, go = (torch.randn(P, 3, generator=gen).to(cuda_dev), torch.randn(P, 4, generator=gen).to(cuda_dev),
                  torch.randn(P, 1, generator=gen).to(cuda_dev))
    ((tsc * gs).sum() + (trot * gr).sum() + (top * go).sum()).backward()
    d_ls, d_rr, d_lg = native.activate_backward(sc, rr.detach(), op, gs, gr, go)
    assert float((d_ls - ls.grad).abs().max()) <= 1e-6 * float(ls.grad.abs().max())
    assert float(((d_rr - rr.grad).abs() / (rr.grad.abs() + 1e-3 * rr.grad.abs().max())).max()) <= 1e-4
    assert float((d_lg - lg.grad).abs().max()) <= 1e-6 * float(lg.grad.abs().max())


def test_grouped_adam_matches_per_group_adam(cuda_dev, native):
    gen = torch.Generator().manual_seed(13)
    sizes = [3001, 4000, 48017, 1000, 3003]
    lrs = [0.0016, 0.01, 0.0025, 0.05, 0.005]
    n = sum(sizes)
    p0 = torch.randn(n, generator=gen).to(cuda_dev)
    a, ma, va = p0.clone(), torch.zeros(n, device=cuda_dev), torch.zeros(n, device=cuda_dev)
    b, mb, vb = p0.clone(), torch.zeros(n, device=cuda_dev), torch.zeros(n, device=cuda_dev)
    for step in range(1, 4):
        g = torch.randn(n, generator=gen).to(cuda_dev)
        native.adam_step_groups(a, g, ma, va, sizes, lrs, 0.9, 0.999, 1e-15, step)
        off = 0
        for sz, lr in zip(sizes, lrs):
            native.adam_step(b[off:off + sz], g[off:off + sz], mb[off:off + sz], vb[off:off + sz], lr, 0.9, 0.999, 1e-15, step)
            off += sz
        assert torch.equal(a, b) and torch.equal(ma, mb) and torch.equal(va, vb)


def test_tile_row_window_equals_full_render_inside(cuda_dev, native):
    """A view rendered only in a band of tile rows: identical pixels inside the band, background outside, fewer
    instances; radii report visibility inside the band."""
    from igs_b200.multiview import rasterize_views
    P, V, W, H = 6000, 2, 176, 200
    state, st = _state_and_settings(cuda_dev, P, V, W, H)
    sc, rot, op = native.activate_forward(state.param("scaling"), state.param("rotation"), state.param("opacity"))
    full = rasterize_views(state.param("xyz"), state.param("shs"), op, sc, rot, st)
    rows = torch.tensor([[3, 9], [0, 5]], dtype=torch.int32, device=cuda_dev)
    band = rasterize_views(state.param("xyz"), state.param("shs"), op, sc, rot, st._replace(tile_rows=rows))
    for v, (a, b) in enumerate(rows.tolist()):
        y0, y1 = 16 * a, min(H, 16 * b)
        for k in (0, 2, 3, 4, 5, 6, 7):   # colour, coord, mcoord, depth, mdepth, alpha, normal
            assert torch.equal(band[k][v][:, y0:y1], full[k][v][:, y0:y1]), (v, k)
        outside = torch.cat([band[0][v][:, :y0], band[0][v][:, y1:]], dim=1)
        assert float(outside.abs().max()) == 0.0 and float(band[6][v].sum() - band[6][v][:, y0:y1].sum()) == 0.0
        assert bool(((band[1][v] > 0) <= (full[1][v] > 0)).all()) and int((band[1][v] > 0).sum()) < int((full[1][v] > 0).sum())


@pytest.mark.parametrize("world", [2, 4])
def test_band_sharded_step_equals_single_rank_step(cuda_dev, native, world):
    """3 views on 2 / 4 ranks do not divide: shard_units cuts the views into bands of tile rows.  The sum over ranks
    of the band losses and packed gradients equals the one-rank step on all views."""
    from igs_b200 import refine
    P, V, W, H = 12000, 3, 176, 200
    state, st_all = _state_and_settings(cuda_dev, P, V, W, H)
    gen = torch.Generator().manual_seed(21)
    gt = torch.rand(V, 3, H, W, generator=gen).to(cuda_dev)
    r_all = refine.refine_iteration(state, st_all, gt, V, do_step=False)
    g_all = state.grad.clone()
    total = torch.zeros_like(g_all)
    loss, rendered, split = 0.0, 0, False
    for rank in range(world):
        units = refine.shard_units(V, H, rank, world)
        split |= any(u.row0 > 0 or u.row1 < H for u in units)
        st = refine.settings_for_units(st_all, units)
        r = refine.refine_iteration(state, st, gt[[u.view for u in units]], V, do_step=False)
        total += state.grad
        loss += float(r.loss)
        rendered += r.num_rendered
    assert split
    assert abs(loss - float(r_all.loss)) <= 1e-5 * abs(loss)
    for name, _ in refine.GROUPS:
        a, b, _k = state.slices[name]
        assert float((total[a:b] - g_all[a:b]).abs().max()) <= 2e-4 * float(g_all[a:b].abs().max()), name
    assert rendered < 1.35 * r_all.num_rendered   # the halo rows are the only duplicated work


def test_step_without_geometry_maps_is_bit_identical(cuda_dev, native):
    """require_coord = require_depth = False inside the fused step (the maps the refine loss never reads are not
    rendered): loss, packed gradients, radii, instance count and screen-space gradients are the same bits."""
    from igs_b200 import refine
    P, V, W, H = 15000, 3, 176, 120
    state, st = _state_and_settings(cuda_dev, P, V, W, H)
    gt = torch.rand(V, 3, H, W, generator=torch.Generator().manual_seed(31)).to(cuda_dev)
    a = refine.refine_iteration(state, st, gt, V, do_step=False)
    g_a = state.grad.clone()
    state.grad.fill_(float("nan"))
    b = refine.refine_iteration(state, st._replace(require_coord=False, require_depth=False), gt, V, do_step=False)
    assert torch.equal(state.grad, g_a) and float(a.loss) == float(b.loss)
    assert a.num_rendered == b.num_rendered and torch.equal(a.radii, b.radii) and torch.equal(a.grad_means2D, b.grad_means2D)
