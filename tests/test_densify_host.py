"""CPU: the densify / prune planner (igs_b200/densify.py) against a line-by-line restatement of the reference's
GaussianModel.densify_and_prune (igs/models/gaussian_model.py:466-666) that rebuilds separate parameter tensors and
Adam moments with torch.cat / boolean indexing like the reference does."""
import math
from pathlib import Path

import numpy as np

import pytest
import torch

from igs_b200 import densify, refine


def build_rotation(r):  # igs/utils/general_utils.py:79-100, with a device argument
    norm = torch.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = torch.zeros((q.size(0), 3, 3), device=r.device)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - r * z); R[:, 0, 2] = 2 * (x * z + r * y)
    R[:, 1, 0] = 2 * (x * y + r * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - r * x)
    R[:, 2, 0] = 2 * (x * z - r * y); R[:, 2, 1] = 2 * (y * z + r * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


class RefModel:
    """The pieces of the reference GaussianModel that densify_and_prune touches (names as in the reference)."""

    def __init__(self, xyz, scaling, rotation, opacity, shs, m, v, percent_dense, max_num):
        self.t = dict(xyz=xyz.clone(), scaling=scaling.clone(), rotation=rotation.clone(), opacity=opacity.clone(), shs=shs.clone())
        self.m = {k: m[k].clone() for k in self.t}
        self.v = {k: v[k].clone() for k in self.t}
        self.percent_dense, self.max_num = percent_dense, max_num

    get_scaling = property(lambda self: torch.exp(self.t["scaling"]))
    get_opacity = property(lambda self: torch.sigmoid(self.t["opacity"]))

    def _prune(self, mask):  # prune_points / _prune_optimizer (:466-523)
        keep = ~mask
        for k in self.t:
            self.t[k] = self.t[k][keep]; self.m[k] = self.m[k][keep]; self.v[k] = self.v[k][keep]

    def _cat(self, d):  # cat_tensors_to_optimizer (:525-556)
        for k in self.t:
            self.t[k] = torch.cat((self.t[k], d[k]), dim=0)
            self.m[k] = torch.cat((self.m[k], torch.zeros_like(d[k])), dim=0)
            self.v[k] = torch.cat((self.v[k], torch.zeros_like(d[k])), dim=0)

    def densify_and_split(self, grads, grad_threshold, scene_extent, N=2):  # :586-615
        n_init_points = self.t["xyz"].shape[0]
        padded_grad = torch.zeros((n_init_points))
        padded_grad[:grads.shape[0]] = grads.squeeze()
        selected = torch.where(padded_grad >= grad_threshold, True, False)
        selected = torch.logical_and(selected, torch.max(self.get_scaling, dim=1).values > self.percent_dense * scene_extent)
        stds = self.get_scaling[selected].repeat(N, 1)
        means = torch.zeros((stds.size(0), 3))
        samples = torch.normal(mean=means, std=stds)
        rots = build_rotation(self.t["rotation"][selected]).repeat(N, 1, 1)
        new_xyz = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + self.t["xyz"][selected].repeat(N, 1)
        new_scaling = torch.log(self.get_scaling[selected].repeat(N, 1) / (0.8 * N))
        d = dict(xyz=new_xyz, scaling=new_scaling, rotation=self.t["rotation"][selected].repeat(N, 1),
                 opacity=self.t["opacity"][selected].repeat(N, 1), shs=self.t["shs"][selected].repeat(N, 1, 1))
        self._cat(d)
        self._prune(torch.cat((selected, torch.zeros(N * int(selected.sum()), dtype=bool))))

    def densify_and_clone(self, grads, grad_threshold, scene_extent):  # :619-638
        selected = torch.where(torch.norm(grads, dim=-1) >= grad_threshold, True, False)
        selected = torch.logical_and(selected, torch.max(self.get_scaling, dim=1).values <= self.percent_dense * scene_extent)
        self._cat({k: self.t[k][selected] for k in self.t})

    def densify_and_prune(self, accum, denom, max_grad, min_opacity, extent, max_screen_size, control_max=True):  # :640-666
        grads = accum / denom
        grads[grads.isnan()] = 0.0
        max_num_add = self.max_num - self.t["xyz"].shape[0]
        selected = torch.where(torch.norm(grads, dim=-1) >= max_grad, True, False)
        if control_max and selected.sum() > max_num_add:
            topk_values, topk_indices = torch.topk(grads, max_num_add, dim=0)
            grads = torch.zeros_like(grads)
            grads.scatter_(0, topk_indices, topk_values)
        self.densify_and_clone(grads, max_grad, extent)
        self.densify_and_split(grads, max_grad, extent)
        prune_mask = (self.get_opacity < min_opacity).squeeze()
        if max_screen_size:
            big_points_vs = torch.zeros(self.t["xyz"].shape[0]) > max_screen_size   # max_radii2D was reset by the postfix
            big_points_ws = self.get_scaling.max(dim=1).values > 0.1 * extent
            prune_mask = torch.logical_or(torch.logical_or(prune_mask, big_points_vs), big_points_ws)
        self._prune(prune_mask)


def _scene(P, seed):
    g = torch.Generator().manual_seed(seed)
    xyz = torch.randn(P, 3, generator=g)
    scaling = torch.randn(P, 3, generator=g) * 0.8 + math.log(0.05)
    rotation = torch.randn(P, 4, generator=g)
    opacity = torch.randn(P, 1, generator=g) * 3
    shs = torch.randn(P, 16, 3, generator=g)
    accum = torch.rand(P, 1, generator=g) * 4e-4 * (torch.rand(P, 1, generator=g) > 0.3)
    denom = torch.randint(0, 4, (P, 1), generator=g).float()     # zeros -> NaN grads -> 0
    m = dict(xyz=torch.randn(P, 3, generator=g), scaling=torch.randn(P, 3, generator=g), rotation=torch.randn(P, 4, generator=g),
             opacity=torch.randn(P, 1, generator=g), shs=torch.randn(P, 16, 3, generator=g))
    v = {k: torch.rand_like(t) for k, t in m.items()}
    return xyz, scaling, rotation, opacity, shs, accum, denom, m, v


@pytest.mark.parametrize("P,max_num,max_screen", [(4000, 150000, None), (4000, 4300, None), (3000, 150000, 20.0), (500, 500, None)])
def test_plan_and_gather_equal_the_reference_rebuild(P, max_num, max_screen):
    xyz, scaling, rotation, opacity, shs, accum, denom, m, v = _scene(P, 7 + P)
    extent, pd, max_grad, min_op = 3.0, 0.01, 0.00015, 0.005
    ref = RefModel(xyz, scaling, rotation, opacity, shs, m, v, pd, max_num)
    torch.manual_seed(123)
    ref.densify_and_prune(accum.clone(), denom.clone(), max_grad, min_op, extent, max_screen)

    state = refine.RefineState(xyz, scaling, rotation, opacity, shs, "cpu")
    for name in ("xyz", "scaling", "rotation", "opacity", "shs"):
        state.view(state.exp_avg, name).copy_(m[name].reshape(state.view(state.exp_avg, name).shape))
        state.view(state.exp_avg_sq, name).copy_(v[name].reshape(state.view(state.exp_avg_sq, name).shape))
    torch.manual_seed(123)
    plan = densify.plan_densify(None, state.param("scaling"), state.param("opacity"), accum.reshape(-1).clone(),
                                denom.reshape(-1).clone(), max_grad, min_op, extent, max_screen, max_num, pd)
    P_new = int(plan.src.numel())
    assert P_new == ref.t["xyz"].shape[0]
    assert (plan.n_clone > 0 and plan.n_split > 0 and plan.n_pruned > 0) or max_num <= P
    flat, ea, eas = densify.apply_plan_torch(state.flat, state.exp_avg, state.exp_avg_sq, P, plan)
    state.resize(P_new, flat, ea, eas)
    for name in ("rotation", "opacity", "shs"):
        assert torch.equal(state.param(name).reshape(ref.t[name].shape), ref.t[name]), name
    for name in ("xyz", "scaling"):
        assert torch.allclose(state.param(name), ref.t[name], rtol=1e-6, atol=1e-6), name
    for name in ("xyz", "scaling", "rotation", "opacity", "shs"):
        assert torch.equal(state.view(state.exp_avg, name).reshape(ref.m[name].shape), ref.m[name]), name
        assert torch.equal(state.view(state.exp_avg_sq, name).reshape(ref.v[name].shape), ref.v[name]), name
    if max_num >= P:
        assert P_new <= max(max_num, P) + 2 * plan.n_split   # bounded densify never adds more than asked


def test_stats_update_matches_reference_semantics():
    g = torch.Generator().manual_seed(3)
    P, V = 257, 5
    g2d = torch.randn(V, P, 3, generator=g)
    radii = torch.randint(-1, 6, (V, P), generator=g, dtype=torch.int32).clamp_min(0)
    st = densify.DensifyStats(P, "cpu")
    st.update_torch(g2d, radii)
    exp_acc = torch.zeros(P); exp_den = torch.zeros(P); exp_r = torch.zeros(P)
    for vv in range(V):                                  # infer_batch.py:312-313 + gaussian_model.py:865-868, per view
        vis = radii[vv] > 0
        exp_r[vis] = torch.max(exp_r[vis], radii[vv][vis].float())
        exp_acc[vis] += torch.norm(g2d[vv][vis, :2], dim=-1)
        exp_den[vis] += 1
    assert torch.equal(st.grad_accum, exp_acc) and torch.equal(st.denom, exp_den) and torch.equal(st.max_radii2D, exp_r)


# ------------------------------------------------------------------------------------------------------------------
# Pin to the reference class itself: goldens written by tools/make_golden_densify.py, which runs the UNMODIFIED
# /root/reference/igs/models/gaussian_model.py (fresh GaussianModel -> load_fromstream -> Adam steps -> statistics ->
# densify_and_prune, the call path of infer_batch.py:266-324) in the build container.
# ------------------------------------------------------------------------------------------------------------------
GOLDEN = Path(__file__).resolve().parent / "golden"
NAMES = ("xyz", "rotation", "shs", "opacity", "scaling")


def _state_from_golden(g, device="cpu"):
    t = {k: torch.from_numpy(g[f"pre_{k}"]) for k in NAMES}
    st = refine.RefineState(t["xyz"], t["scaling"], t["rotation"], t["opacity"], t["shs"], device)
    for k in NAMES:
        st.view(st.exp_avg, k).copy_(torch.from_numpy(g[f"pre_m_{k}"]).reshape(st.view(st.exp_avg, k).shape))
        st.view(st.exp_avg_sq, k).copy_(torch.from_numpy(g[f"pre_v_{k}"]).reshape(st.view(st.exp_avg_sq, k).shape))
    return st


def _check_against_golden(st, g, exact_moments=True):
    P_new = g["post_xyz"].shape[0]
    assert st.P == P_new
    for k in NAMES:
        got = st.param(k).detach().cpu().reshape(g[f"post_{k}"].shape)
        ref = torch.from_numpy(g[f"post_{k}"])
        if k in ("xyz", "scaling"):       # split children: R(q) * sample + xyz and log(exp(s) / 1.6), fp32 re-association
            assert torch.allclose(got, ref, rtol=2e-6, atol=2e-6), k
        else:
            assert torch.equal(got, ref), k
        for buf, tag in ((st.exp_avg, "m"), (st.exp_avg_sq, "v")):
            got = st.view(buf, k).detach().cpu().reshape(g[f"post_{tag}_{k}"].shape)
            assert torch.equal(got, torch.from_numpy(g[f"post_{tag}_{k}"])), (tag, k)


@pytest.mark.parametrize("case", ["refine", "bounded", "cloned"])
def test_statistics_and_densify_match_the_reference_class(case):
    g = np.load(GOLDEN / f"ref_densify_{case}.npz")
    P = int(g["cfg_P"])
    # statistics: three single-view updates (add_densification_stats + the max_radii2D line of infer_batch.py:312)
    stats = densify.DensifyStats(P, "cpu")
    for i in range(3):
        radii = torch.where(torch.from_numpy(g[f"stat{i}_vis"]), torch.from_numpy(g[f"stat{i}_radii"]), torch.zeros(P)).to(torch.int32)
        stats.update_torch(torch.from_numpy(g[f"stat{i}_grad"]).unsqueeze(0), radii.unsqueeze(0))
    assert torch.equal(stats.grad_accum, torch.from_numpy(g["grad_accum"]).reshape(-1))
    assert torch.equal(stats.denom, torch.from_numpy(g["denom"]).reshape(-1))
    assert torch.equal(stats.max_radii2D, torch.from_numpy(g["max_radii2D"]).reshape(-1))
    # densify_and_prune with the reference's defaults for this path (percent_dense is whatever the golden's model had)
    st = _state_from_golden(g)
    gen = torch.Generator().manual_seed(int(g["cfg_rng_seed"]))   # the reference draws from the global CPU generator
    plan = densify.plan_densify(None, st.param("scaling"), st.param("opacity"), stats.grad_accum, stats.denom,
                                float(g["cfg_thr"]), 0.005, float(g["cfg_extent"]), None, int(g["cfg_max_num"]),
                                float(g["cfg_percent_dense"]), True, generator=gen)
    if case == "refine":
        assert plan.n_clone == 0 and plan.n_split > 0      # percent_dense = 0: the refine loop never clones
    if case == "cloned":
        assert plan.n_clone > 0 and plan.n_split > 0
    flat, m, v = densify.apply_plan_torch(st.flat, st.exp_avg, st.exp_avg_sq, P, plan)
    st.resize(int(plan.src.numel()), flat, m, v)
    _check_against_golden(st, g)


def test_default_percent_dense_is_the_refine_paths_zero():
    import inspect
    from igs_b200 import stream
    assert stream.StreamConfig().percent_dense == 0.0
    assert inspect.signature(densify.densify_and_prune).parameters["percent_dense"].default == 0.0
