"""Densify / prune of the refine loop on the packed parameter record.

Reference: igs/models/gaussian_model.py:586-666 (`densify_and_split`, `densify_and_clone`, `densify_and_prune`),
:466-557 (`_prune_optimizer`, `prune_points`, `cat_tensors_to_optimizer`), :558-584 (`densification_postfix`),
:865-868 (`add_densification_stats`); call site infer_batch.py:308-321.

The reference rebuilds every nn.Parameter and both Adam moments three times per call (clone -> cat, split -> cat +
boolean prune, opacity prune) and finishes with torch.cuda.empty_cache().  Here the three stages are planned on
[P]-sized vectors only (`plan_densify`: which old Gaussian every new slot copies, and whether it is kept, cloned or a
split child), and ONE kernel (`igs_densify_gather`) writes the new packed parameters and moments in a single pass.
The order of the surviving Gaussians, the Adam-state carry (kept: carried; new: zero), the split sampling
(torch.normal with the caller's generator, so the random stream is the reference's) and the max-point-bounded top-k
are the reference's.  Statistics follow the original 3DGS semantics after a densification: zeros of the new size (the
reference's `prune_points` leaves them at the stale pre-prune length).
"""
from __future__ import annotations

from typing import NamedTuple, Optional

import torch

from .refine import GROUPS, RefineState

_ROLE_ORDER = ("xyz", "rotation", "shs", "opacity", "scaling")   # the kernel's component order


class DensifyStats:
    """xyz_gradient_accum / denom / max_radii2D of GaussianModel, as flat [P] device vectors."""

    def __init__(self, P: int, device):
        self.grad_accum = torch.zeros(P, device=device)
        self.denom = torch.zeros(P, device=device)
        self.max_radii2D = torch.zeros(P, device=device)

    def update(self, grad_means2D: torch.Tensor, radii: torch.Tensor, grad_scale: float = 1.0) -> None:
        """All V views of one step at once (the reference calls add_densification_stats once per single-view
        iteration, infer_batch.py:312-313): grad_means2D [V,P,3], radii [V,P] int32.

        grad_scale: the reference accumulates the norm of the gradient of ONE view's undivided loss
        (infer_batch.py:302-313); a step whose loss is the mean over `total_views` views hands in gradients that are
        1/total_views of that, so the caller passes total_views and `densify_grad_threshold` keeps its meaning.  The
        views must be whole images: a view rendered as several tile-row bands would be counted once per band."""
        from . import _C
        _C.densify_stats(grad_means2D, radii, self.grad_accum, self.denom, self.max_radii2D, float(grad_scale))

    def all_reduce(self, group=None) -> None:
        """View-sharded refine: every rank saw only its views.  SUM the accumulators and MAX the radii over the
        ranks, so that every rank plans the identical densification (same statistics, same parameters, same seeded
        generator) and no broadcast of the new parameters is needed."""
        if group is None or torch.distributed.get_world_size(group) <= 1:
            return
        torch.distributed.all_reduce(self.grad_accum, group=group)
        torch.distributed.all_reduce(self.denom, group=group)
        torch.distributed.all_reduce(self.max_radii2D, op=torch.distributed.ReduceOp.MAX, group=group)

    def update_torch(self, grad_means2D: torch.Tensor, radii: torch.Tensor, grad_scale: float = 1.0) -> None:
        """The same update with torch ops (restatement used by the tests)."""
        for v in range(radii.shape[0]):
            vis = radii[v] > 0
            self.grad_accum[vis] += grad_scale * torch.norm(grad_means2D[v][vis, :2], dim=-1)
            self.denom[vis] += 1
            self.max_radii2D[vis] = torch.max(self.max_radii2D[vis], radii[v][vis].to(self.max_radii2D.dtype))


class DensifyPlan(NamedTuple):
    src: torch.Tensor       # [P_new] int32: old Gaussian each new slot copies
    kind: torch.Tensor      # [P_new] uint8: 0 keep, 1 clone, 2 split child
    samples: torch.Tensor   # [P_new,3] float32: N(0, scale) draws of the split children (zeros elsewhere)
    split_div: float        # 0.8 * N
    n_clone: int
    n_split: int            # parents split (each replaced by N children)
    n_pruned: int


def plan_densify(xyz_unused, log_scale: torch.Tensor, opacity_logit: torch.Tensor, grad_accum: torch.Tensor,
                 denom: torch.Tensor, max_grad: float, min_opacity: float, extent: float,
                 max_screen_size: Optional[float], max_num: int, percent_dense: float, control_max: bool = True,
                 N: int = 2, generator: Optional[torch.Generator] = None) -> DensifyPlan:
    """densify_and_prune (gaussian_model.py:640-666) on index vectors.  Pure torch: runs on any device."""
    P = log_scale.shape[0]
    dev = log_scale.device
    grads = grad_accum.reshape(P, 1) / denom.reshape(P, 1)
    grads[grads.isnan()] = 0.0
    max_num_add = max_num - P
    selected = torch.norm(grads, dim=-1) >= max_grad
    if control_max and int(selected.sum()) > max_num_add:          # max-points bounded densify (:648-653)
        k = max(max_num_add, 0)
        topk_values, topk_indices = torch.topk(grads, k, dim=0)
        grads = torch.zeros_like(grads)
        grads.scatter_(0, topk_indices, topk_values)
    scale_max = torch.exp(log_scale).max(dim=1).values               # get_scaling.max(dim=1)
    thr = percent_dense * extent
    # --- densify_and_clone (:619-638) ---
    clone_mask = (torch.norm(grads, dim=-1) >= max_grad) & (scale_max <= thr)
    clone_idx = torch.nonzero(clone_mask).reshape(-1)
    n_clone = int(clone_idx.numel())
    seq_src = torch.cat([torch.arange(P, device=dev), clone_idx])
    seq_kind = torch.cat([torch.zeros(P, dtype=torch.uint8, device=dev), torch.ones(n_clone, dtype=torch.uint8, device=dev)])
    # --- densify_and_split (:586-615) on the P + n_clone points ---
    padded_grad = torch.zeros(P + n_clone, device=dev)
    padded_grad[:P] = grads.squeeze(-1)
    split_sel = (padded_grad >= max_grad) & (scale_max[seq_src] > thr)
    split_idx = torch.nonzero(split_sel).reshape(-1)
    n_split = int(split_idx.numel())
    parents = seq_src[split_idx]
    stds = torch.exp(log_scale[parents]).repeat(N, 1)
    means = torch.zeros((stds.size(0), 3), device=dev)
    samples = torch.normal(mean=means, std=stds, generator=generator) if stds.numel() else means
    seq_src = torch.cat([seq_src, parents.repeat(N)])
    seq_kind = torch.cat([seq_kind, torch.full((N * n_split,), 2, dtype=torch.uint8, device=dev)])
    seq_samples = torch.cat([torch.zeros(P + n_clone, 3, device=dev), samples])
    keep = ~torch.cat([split_sel, torch.zeros(N * n_split, dtype=torch.bool, device=dev)])   # prune the split parents
    seq_src, seq_kind, seq_samples = seq_src[keep], seq_kind[keep], seq_samples[keep]
    # --- final prune (:658-664) ---
    prune = torch.sigmoid(opacity_logit.reshape(-1)[seq_src]) < min_opacity
    if max_screen_size:
        # max_radii2D was reset to zeros by densification_postfix, so `big_points_vs` is (0 > max_screen_size)
        big_vs = torch.zeros_like(prune) if max_screen_size >= 0 else torch.ones_like(prune)
        ent_scale = torch.exp(log_scale[seq_src])
        child = (seq_kind == 2).unsqueeze(-1)
        ent_scale = torch.where(child, torch.exp(torch.log(ent_scale / (0.8 * N))), ent_scale)
        big_ws = ent_scale.max(dim=1).values > 0.1 * extent
        prune = prune | big_vs | big_ws
    n_pruned = int(prune.sum())
    live = ~prune
    return DensifyPlan(seq_src[live].to(torch.int32).contiguous(), seq_kind[live].contiguous(),
                       seq_samples[live].to(torch.float32).contiguous(), 0.8 * N, n_clone, n_split, n_pruned)


def _offsets(P: int):
    off, out = 0, {}
    for name, k in GROUPS:
        out[name] = off
        off += P * k
    return [out[n] for n in _ROLE_ORDER]


def apply_plan_torch(flat: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor, P_old: int, plan: DensifyPlan):
    """Torch restatement of igs_densify_gather (any device): returns (flat, exp_avg, exp_avg_sq) of the new size."""
    widths = dict(GROUPS)
    P_new = int(plan.src.numel())
    src = plan.src.long()
    new = [torch.empty(P_new * 59, dtype=flat.dtype, device=flat.device) for _ in range(3)]
    oo, on = dict(zip(_ROLE_ORDER, _offsets(P_old))), dict(zip(_ROLE_ORDER, _offsets(P_new)))
    fresh = (plan.kind != 0)
    child = (plan.kind == 2)
    old = {}
    for name in _ROLE_ORDER:
        w = widths[name]
        old[name] = [b[oo[name]:oo[name] + P_old * w].view(P_old, w) for b in (flat, exp_avg, exp_avg_sq)]
    for name in _ROLE_ORDER:
        w = widths[name]
        val = old[name][0][src].clone()
        if name == "xyz" and bool(child.any()):
            q = old["rotation"][0][src][child]
            q = q / torch.sqrt((q * q).sum(dim=1, keepdim=True))
            r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
            R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                             2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                             2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).view(-1, 3, 3)
            val[child] = torch.bmm(R, plan.samples[child].unsqueeze(-1)).squeeze(-1) + val[child]
        if name == "scaling" and bool(child.any()):
            val[child] = torch.log(torch.exp(val[child]) / plan.split_div)
        new[0][on[name]:on[name] + P_new * w] = val.reshape(-1)
        for b in (1, 2):
            mv = old[name][b][src].clone()
            mv[fresh] = 0.0
            new[b][on[name]:on[name] + P_new * w] = mv.reshape(-1)
    return tuple(new)


def densify_and_prune(state: RefineState, stats: DensifyStats, max_grad: float, min_opacity: float, extent: float,
                      max_screen_size: Optional[float] = None, max_num: int = 150000, percent_dense: float = 0.0,
                      control_max: bool = True, N: int = 2, generator: Optional[torch.Generator] = None) -> DensifyPlan:
    """GaussianModel.densify_and_prune on a RefineState: plans on [P] vectors, then ONE gather kernel rebuilds the
    packed parameters and both Adam moments.  `state` and `stats` are updated in place; returns the plan.

    percent_dense defaults to 0: that is its value on the refine path — GaussianModel.__init__ sets it to 0
    (gaussian_model.py:84) and only training_setup(), which load_fromstream / infer_batch.py never call, changes it —
    so the reference never clones there and splits every selected Gaussian."""
    from . import _C
    P_old = state.P
    plan = plan_densify(None, state.param("scaling"), state.param("opacity"), stats.grad_accum, stats.denom, max_grad,
                        min_opacity, extent, max_screen_size, max_num, percent_dense, control_max, N, generator)
    P_new = int(plan.src.numel())
    flat, m, v = _C.densify_gather(state.flat, state.exp_avg, state.exp_avg_sq, P_old, plan.src, plan.kind, plan.samples,
                                   float(plan.split_div), _offsets(P_old), _offsets(P_new))
    state.resize(P_new, flat, m, v)
    stats.__init__(P_new, state.device)
    return plan
