"""Key-frame guided streaming refine: the per-key-frame loop of IGS on the B200-native rasterizer.

Reference: infer_batch.py:244-324.  For every key frame the reference converts the streamed Gaussians into a
GaussianModel with a FRESH Adam optimizer (`load_fromstream`, igs/models/gaussian_model.py:265-348), then runs
`refine_iterations` single-view iterations: views are drawn without replacement from the key frame's camera list
(`randint` + `pop`, refilled when empty, :283-292), each iteration renders the view, takes
lambda*L1 + (1-lambda)*(1-SSIM), calls backward, updates the densification statistics / densifies
(:308-321) and steps Adam (:323).  After the loop the key frame's output view is rendered once more for the PSNR
(:352-362).

Here one iteration is ONE native call (`refine.refine_iteration`), the statistics update is one kernel over the
step's views, densify/prune is the single-gather rebuild of `igs_b200.densify`, and the key frame's target images are
uploaded once (pinned host -> device on a copy stream) instead of one `.to("cuda")` per iteration.
`views_per_step=1` is the reference's schedule; `views_per_step=V` is BASELINE's batched schedule.
"""
from __future__ import annotations

import random
from typing import List, NamedTuple, Optional

import torch

from . import refine
from .densify import DensifyStats, densify_and_prune
from .multiview import MultiViewSettings


class StreamConfig(NamedTuple):
    """`opt` fields of the reference's yaml that steer the refine loop (configs/demo.yaml:52-72)."""
    refine_iterations: int = 50
    lambda_l1: float = 0.8
    use_densify: bool = False
    densify_from_iter: int = 0
    densify_until_iter: int = 100
    densification_interval: int = 20
    densify_grad_threshold: float = 0.00015
    max_num: int = 150000
    control_max: bool = True
    min_opacity: float = 0.005          # literal at infer_batch.py:320
    radius: float = 1.0                 # batch["radius"][0]: scene extent used by densify_and_prune
    percent_dense: float = 0.0          # GaussianModel.__init__ (gaussian_model.py:84); load_fromstream never changes it
    views_per_step: int = 1
    lrs: Optional[dict] = None          # training_lr; None -> refine.DEFAULT_LRS


def view_schedule(n_views: int, iterations: int, rng: random.Random, views_per_step: int = 1) -> List[List[int]]:
    """The reference's draw: `pick = randint(0, len(list)-1); list.pop(pick)`, the list refilled with all views when
    it runs empty (infer_batch.py:283-292).  Returns one list of view indices per iteration."""
    out, pool = [], []
    for _ in range(iterations):
        step = []
        for _k in range(views_per_step):
            if not pool:
                pool = list(range(n_views))
            step.append(pool.pop(rng.randint(0, len(pool) - 1)))
        out.append(step)
    return out


def select_views(all_views: MultiViewSettings, idx: List[int]) -> MultiViewSettings:
    """Settings of a step that renders the views `idx` of `all_views` (a contiguous slice is a view, no copy)."""
    if len(idx) == 1 or idx == list(range(idx[0], idx[0] + len(idx))):
        a, b = idx[0], idx[0] + len(idx)
        return all_views._replace(viewmatrices=all_views.viewmatrices[a:b], projmatrices=all_views.projmatrices[a:b],
                                  campos=all_views.campos[a:b])
    t = torch.tensor(idx, dtype=torch.long, device=all_views.viewmatrices.device)
    return all_views._replace(viewmatrices=all_views.viewmatrices[t].contiguous(),
                              projmatrices=all_views.projmatrices[t].contiguous(), campos=all_views.campos[t].contiguous())


class KeyframeResult(NamedTuple):
    losses: torch.Tensor        # [iterations] device tensor (no host sync inside the loop)
    schedule: List[List[int]]
    P: int                      # Gaussians after the loop (densify / prune change it)
    densified_at: List[int]


def refine_keyframe(state: refine.RefineState, views: MultiViewSettings, gt: torch.Tensor, cfg: StreamConfig,
                    rng: random.Random, group=None, generator: Optional[torch.Generator] = None) -> KeyframeResult:
    """One key frame of infer_batch.py:266-324 on `state` (fresh Adam moments, like `load_fromstream`).

    views: all V cameras of the key frame; gt [V,3,H,W] on the device.  With a process group the step's views are the
    same on every rank and `views_per_step` must be a multiple of the world size: rank r renders every world-th view
    of the step and gradients / densification statistics are all-reduced."""
    V = int(views.viewmatrices.shape[0])
    world = torch.distributed.get_world_size(group) if group is not None else 1
    rank = torch.distributed.get_rank(group) if group is not None else 0
    if cfg.views_per_step % world != 0:
        raise ValueError("views_per_step must be a multiple of the number of ranks")
    state.exp_avg.zero_()
    state.exp_avg_sq.zero_()
    state.step = 0
    if cfg.lrs is not None:
        state.lrs = dict(cfg.lrs)
    stats = DensifyStats(state.P, state.device) if cfg.use_densify else None
    sched = view_schedule(V, cfg.refine_iterations, rng, cfg.views_per_step)
    losses = torch.zeros(cfg.refine_iterations, device=state.device)
    densified_at = []
    for it, step_views in enumerate(sched):
        mine = step_views[rank::world]
        st = select_views(views, mine)
        gt_step = gt[mine[0]:mine[0] + 1] if len(mine) == 1 else gt[torch.tensor(mine, device=gt.device)]
        r = refine.refine_iteration(state, st, gt_step, len(step_views), group=group, lambda_l1=cfg.lambda_l1,
                                    do_step=False)
        losses[it] = r.loss
        densified = False
        if cfg.use_densify and it < cfg.densify_until_iter:
            if st.tile_rows is not None:
                raise ValueError("densification statistics need whole views (tile-row bands would be counted twice)")
            stats.update(r.grad_means2D, r.radii, grad_scale=float(len(step_views)))   # back on the per-view scale
            if it > cfg.densify_from_iter and it % cfg.densification_interval == 0:
                stats.all_reduce(group)
                densify_and_prune(state, stats, cfg.densify_grad_threshold, cfg.min_opacity, cfg.radius, None,
                                  cfg.max_num, cfg.percent_dense, cfg.control_max, generator=generator)
                densified = True
                densified_at.append(it)
        # The reference's densify_and_prune replaces every nn.Parameter, so the `optimizer.step()` that follows it in
        # the same iteration sees grad=None and updates nothing (infer_batch.py:320-324): no Adam step here either.
        if not densified:
            state.adam_step()
    return KeyframeResult(losses, sched, state.P, densified_at)


def render_view(state: refine.RefineState, views: MultiViewSettings, v: int) -> torch.Tensor:
    """Forward-only render of view `v` (the key frame's output view, infer_batch.py:352-356): colour [3,H,W]."""
    with torch.no_grad():
        return refine.forward_views(state, select_views(views, [v]))[0][0]


def psnr(pred: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """-10 log10 mean((clamp(pred,0,1) - gt)^2)  (infer_batch.py:358-361)."""
    return -10.0 * torch.log10(torch.mean((pred.clamp(0, 1) - gt) ** 2))


class StreamResult(NamedTuple):
    psnrs: List[float]
    points_num: List[int]
    keyframe_ms: List[float]    # device time of each key frame's refine (upload + iterations + final render)
    iterations: int


def run_stream(state: refine.RefineState, views: MultiViewSettings, keyframe_images, cfg: StreamConfig, seed: int = 0,
               group=None, eval_view: int = 0) -> StreamResult:
    """Streams over key frames: `keyframe_images` yields one pinned-host [V,3,H,W] tensor per key frame (the
    refine_dataset entry of infer_batch.py:253-256).  The next key frame's images are uploaded on a copy stream while
    the current one refines.  Gaussians are carried from key frame to key frame (in the reference AGM-Net moves them
    in between; that network is outside this path)."""
    dev = state.device
    rng = random.Random(seed)
    gen = torch.Generator(device=dev).manual_seed(seed)
    copy_stream = torch.cuda.Stream(device=dev)
    frames = list(keyframe_images)
    bufs = [torch.empty_like(frames[0], device=dev) for _ in range(2)] if frames else []
    ready = [torch.cuda.Event() for _ in range(2)]
    free = [torch.cuda.Event() for _ in range(2)]
    cur = torch.cuda.current_stream(dev)

    def upload(k):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(free[k % 2])
            bufs[k % 2].copy_(frames[k], non_blocking=True)
            ready[k % 2].record(copy_stream)

    for e in free:
        e.record(cur)
    if frames:
        upload(0)
    psnrs, points, times = [], [], []
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in frames]
    vals = []
    for k in range(len(frames)):
        ev[k][0].record(cur)
        if k + 1 < len(frames):
            upload(k + 1)
        cur.wait_event(ready[k % 2])
        gt = bufs[k % 2]
        refine_keyframe(state, views, gt, cfg, rng, group=group, generator=gen)
        img = render_view(state, views, eval_view)
        vals.append(psnr(img, gt[eval_view]))
        free[k % 2].record(cur)
        ev[k][1].record(cur)
        points.append(state.P)
    torch.cuda.synchronize(dev)
    for k in range(len(frames)):
        times.append(ev[k][0].elapsed_time(ev[k][1]))
        psnrs.append(float(vals[k]))
    return StreamResult(psnrs, points, times, cfg.refine_iterations * len(frames))
