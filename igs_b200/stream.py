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
    # "exact": every forward reads its instance count back (the reference's behaviour, rasterizer_impl.cu:354);
    # "async": capacity mode after the key frame's first iteration, no host round-trip (refine.RefineState);
    # "graph": as "async", and the whole iteration is ONE CUDA-graph replay (single rank, densify off)
    mode: str = "exact"
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
    redone: int = 0             # capacity-mode iterations whose instance count overflowed and were run again


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
    if cfg.mode not in ("exact", "async", "graph"):
        raise ValueError(f"unknown stream mode {cfg.mode!r}")
    if cfg.mode == "graph" and (world > 1 or cfg.use_densify):
        raise ValueError('mode="graph" is single-rank and needs densify off (the Gaussian count must not change)')
    state.enable_async_binning(cfg.mode == "async")
    if cfg.mode == "graph":
        return _refine_keyframe_graphed(state, views, gt, cfg, sched, losses)
    overflows0 = state.overflows
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
            state.adam_step(guarded=state.last_guarded)
    redone = 0
    if state.async_binning:
        # an iteration whose instance count exceeded its capacity skipped its own Adam step on the device; the count
        # is known by now (one synchronisation per key frame), so the same number of iterations is run again
        state.poll_status(wait=True)
        redone = state.overflows - overflows0
        for k in range(redone):
            mine = sched[-1 - k % len(sched)][rank::world]
            st = select_views(views, mine)
            gt_step = gt[mine[0]:mine[0] + 1] if len(mine) == 1 else gt[torch.tensor(mine, device=gt.device)]
            state.enable_async_binning(False)
            refine.refine_iteration(state, st, gt_step, len(sched[-1]), group=group, lambda_l1=cfg.lambda_l1)
            state.enable_async_binning(True)
    return KeyframeResult(losses, sched, state.P, densified_at, redone)


def _refine_keyframe_graphed(state, views, gt, cfg, sched, losses) -> KeyframeResult:
    """The key frame's iterations as CUDA-graph replays: iteration 0 runs eagerly in exact mode (it tells the instance
    count), the others replay ONE captured iteration with the step's cameras / targets copied into its static buffers."""
    def inputs(step_views):
        if len(step_views) == 1:
            v = step_views[0]
            return select_views(views, step_views), gt[v:v + 1]
        return select_views(views, step_views), gt[torch.tensor(step_views, device=gt.device)]

    st0, gt0 = inputs(sched[0])
    r = refine.refine_iteration(state, st0, gt0, len(sched[0]), lambda_l1=cfg.lambda_l1)
    losses[0] = r.loss
    redone = 0
    if len(sched) > 1:
        # Capturing costs far more than a key frame (torch.cuda.graph synchronises, collects garbage and empties the
        # allocator cache), so the captured iteration is kept on the state and reused by later key frames as long as the
        # shapes, the parameter storage and the capacity still fit.
        key = (state.P, len(sched[0]), int(views.image_height), int(views.image_width), float(cfg.lambda_l1),
               state.flat.data_ptr(), state.exp_avg.data_ptr(), tuple(sorted(state.lrs.items())))
        g = getattr(state, "_graphed", None)
        if g is None or g.key != key or g.capacity < int(r.num_rendered * 1.1):
            cap = (int(r.num_rendered * 1.3) + 4095) // 4096 * 4096 + 4096
            g = refine.GraphedIteration(state, st0, gt0, len(sched[0]), cap, cfg.lambda_l1)
            g.key = key
            state._graphed = g
        g.begin(state.step)
        for it in range(1, len(sched)):
            st, gts = inputs(sched[it])
            losses[it] = g.run(st.viewmatrices, st.projmatrices, st.campos, gts)
        redone = g.finish()
        if redone:
            state._graphed = None   # the capacity was too small: capture again (with the new count) next time
        for k in range(redone):   # overflowed replays skipped their step: run as many eager iterations
            st, gts = inputs(sched[-1 - k % len(sched)])
            refine.refine_iteration(state, st, gts, len(sched[-1]), lambda_l1=cfg.lambda_l1)
    return KeyframeResult(losses, sched, state.P, [], redone)


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
