"""One refine iteration of IGS's per-key-frame loop on the B200-native rasterizer.

Reference: infer_batch.py:279-324 (render -> 0.8*L1 + 0.2*(1-SSIM) -> backward -> Adam(eps=1e-15)),
igs/models/gaussian_model.py:89-127 (activations) and :265-348 (parameter groups / learning rates),
igs/utils/loss_utils.py:16-62 (l1_loss, ssim).  The reference renders ONE view per Adam step; the
schedule here is BASELINE.json's: V views per step, sharded by view over the ranks of a process
group, per-Gaussian gradients summed with one NCCL all-reduce over a packed [P*59] buffer, then an
identical fused Adam step on every rank.  `views_per_step=1` on one rank is the
reference's schedule.

The loss can run through torch (`loss_impl="torch"`, a restatement of loss_utils.py with grouped
conv2d) or through the hand-written fused L1+SSIM kernel (`loss_impl="fused"`, csrc/loss.cu) that
writes dL/dcolor directly; both give the same value/gradient within fp32 rounding (tests).
"""
from __future__ import annotations

import math
from typing import Dict, NamedTuple, Optional

import torch
import torch.nn.functional as F

from . import _C
from .multiview import MultiViewSettings, rasterize_views

# configs/demo.yaml:66-71 `training_lr`
DEFAULT_LRS = dict(xyz=0.0016, shs=0.0025, opacity=0.05, scaling=0.005, rotation=0.01)
# packing order of the [P*59] gradient / parameter record
GROUPS = (("xyz", 3), ("rotation", 4), ("shs", 48), ("opacity", 1), ("scaling", 3))
FLOATS_PER_GAUSSIAN = sum(n for _, n in GROUPS)  # 59


class RefineState:
    """Refine parameters (pre-activation, like GaussianModel._xyz/_shs/_opacity/_scaling/_rotation),
    Adam moments and the packed gradient buffer, all resident on one device."""

    def __init__(self, xyz, log_scale, rot, opacity_logit, sh, device, lrs: Optional[Dict[str, float]] = None):
        P = xyz.shape[0]
        self.P = P
        self.device = torch.device(device)
        self.lrs = dict(DEFAULT_LRS if lrs is None else lrs)
        src = dict(xyz=xyz, rotation=rot, shs=sh.reshape(P, 48), opacity=opacity_logit.reshape(P, 1), scaling=log_scale)
        # one flat buffer each for params / grads / exp_avg / exp_avg_sq, grouped (not interleaved)
        n = P * FLOATS_PER_GAUSSIAN
        self.flat = torch.empty(n, device=self.device, dtype=torch.float32)
        self.grad = torch.zeros(n, device=self.device, dtype=torch.float32)
        self.exp_avg = torch.zeros(n, device=self.device, dtype=torch.float32)
        self.exp_avg_sq = torch.zeros(n, device=self.device, dtype=torch.float32)
        self.slices = {}
        off = 0
        for name, k in GROUPS:
            self.slices[name] = (off, off + P * k, k)
            self.flat[off:off + P * k].copy_(src[name].reshape(-1).to(self.device, non_blocking=True))
            off += P * k
        self.step = 0
        # Capacity mode of the forward (no host read of the instance count, igs_forward_args.instance_capacity):
        # off until enable_async_binning(); the high-water mark of the observed instance counts sizes the buffers.
        self.async_binning = False
        self.capacity_margin = 1.2
        self._r_seen = {}             # number of local views -> largest instance count observed
        self._pending = []            # (n_views, pinned host mirror of `status`, event) not yet looked at
        self.overflows = 0            # forwards whose capacity was exceeded (their Adam step was skipped on the device)
        self.status = torch.zeros(2, dtype=torch.int32, device=self.device)
        self.last_guarded = False     # did the latest refine_iteration run its forward in capacity mode?
        self.fused = None             # symmetric buffers + handles of the fused exchange (enable_fused_exchange)

    # ---- fused exchange over NVLink / NVSwitch multicast ------------------------------------------------------------
    def enable_fused_exchange(self, group) -> bool:
        """Moves the parameter and gradient records into SYMMETRIC memory bound to a multicast object
        (torch.distributed._symmetric_memory: plumbing) so that the exchange step of a view-sharded iteration is ONE
        kernel per rank — reduce-scatter of the gradients in the switch, Adam on the rank's 1/N slice, broadcast of the
        updated slice (csrc/exchange.cu, igs_fused_reduce_adam) — instead of ncclAllReduce + a replicated Adam launch
        (+ two 4-byte all-reduces for the loss and the overflow flag).  Returns False, leaving everything as it was,
        where multicast is not available (then `reduce_gradients` + `adam_step` stay in use).  Collective: every rank
        of `group` must call it."""
        import torch.distributed as dist
        self.fused = None
        world = dist.get_world_size(group) if group is not None else 1
        if world <= 1:
            return False
        try:
            import torch.distributed._symmetric_memory as symm_mem
            n = self.P * FLOATS_PER_GAUSSIAN
            n_pad = (n + 3) // 4 * 4
            flat = symm_mem.empty(n_pad, dtype=torch.float32, device=self.device)
            grad = symm_mem.empty(n_pad + 4, dtype=torch.float32, device=self.device)   # + tail slot {loss, overflow, 0, 0}
            h_flat = symm_mem.rendezvous(flat, group)
            h_grad = symm_mem.rendezvous(grad, group)
            ok = int(h_flat.multicast_ptr) != 0 and int(h_grad.multicast_ptr) != 0   # 0: no NVLS multicast on this system
        except Exception as ex:   # no symmetric memory / multicast on this system: keep the NCCL path
            self.fused_error = f"{type(ex).__name__}: {ex}"
            ok = False
        flag = torch.tensor([1.0 if ok else 0.0], device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if float(flag.item()) == 0.0:
            return False
        flat.zero_(); grad.zero_()
        flat[:n].copy_(self.flat)
        grad[:n].copy_(self.grad)
        self.flat, self.grad = flat[:n], grad[:n]
        def pad(t):
            o = torch.zeros(n_pad, device=self.device)
            o[:n].copy_(t)
            return o
        self._m_pad, self._v_pad = pad(self.exp_avg), pad(self.exp_avg_sq)
        self.exp_avg, self.exp_avg_sq = self._m_pad[:n], self._v_pad[:n]
        rank = dist.get_rank(group)
        chunk = ((n_pad // 4) + world - 1) // world * 4
        self.fused = dict(group=group, flat=flat, grad=grad, h_flat=h_flat, h_grad=h_grad, n=n, n_pad=n_pad,
                          begin=min(n_pad, rank * chunk), end=min(n_pad, (rank + 1) * chunk),
                          loss=torch.zeros(1, device=self.device), skipped=torch.zeros(1, dtype=torch.int32, device=self.device))
        h_grad.barrier(channel=0)
        return True

    def fused_exchange_step(self, partial_loss: torch.Tensor, guarded: bool) -> torch.Tensor:
        """gradients of all ranks -> Adam on this rank's slice -> parameters of all ranks, one kernel (see
        enable_fused_exchange).  Returns the step's loss (sum of the ranks' partial losses), a device scalar."""
        f = self.fused
        n, n_pad = f["n"], f["n_pad"]
        tail = f["grad"][n_pad:n_pad + 2]
        tail[0:1].copy_(partial_loss.reshape(1))
        if guarded:
            tail[1:2].copy_(self.status[1:2])          # int32 0 / 1 -> float
        else:
            tail[1:2].zero_()
        self.step += 1
        sizes = [self.P * k for _, k in GROUPS]
        sizes[-1] += n_pad - n                           # the pad floats ride in the last group (zero gradient: no-op)
        f["h_grad"].barrier(channel=0)                   # every rank's gradients and tail slot are complete
        _C.fused_reduce_adam(int(f["h_grad"].multicast_ptr), int(f["h_flat"].multicast_ptr), f["flat"], self._m_pad, self._v_pad,
                             n_pad, f["begin"], f["end"], sizes, [float(self.lrs[k]) for k, _ in GROUPS], 0.9, 0.999, 1e-15,
                             self.step, f["loss"], f["skipped"])
        f["h_grad"].barrier(channel=1)                   # every rank's broadcast has landed; gradients may be rewritten
        return f["loss"][0]

    # ---- capacity mode bookkeeping -------------------------------------------------------------------------------
    def enable_async_binning(self, on: bool = True, margin: float = 1.2) -> None:
        """From the next iteration on, forwards for a view count whose instance total has been seen once run without
        the host round-trip (reference rasterizer_impl.cu:354): buffers are sized margin x the largest total seen, the
        true total stays on the device, and an overflowing iteration turns its own Adam step into a no-op."""
        self.async_binning = bool(on)
        self.capacity_margin = float(margin)

    def note_instances(self, n_views: int, R: int) -> None:
        if R > self._r_seen.get(n_views, 0):
            self._r_seen[n_views] = int(R)

    def poll_status(self, wait: bool = False) -> None:
        """Looks at the status words of finished capacity-mode iterations (no synchronisation unless `wait`)."""
        keep = []
        for n_views, host, ev in self._pending:
            if wait:
                ev.synchronize()
            if ev.query():
                self.note_instances(n_views, int(host[0]))
                if int(host[1]) != 0:
                    self.overflows += 1
            else:
                keep.append((n_views, host, ev))
        self._pending = keep

    def instance_capacity(self, n_views: int) -> int:
        """0 = exact mode (count read back by the host); otherwise the capacity for a forward over n_views views."""
        if not self.async_binning:
            return 0
        self.poll_status()
        seen = self._r_seen.get(n_views, 0)
        if seen <= 0:
            return 0
        return min(0x7FFFFFFF, (int(seen * self.capacity_margin) + 4095) // 4096 * 4096)

    def track_status(self, n_views: int) -> None:
        host = torch.empty(2, dtype=torch.int32, pin_memory=True)
        host.copy_(self.status, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._pending.append((n_views, host, ev))
        if len(self._pending) > 64:
            self.poll_status(wait=True)

    def resize(self, P: int, flat: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor) -> None:
        """Adopts re-laid-out records of a new size (densify / prune, igs_b200/densify.py); Adam's step count is kept,
        like the reference, which keeps `state["step"]` of every surviving parameter group."""
        n = P * FLOATS_PER_GAUSSIAN
        assert flat.numel() == n and exp_avg.numel() == n and exp_avg_sq.numel() == n
        self.P = P
        self.flat, self.exp_avg, self.exp_avg_sq = flat, exp_avg, exp_avg_sq
        self.grad = torch.zeros(n, device=self.device, dtype=torch.float32)
        self._r_seen = {}             # instance counts of the old Gaussian set say nothing about the new one
        self.fused = None             # the symmetric buffers had the old size: call enable_fused_exchange again
        off = 0
        for name, k in GROUPS:
            self.slices[name] = (off, off + P * k, k)
            off += P * k

    def view(self, buf: torch.Tensor, name: str) -> torch.Tensor:
        a, b, k = self.slices[name]
        t = buf[a:b]
        return t.view(self.P, 16, 3) if name == "shs" else t.view(self.P, k)

    def param(self, name: str) -> torch.Tensor:
        return self.view(self.flat, name)

    def load_flat(self, host_flat: torch.Tensor) -> None:
        """H2D copy of all parameters from a (pinned) host buffer laid out like `self.flat`."""
        self.flat.copy_(host_flat, non_blocking=True)

    def adam_step(self, guarded: bool = False) -> None:
        """torch.optim.Adam(lr per group, betas=(0.9,0.999), eps=1e-15) — gaussian_model.py:348 — as one
        fused kernel launch over all parameter groups (igs_adam_step_groups).  guarded: the step becomes a no-op on
        the device when this iteration's capacity-mode forward overflowed (self.status[1])."""
        self.step += 1
        _C.adam_step_groups(self.flat, self.grad, self.exp_avg, self.exp_avg_sq,
                            [self.P * k for _, k in GROUPS], [float(self.lrs[n]) for n, _ in GROUPS],
                            0.9, 0.999, 1e-15, self.step, skip=self.status[1:2] if guarded else None)

    def offsets(self):
        """Element offsets of (xyz, rotation, shs, opacity, scaling) in the packed record (refine_grad's `offs`)."""
        return [self.slices[n][0] for n in ("xyz", "rotation", "shs", "opacity", "scaling")]


# ------------------------------------------------------------------------------------------------
# loss (reference igs/utils/loss_utils.py:16-62), torch restatement
# ------------------------------------------------------------------------------------------------
_WINDOWS: Dict[tuple, torch.Tensor] = {}


def _window(channel: int, device, dtype) -> torch.Tensor:
    key = (channel, str(device), dtype)
    if key not in _WINDOWS:
        g = torch.tensor([math.exp(-(x - 11 // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
        g = (g / g.sum()).unsqueeze(1)
        w2 = g.mm(g.t()).float().unsqueeze(0).unsqueeze(0)
        _WINDOWS[key] = w2.expand(channel, 1, 11, 11).contiguous().to(device=device, dtype=dtype)
    return _WINDOWS[key]


def ssim_map_torch(img1: torch.Tensor, img2: torch.Tensor) -> torch.Tensor:
    """[V,3,H,W] x [V,3,H,W] -> SSIM map [V,3,H,W] (loss_utils.py:43-57)."""
    ch = img1.shape[-3]
    w = _window(ch, img1.device, img1.dtype)
    mu1 = F.conv2d(img1, w, padding=5, groups=ch)
    mu2 = F.conv2d(img2, w, padding=5, groups=ch)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = F.conv2d(img1 * img1, w, padding=5, groups=ch) - mu1_sq
    s2 = F.conv2d(img2 * img2, w, padding=5, groups=ch) - mu2_sq
    s12 = F.conv2d(img1 * img2, w, padding=5, groups=ch) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return ((2 * mu1_mu2 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))


def refine_loss_torch(pred: torch.Tensor, gt: torch.Tensor, lambda_l1: float = 0.8) -> torch.Tensor:
    """Per-view loss [V]: lambda*L1 + (1-lambda)*(1-SSIM) (infer_batch.py:302-305)."""
    l1 = (pred - gt).abs().mean(dim=(1, 2, 3))
    ss = ssim_map_torch(pred, gt).mean(dim=(1, 2, 3))
    return lambda_l1 * l1 + (1.0 - lambda_l1) * (1.0 - ss)


class _FusedLoss(torch.autograd.Function):
    """Sum over views of the per-view refine loss, value and dL/dpred from csrc/loss.cu."""

    @staticmethod
    def forward(ctx, pred, gt, lambda_l1, loss_rows=None):
        loss_sum, dpred = _C.refine_loss(pred, gt, float(lambda_l1), 1.0, list(loss_rows) if loss_rows else [])
        ctx.save_for_backward(dpred)
        return loss_sum

    @staticmethod
    def backward(ctx, g):
        (dpred,) = ctx.saved_tensors
        return dpred * g, None, None, None


def refine_loss_sum(pred, gt, lambda_l1=0.8, impl="fused", loss_rows=None):
    if impl == "fused":
        return _FusedLoss.apply(pred.contiguous(), gt.contiguous(), lambda_l1, loss_rows)
    if loss_rows:
        raise ValueError("row-restricted losses need the fused loss kernel")
    return refine_loss_torch(pred, gt, lambda_l1).sum()


# ------------------------------------------------------------------------------------------------
# the iteration
# ------------------------------------------------------------------------------------------------
class IterationResult(NamedTuple):
    loss: torch.Tensor           # scalar, mean over ALL views of the step (after all-reduce)
    num_rendered: int
    radii: torch.Tensor          # [V_local,P]
    grad_means2D: Optional[torch.Tensor]  # [V_local,P,3] screen-space grads (densification statistics)


def forward_views(state: RefineState, settings: MultiViewSettings):
    """Activations (gaussian_model.py:89-127) + batched rasterization; returns the 8 output maps."""
    xyz = state.param("xyz")
    scales = torch.exp(state.param("scaling"))
    rot = F.normalize(state.param("rotation"))
    opac = torch.sigmoid(state.param("opacity"))
    return rasterize_views(xyz, state.param("shs"), opac, scales, rot, settings)


def refine_iteration(state: RefineState, settings: MultiViewSettings, gt: torch.Tensor, total_views: int,
                     group=None, lambda_l1: float = 0.8, loss_impl: str = "fused", do_step: bool = True,
                     impl: str = "native") -> IterationResult:
    """One Adam step on `total_views` views, of which this rank renders `settings.viewmatrices.shape[0]`.

    loss = (1/total_views) * sum_v [lambda*L1_v + (1-lambda)*(1-SSIM_v)]; gradients are summed over ranks.

    impl="native" (default with the fused loss) runs activations -> forward -> loss -> backward -> activation chain
    rule in ONE native call (`_C.refine_grad`) that writes the packed gradient record in place: no autograd graph,
    no per-op Python.  impl="autograd" is the same computation through torch autograd over `rasterize_views`
    (what a reference user's loop looks like); the two agree to fp32 rounding (tests/test_refine_gpu.py).
    """
    Vl = int(settings.viewmatrices.shape[0]) if settings is not None else 0
    if impl == "native" and loss_impl == "fused" and Vl > 0:
        s = settings
        cap = state.instance_capacity(Vl)
        loss, R, radii, g2d, _status = _C.refine_grad(
            state.flat, state.grad, state.offsets(), state.P, s.bg, s.viewmatrices, s.projmatrices, s.campos,
            float(s.tanfovx), float(s.tanfovy), float(s.kernel_size), float(s.scale_modifier), int(s.image_height),
            int(s.image_width), int(s.sh_degree), gt, int(total_views), float(lambda_l1), bool(s.require_coord),
            bool(s.require_depth), bool(s.debug), s.tile_rows,
            list(s.loss_rows) if s.loss_rows else [], cap, state.status)
        if cap == 0:
            state.note_instances(Vl, int(R))     # exact mode: R is the true count
        else:
            state.track_status(Vl)               # capacity mode: the count is looked at later, without a sync
        guarded = cap != 0
        state.last_guarded = guarded
        if state.fused is not None and do_step and group is not None and state.fused["group"] is group:
            loss = state.fused_exchange_step(loss, guarded)
            return IterationResult(loss, int(R), radii, g2d)
        if guarded and group is not None and torch.distributed.get_world_size(group) > 1:
            # one rank's overflow must skip the step on EVERY rank: fold the flag into the exchange
            torch.distributed.all_reduce(state.status[1:2], op=torch.distributed.ReduceOp.MAX, group=group)
        loss = reduce_gradients(state.grad, loss, group)
        if do_step:
            state.adam_step(guarded=guarded)
        return IterationResult(loss, int(R), radii, g2d)
    state.last_guarded = False
    names = [n for n, _ in GROUPS]
    leaves = []
    for n in names:
        p = state.param(n).detach().requires_grad_(True)
        leaves.append(p)
    lv = dict(zip(names, leaves))
    if Vl > 0:
        scales = torch.exp(lv["scaling"])
        rot = F.normalize(lv["rotation"])
        opac = torch.sigmoid(lv["opacity"])
        outs = rasterize_views(lv["xyz"], lv["shs"], opac, scales, rot, settings)
        color, radii = outs[0], outs[1]
        loss = refine_loss_sum(color, gt, lambda_l1, loss_impl, settings.loss_rows) / float(total_views)
        grads = torch.autograd.grad(loss, leaves, allow_unused=True)
        for n, g in zip(names, grads):
            a, b, _k = state.slices[n]
            if g is None:
                state.grad[a:b].zero_()
            else:
                state.grad[a:b].copy_(g.reshape(-1))
        loss = loss.detach()
        fn = color.grad_fn
        g2d = getattr(fn, "grad_means2D", None)
        R = int(getattr(fn, "num_rendered", 0))
    else:  # a rank without views still takes part in the all-reduce and the Adam step
        state.grad.zero_()
        loss = torch.zeros((), device=state.device)
        radii, g2d, R = torch.zeros(0, state.P, dtype=torch.int32, device=state.device), None, 0
    loss = reduce_gradients(state.grad, loss, group)
    if do_step:
        state.adam_step()
    return IterationResult(loss, R, radii, g2d)


def reduce_gradients(flat_grad: torch.Tensor, loss: torch.Tensor, group=None) -> torch.Tensor:
    """The one exchange step of the view-sharded iteration: SUM all-reduce of the packed [P*59] gradient
    buffer (in place) and of the already 1/total_views-scaled partial loss.  Backend-agnostic (NCCL on the GPU
    box, gloo in the CPU tests)."""
    if group is None or torch.distributed.get_world_size(group) <= 1:
        return loss
    torch.distributed.all_reduce(flat_grad, group=group)
    loss = loss.clone()
    torch.distributed.all_reduce(loss, group=group)  # 4 bytes; kept separate so `flat_grad` stays the plain record
    return loss


def pack_gradients(grads: dict, P: int) -> torch.Tensor:
    """Packs per-parameter gradients into the [P*59] record order of GROUPS."""
    return torch.cat([grads[n].reshape(P * k) for n, k in GROUPS])


def shard_views_balanced(costs, rank: int, world: int):
    """Longest-processing-time-first packing of views onto ranks by a per-view cost (e.g. instance counts of the
    previous iteration); deterministic, so every rank derives the same assignment.  Pixel work is equal per view,
    so each view's cost gets a constant pixel term added by the caller if wanted."""
    order = sorted(range(len(costs)), key=lambda v: (-costs[v], v))
    load = [0.0] * world
    mine = [[] for _ in range(world)]
    for v in order:
        r = min(range(world), key=lambda i: (load[i], i))
        load[r] += float(costs[v])
        mine[r].append(v)
    return sorted(mine[rank])


def shard_views(total_views: int, rank: int, world: int):
    """View v -> rank v mod world: 20 views on 8 ranks = 3,3,3,3,2,2,2,2."""
    return list(range(rank, total_views, world))


# ------------------------------------------------------------------------------------------------
# finer than whole views: bands of tile rows
# ------------------------------------------------------------------------------------------------
SSIM_RADIUS = 5   # 11x11 window: the loss terms of a pixel row read rendered rows +- 5


class Unit(NamedTuple):
    """A band of one view owned by a rank: its loss terms are those of pixel rows [row0, row1)."""
    view: int
    row0: int
    row1: int


def shard_units(total_views: int, height: int, rank: int, world: int):
    """Whole views when `world` divides `total_views`; otherwise every view is cut into world/gcd(total_views, world)
    bands of tile rows so that all ranks get the same number of bands (20 views on 8 ranks: 40 half-views, 5 per rank
    = 2.5 views each instead of 3,3,3,3,2,2,2,2).  A rank's bands are contiguous in (view, band) order and adjacent
    bands of one view are merged.  Each band's tile-row window is widened by the SSIM radius (`unit_windows`), and
    per-pixel results inside a window are identical to the full render, so the sum over ranks of the band losses /
    gradients equals the single-rank step (tests/test_refine_gpu.py)."""
    if world <= 1 or total_views % world == 0:
        return [Unit(v, 0, height) for v in range(rank, total_views, world)] if world > 1 else \
               [Unit(v, 0, height) for v in range(total_views)]
    parts = world // math.gcd(total_views, world)
    tile_rows = (height + 15) // 16
    if parts > tile_rows:
        return [Unit(v, 0, height) for v in range(rank, total_views, world)]
    n_units = total_views * parts
    lo, hi = rank * n_units // world, (rank + 1) * n_units // world
    units = []
    for u in range(lo, hi):
        v, k = divmod(u, parts)
        r0 = min(height, 16 * (k * tile_rows // parts))
        r1 = min(height, 16 * ((k + 1) * tile_rows // parts))
        if units and units[-1].view == v and units[-1].row1 == r0:
            units[-1] = Unit(v, units[-1].row0, r1)
        else:
            units.append(Unit(v, r0, r1))
    return units


def unit_windows(units, height: int):
    """(tile_rows [U,2] int32 CPU tensor, loss_rows flat list) for `units`: the tile-row window that must be rendered
    so that rows [row0 - 5, row1 + 5) are exact, and the rows whose loss terms the unit owns.  (None, None) when every
    unit is a whole view."""
    if all(u.row0 == 0 and u.row1 >= height for u in units):
        return None, None
    t_all = (height + 15) // 16
    tr, lr = [], []
    for u in units:
        tr.append([max(0, (u.row0 - SSIM_RADIUS) // 16), min(t_all, (u.row1 + SSIM_RADIUS + 15) // 16)])
        lr += [u.row0, u.row1]
    return torch.tensor(tr, dtype=torch.int32), lr


def settings_for_units(all_views: MultiViewSettings, units) -> MultiViewSettings:
    """The MultiViewSettings of a rank that owns `units` of the views in `all_views` (one entry per unit: a view that
    is split appears once per band).  The matching targets are `gt_all[[u.view for u in units]]`."""
    idx = torch.tensor([u.view for u in units], dtype=torch.long, device=all_views.viewmatrices.device)
    tr, lr = unit_windows(units, int(all_views.image_height))
    return all_views._replace(viewmatrices=all_views.viewmatrices[idx].contiguous(),
                              projmatrices=all_views.projmatrices[idx].contiguous(),
                              campos=all_views.campos[idx].contiguous(),
                              tile_rows=None if tr is None else tr.to(all_views.viewmatrices.device),
                              loss_rows=lr)


# ------------------------------------------------------------------------------------------------
# the single-rank iteration as ONE CUDA graph
# ------------------------------------------------------------------------------------------------
class GraphedIteration:
    """activations -> forward -> loss -> backward -> Adam of one refine iteration captured in a CUDA graph.

    Possible because the capacity mode of the forward needs no host read of the instance count (the reference's
    blocking cudaMemcpy, rasterizer_impl.cu:354, is what makes its iteration un-capturable) and because the Adam step
    count lives on the device (`igs_adam_step_groups_guarded`).  The cameras and the target image are static buffers:
    `run(views, gt)` copies the iteration's inputs into them and replays.  An iteration whose instance count exceeds
    the captured capacity skips its own Adam step on the device; `overflowed()` tells (it synchronises).

    Single rank, fixed number of views per step, no tile-row bands.  Built by `stream.refine_keyframe(graph=True)`.
    """

    def __init__(self, state: RefineState, settings: MultiViewSettings, gt: torch.Tensor, total_views: int,
                 capacity: int, lambda_l1: float = 0.8):
        assert settings.tile_rows is None and not settings.loss_rows, "bands are not supported in the graphed iteration"
        self.state = state
        self.s = settings._replace(viewmatrices=settings.viewmatrices.clone(), projmatrices=settings.projmatrices.clone(),
                                   campos=settings.campos.clone())
        self.gt = gt.clone()
        self.total_views = int(total_views)
        self.capacity = int(capacity)
        self.lambda_l1 = float(lambda_l1)
        self.step_dev = torch.full((1,), state.step, dtype=torch.int32, device=state.device)
        self.status = torch.zeros(2, dtype=torch.int32, device=state.device)
        self.n_overflow = torch.zeros(1, dtype=torch.int32, device=state.device)
        self.loss = torch.zeros((), device=state.device)
        self.radii = None
        self.g2d = None
        self.replays = 0
        side = torch.cuda.Stream(device=state.device)
        side.wait_stream(torch.cuda.current_stream(state.device))
        with torch.cuda.stream(side):           # warm-up outside the capture (lazy initialisation, allocator pools)
            snap = [t.clone() for t in (state.flat, state.exp_avg, state.exp_avg_sq)]
            self._body()
            for t, c in zip((state.flat, state.exp_avg, state.exp_avg_sq), snap):
                t.copy_(c)
            self.step_dev.fill_(state.step)
            self.n_overflow.zero_()
        torch.cuda.current_stream(state.device).wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, capture_error_mode="relaxed"):
            self._body()

    def _body(self):
        st, s = self.state, self.s
        loss, _R, radii, g2d, _ = _C.refine_grad(
            st.flat, st.grad, st.offsets(), st.P, s.bg, s.viewmatrices, s.projmatrices, s.campos, float(s.tanfovx),
            float(s.tanfovy), float(s.kernel_size), float(s.scale_modifier), int(s.image_height), int(s.image_width),
            int(s.sh_degree), self.gt, self.total_views, self.lambda_l1, bool(s.require_coord), bool(s.require_depth),
            False, None, [], self.capacity, self.status)
        self.loss.copy_(loss)
        self.radii, self.g2d = radii, g2d
        self.n_overflow.add_(self.status[1:2])
        _C.adam_step_groups(st.flat, st.grad, st.exp_avg, st.exp_avg_sq, [st.P * k for _, k in GROUPS],
                            [float(st.lrs[n]) for n, _ in GROUPS], 0.9, 0.999, 1e-15, 1, skip=self.status[1:2],
                            step_dev=self.step_dev)

    def begin(self, step: int) -> None:
        """Start of a run of replays: the device-side Adam step count takes the state's, the overflow counter restarts."""
        self.step_dev.fill_(int(step))
        self.n_overflow.zero_()

    def run(self, viewmatrices, projmatrices, campos, gt) -> torch.Tensor:
        """One iteration on the given cameras / targets (device tensors of the captured shapes); returns the loss
        (a device scalar that the next run overwrites)."""
        self.s.viewmatrices.copy_(viewmatrices)
        self.s.projmatrices.copy_(projmatrices)
        self.s.campos.copy_(campos)
        self.gt.copy_(gt)
        self.graph.replay()
        self.replays += 1
        return self.loss

    def finish(self) -> int:
        """Folds the device-side step count back into the state; returns the number of overflowed (skipped) iterations."""
        self.state.step = int(self.step_dev.item())
        return int(self.n_overflow.item())
