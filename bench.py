#!/usr/bin/env python
"""bench.py — refine-iteration benchmark (BASELINE.json metric) for the B200-native splat rasterizer.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C3]

One "step" = one refine iteration on the N3DV-shaped workload (BASELINE.json configs[2], SURVEY §8d
"C3"): 300k Gaussians SH3, 20 views 1352x1014, activations -> rasterize fwd -> 0.8*L1+0.2*(1-SSIM)
-> backward -> [NCCL all-reduce of the packed per-Gaussian gradients] -> fused Adam.  With N ranks the
20 views are sharded v -> rank v mod N (strong scaling: the work of one step is fixed).

Printed JSON line (rank 0):
  value/ms_per_step  device-timed ms per refine iteration with every input resident in HBM
  e2e                the same iteration with the cameras and target images of the step's views copied from
                     pinned HOST memory inside the timed region and the loss copied back to the host every
                     step (asynchronously, consumed one step later; `value_with_blocking_readback` = the same
                     with a blocking .item() after every step)
  fwd_mpix_s         forward-only throughput (the metric's second half)
  roofline           dominant stage of the step: algorithmic bytes / CUDA-event duration
  cpu_baseline       the CPU oracle (oracle/raster_oracle.c, a restatement: the reference has no CPU
                     path) timed on this box's host cores on a bounded sample, N=1 only
  ref_cuda           (extra) the UNMODIFIED reference CUDA rasterizer (baseline/_ref) on the same views
                     with the same torch loss, when its build travelled to this box
`--impl reference` times the CPU oracle port alone (the reference rasterizer is CUDA-only, so its
"CPU implementation" is the restatement; kind = "port").
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "refine_iter_ms"
UNIT = "ms"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--P", type=int, default=0, help="override the number of Gaussians (debug)")
    ap.add_argument("--views", type=int, default=0, help="override the number of views (debug)")
    ap.add_argument("--loss", default="auto", choices=["auto", "fused", "torch"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-sample-views", type=int, default=2)
    ap.add_argument("--no-bands", dest="bands", action="store_false",
                    help="shard whole views only (v mod N) even when N does not divide the number of views")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="N > 1: fused = one kernel per rank over NVLS multicast (reduce-scatter + Adam on the slice + "
                         "broadcast, csrc/exchange.cu; falls back to nccl where multicast is unavailable); nccl = "
                         "ncclAllReduce + replicated Adam")
    ap.add_argument("--exact-binning", action="store_true",
                    help="read the instance count back on the host in every forward (reference behaviour) instead of the "
                         "capacity mode (igs_forward_args.instance_capacity)")
    ap.add_argument("--no-balance", dest="balance", action="store_false",
                    help="assign view v to rank v mod N instead of balancing by per-view instance counts")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# clocks sampled DURING the timed region (B200_PROFILING.md recipe)
# ------------------------------------------------------------------------------------------------
sys.path.insert(0, str(ROOT / "tools"))
from clocks import ClockSampler  # noqa: E402  (tools/clocks.py: NVML / nvidia-smi sampler shared with tools/bench_*.py)


# ------------------------------------------------------------------------------------------------
# CPU oracle leg (cpu_baseline and --impl reference)
# ------------------------------------------------------------------------------------------------
def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_oracle_ms(cfg, view_ids, with_loss=True):
    """Times the CPU arm on the views `view_ids` of the workload: oracle forward -> torch-CPU L1+SSIM and its gradient
    -> oracle backward per view, plus ONE torch-CPU Adam step over the 17.7 M parameters per call (the same work as
    one GPU step restricted to those views).  The oracle's OpenMP thread count is set to every core this process may
    use — torchrun exports OMP_NUM_THREADS=1 to its ranks, which would otherwise time one core.
    Returns (ms per view, ms of the Adam step, threads used)."""
    import numpy as np
    import torch
    from igs_b200 import refine, scenes
    sys.path.insert(0, str(ROOT / "oracle"))
    import oracle_py  # noqa: E402  (ctypes wrapper over oracle/raster_oracle.c; test infrastructure)

    cores = host_cores()
    threads = oracle_py.set_num_threads(cores)
    torch.set_num_threads(cores)
    P, W, H, V = cfg["P"], cfg["W"], cfg["H"], cfg["V"]
    sc = scenes.make_gaussians(P, 0, cfg["log_scale_mean"])
    cams = scenes.make_cameras(V, W, H, 1)
    scales, rot, opac = scenes.activated(sc)
    times = []
    for v in view_ids:
        gt = torch.rand(1, 3, H, W, generator=torch.Generator().manual_seed(2 + v))
        t0 = time.perf_counter()
        fw = oracle_py.forward(sc.xyz.numpy(), scales.numpy(), rot.numpy(), opac.numpy().reshape(-1), sc.sh.numpy(),
                               cams.viewmatrix[v].numpy(), cams.projmatrix[v].numpy(), cams.campos[v].numpy(),
                               W, H, cams.tanfovx, cams.tanfovy, sh_degree=3)
        if with_loss:
            pred = torch.from_numpy(np.ascontiguousarray(fw["image"]["color"])).reshape(1, 3, H, W).requires_grad_(True)
            refine.refine_loss_torch(pred, gt).sum().backward()
            dcol = pred.grad[0].numpy()
        else:
            dcol = torch.randn(3, H, W, generator=torch.Generator().manual_seed(3)).numpy()
        oracle_py.backward(fw, dL_color=dcol)
        times.append((time.perf_counter() - t0) * 1e3)
    n = P * 59
    par = torch.zeros(n, requires_grad=True)
    par.grad = torch.full((n,), 1e-3)
    opt = torch.optim.Adam([par], lr=1e-3, eps=1e-15)
    opt.step()
    t0 = time.perf_counter()
    opt.step()
    adam_ms = (time.perf_counter() - t0) * 1e3
    return float(np.mean(times)), adam_ms, threads


def cpu_sample_views(V, step, n):
    """n distinct views for CPU step `step`: evenly spread over the arc and rotated with the step, so that consecutive
    steps time different views (round 1 timed view 0 only)."""
    n = max(1, min(n, V))
    return [int((step + k * V // n) % V) for k in range(n)]


def run_reference_arm(args, cfg, rank):
    if rank != 0:
        return 0
    sample_views = max(2, args.cpu_sample_views)
    V = cfg["V"]
    # warm-up + steps, each a bounded sample (sample_views of the V views), scaled to the metric's unit
    vals = []
    threads = 1
    for i in range(args.warmup + args.steps):
        ms_view, adam_ms, threads = cpu_oracle_ms(cfg, cpu_sample_views(V, i, sample_views))
        if i >= args.warmup:
            vals.append(ms_view * V + adam_ms)
    ms = sum(vals) / len(vals)
    sample = (f"{sample_views} of {V} views per step (views rotate with the step: {cpu_sample_views(V, 0, sample_views)}, "
              f"{cpu_sample_views(V, 1, sample_views)}, ...): oracle fwd + torch-CPU L1+SSIM fwd/bwd + oracle bwd, scaled "
              f"x{V}/{sample_views}, + one torch-CPU Adam step over {cfg['P'] * 59} parameters")
    line = {
        "impl": "reference", "metric": METRIC, "value": ms, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, cfg),
        "cpu_baseline": {"value": ms, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": ms, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference rasterizer is CUDA-only; this arm is the CPU restatement (oracle/) on host cores",
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(args, cfg):
    return {"workload": f"{args.workload}: N3DV-shaped refine iteration, {cfg['P']} Gaussians SH3, {cfg['V']} views "
                        f"{cfg['W']}x{cfg['H']}, fwd+loss+bwd+Adam, require_coord=require_depth=True",
            "P": cfg["P"], "views": cfg["V"], "width": cfg["W"], "height": cfg["H"],
            "parallelism": f"view-sharded x{args.gpus}", "l2": "inputs larger than L2 (working set > 1 GB per step)",
            "stationary": "parameters restored to the seeded scene at the start of every step (Adam still runs)"}


# ------------------------------------------------------------------------------------------------
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    from igs_b200 import scenes
    cfg = dict(scenes.CONFIGS[args.workload])
    if args.P:
        cfg["P"] = args.P
    if args.views:
        cfg["V"] = args.views

    if args.impl == "reference":
        return run_reference_arm(args, cfg, rank)

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this library has no CPU path (use --impl reference for the CPU oracle)")
    import torch.distributed as dist
    from igs_b200 import _C
    from igs_b200.multiview import MultiViewSettings
    from igs_b200 import refine

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD

    P, W, H, V = cfg["P"], cfg["W"], cfg["H"], cfg["V"]
    N = W * H
    loss_impl = args.loss
    if loss_impl == "auto":
        loss_impl = "fused" if hasattr(_C, "refine_loss") else "torch"

    # ---- synthetic, seeded inputs, generated on the host ----
    sc = scenes.make_gaussians(P, 0, cfg["log_scale_mean"])
    cams = scenes.make_cameras(V, W, H, 1)
    assignment = "v mod N"
    my_views = refine.shard_views(V, rank, world)
    if world > 1 and args.balance:
        # cost-balanced assignment: every rank measures the instance count R_v of each view with a forward-only
        # pass (deterministic, so all ranks derive the same table) and views are packed greedily (LPT)
        costs = measure_view_costs(cfg, cams, sc, torch.device("cuda", local_rank))
        my_views = refine.shard_views_balanced(costs, rank, world)
        assignment = "LPT on per-view instance counts"
    # N does not divide the views (20 views on 8 GPUs): cut the views into bands of tile rows so that every rank
    # renders the same share (2.5 views) instead of 3,3,3,3,2,2,2,2 (igs_b200.refine.shard_units)
    tile_rows_dev, loss_rows = None, None
    band_share = None
    if world > 1 and args.bands and V % world != 0:
        units = refine.shard_units(V, H, rank, world)
        my_views = [u.view for u in units]
        tr, loss_rows = refine.unit_windows(units, H)
        if tr is not None:
            tile_rows_dev = tr.to(dev)
            band_share = sum(u.row1 - u.row0 for u in units) / float(H)
            assignment = f"bands of tile rows: {len(units)} units = {band_share:.2f} views per rank (shard_units)"
    Vl = len(my_views)
    g = torch.Generator().manual_seed(2 + 1000 * rank)
    gt_host = torch.rand(max(Vl, 1), 3, H, W, generator=g).pin_memory()
    state = refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, dev)
    cam_host = torch.cat([cams.viewmatrix[my_views].reshape(-1), cams.projmatrix[my_views].reshape(-1),
                          cams.campos[my_views].reshape(-1)]).pin_memory() if Vl else torch.zeros(0).pin_memory()
    cam_dev = torch.empty_like(cam_host, device=dev)
    gt_dev = torch.empty(max(Vl, 1), 3, H, W, device=dev)
    bg = torch.zeros(3, device=dev)

    def settings_from(cam_flat):
        if Vl == 0:
            return None
        vm = cam_flat[:Vl * 16].view(Vl, 4, 4)
        pm = cam_flat[Vl * 16:Vl * 32].view(Vl, 4, 4)
        cp = cam_flat[Vl * 32:].view(Vl, 3)
        return MultiViewSettings(image_height=H, image_width=W, tanfovx=cams.tanfovx, tanfovy=cams.tanfovy,
                                 kernel_size=0.0, bg=bg, scale_modifier=1.0, viewmatrices=vm, projmatrices=pm,
                                 sh_degree=3, campos=cp, require_depth=True, require_coord=True, debug=False,
                                 tile_rows=tile_rows_dev, loss_rows=loss_rows)

    cam_dev.copy_(cam_host)
    gt_dev.copy_(gt_host)
    st = settings_from(cam_dev)
    gt_local = gt_dev[:Vl]

    # Stationary workload: against random targets Adam quickly drives opacities down, so the cost of a step would
    # depend on how many steps ran before it.  Every step therefore starts from the same parameters (one 70.8 MB
    # device-to-device copy inside the timed region); all of the step's work, Adam included, still runs.
    exchange = "none"
    if world > 1:
        exchange = "fused (NVLS multicast kernel)" if (args.exchange == "fused" and state.enable_fused_exchange(group)) else \
                   "nccl all-reduce + replicated Adam"
    snapshot = state.flat.clone()
    state.enable_async_binning(not args.exact_binning)   # the first step per view count runs exact and sizes the rest

    def step_resident():
        state.flat.copy_(snapshot)
        return refine.refine_iteration(state, st, gt_local, V, group=group, loss_impl=loss_impl)

    # End-to-end step: this step's INPUTS — the cameras and target images of its views, what the reference's loop moves
    # with `.to("cuda")` every iteration (infer_batch.py:287-288) — come from pinned HOST memory and the loss goes back
    # to the host; the Gaussians are model state and stay on the device, as in the reference.  The copies run on a side
    # stream into double-buffered staging, so the H2D of step k+1 overlaps the kernels of step k (every step still
    # copies all of its own inputs inside the timed region).
    copy_stream = torch.cuda.Stream(device=dev)
    stage = [dict(cams=torch.empty_like(cam_dev), gt=torch.empty_like(gt_dev),
                  ready=torch.cuda.Event(), free=torch.cuda.Event()) for _ in range(2)]
    e2e_state = {"k": 0, "primed": False, "loss": float("nan")}
    # The step's result (the loss) goes back to the host EVERY step, as an asynchronous 4-byte copy into pinned memory that
    # the host consumes one step later: the host stays one step ahead of the device instead of draining the stream after
    # every iteration (the reference's loop reads no loss at all inside the iteration, infer_batch.py:279-324).
    loss_pin = torch.empty(2, dtype=torch.float32).pin_memory()
    loss_ev = [torch.cuda.Event() for _ in range(2)]

    def prefetch(slot):
        b = stage[slot]
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(b["free"])          # the step that last used this slot has finished reading it
            b["cams"].copy_(cam_host, non_blocking=True)
            b["gt"].copy_(gt_host, non_blocking=True)
            b["ready"].record(copy_stream)

    def step_e2e():
        k = e2e_state["k"]
        if not e2e_state["primed"]:
            for b in stage:
                b["free"].record(torch.cuda.current_stream())
            prefetch(k % 2)
            e2e_state["primed"] = True
        b = stage[k % 2]
        prefetch((k + 1) % 2)                          # next step's inputs, overlapped with this step's kernels
        cur = torch.cuda.current_stream()
        cur.wait_event(b["ready"])
        state.flat.copy_(snapshot)                     # model state lives on the device, as in the reference's loop
        r = refine.refine_iteration(state, settings_from(b["cams"]), b["gt"][:Vl], V, group=group, loss_impl=loss_impl)
        b["free"].record(cur)
        e2e_state["k"] = k + 1
        if e2e_state.get("blocking"):
            return float(r.loss.item())                # (comparison protocol: drain the stream after every step)
        loss_pin[k % 2:k % 2 + 1].copy_(r.loss.detach().reshape(1), non_blocking=True)   # D2H read of the step's result
        loss_ev[k % 2].record(cur)
        if k > 0:
            loss_ev[(k - 1) % 2].synchronize()
            e2e_state["loss"] = float(loss_pin[(k - 1) % 2])
        return e2e_state["loss"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    step_stats = {}

    def timed(fn, steps, warmup, tag=None):
        for _ in range(warmup):
            fn()
        barrier()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        evs[0].record()
        last = None
        for i in range(steps):
            last = fn()
            evs[i + 1].record()
        barrier()
        ms = torch.tensor([evs[0].elapsed_time(evs[steps])], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if tag is not None:   # per-step device times of this rank (diagnostic: exposes one-off host stalls)
            per = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(steps))
            step_stats[tag] = {"min": per[0], "median": per[len(per) // 2], "max": per[-1]}
        return float(ms.item()) / steps, last

    # ---- (1) device-resident timing with stage timers + clocks ----
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    _C.stage_timing_enable(True)   # on during warm-up too, so the library's event pool is populated before timing
    for _ in range(max(args.warmup, 3)):
        step_resident()
    barrier()
    _C.stage_timing_read(True)
    launches0 = _C.kernel_launches()
    sampler.begin()
    ms_step, last = timed(step_resident, args.steps, 0, tag="resident")
    sampler.end()
    # A one-off host stall (a single step several times the median) is not the workload: like a throttled run,
    # such a measurement is rejected and re-measured ONCE; both attempts are reported.
    first_attempt = None
    st0 = step_stats["resident"]
    flag = torch.tensor([1.0 if st0["max"] > 1.5 * st0["median"] else 0.0], device=dev)
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MAX)
    if float(flag.item()) > 0:
        first_attempt = {"ms_per_step": ms_step, "step_ms": dict(st0)}
        _C.stage_timing_read(True)
        launches0 = _C.kernel_launches()
        sampler.begin()
        ms_step, last = timed(step_resident, args.steps, 0, tag="resident")
        sampler.end()
    clocks = sampler.stop() if rank == 0 else None
    launches = _C.kernel_launches() - launches0
    stages = _C.stage_timing_read(True)
    _C.stage_timing_enable(False)
    R_local = last.num_rendered
    loss_val = float(last.loss.item())
    # per-rank load (views, instances, device ms spent in this library's stages per step)
    my_ms = sum(ms for _n, ms, c in stages if c > 0) / args.steps
    per_rank = torch.tensor([float(band_share if band_share is not None else Vl), float(R_local), my_ms], device=dev,
                            dtype=torch.float64)
    if world > 1:
        allr = [torch.zeros_like(per_rank) for _ in range(world)]
        dist.all_gather(allr, per_rank)
    else:
        allr = [per_rank]
    per_rank = [{"views": round(float(t[0]), 2), "R": int(t[1]), "stage_ms": round(float(t[2]), 3)} for t in allr]

    # ---- (1b) the same step without the geometry maps nobody reads (extra; NOT the headline) ----
    # The refine loop's loss reads the colour map only (infer_batch.py:300-305); the reference still renders
    # coord / depth / normal because its rasterizer is called with require_coord = require_depth = True.  Inside the
    # fused native step those maps never leave the call, so eliding them changes no result (loss, gradients and radii
    # are bit-identical: tests/test_refine_gpu.py); `value` keeps them, this key shows the step without them.
    ms_lean = None
    if Vl:
        st_lean = st._replace(require_coord=False, require_depth=False)

        def step_lean():
            state.flat.copy_(snapshot)
            return refine.refine_iteration(state, st_lean, gt_local, V, group=group, loss_impl=loss_impl)
        ms_lean, _ = timed(step_lean, args.steps, 3)

    # ---- (2) forward-only throughput ----
    def fwd_only():
        with torch.no_grad():
            return refine.forward_views(state, st) if Vl else None
    ms_fwd, _ = timed(fwd_only, max(3, args.steps // 2), 2)
    fwd_mpix = V * N / 1e6 / (ms_fwd / 1e3)

    # ---- (2a) work counters (outside any timed region): E = evaluated (pixel, instance) pairs = sum of n_contrib[0]
    # (SURVEY 8d), R = instances; and, for N > 1, that every rank holds bit-identical parameters after the steps ----
    work = None
    if Vl:
        with torch.no_grad():
            sc_a, rot_a, op_a = _C.activate_forward(state.param("scaling"), state.param("rotation"), state.param("opacity"))
            e_ = torch.empty(0, device=dev)
            o = _C.rasterize_views(st.bg, state.param("xyz"), e_, op_a, sc_a, rot_a, 1.0, e_, st.viewmatrices, st.projmatrices,
                                   st.tanfovx, st.tanfovy, 0.0, H, W, state.param("shs"), 3, st.campos, False, True, True, False,
                                   st.tile_rows)
            dec = _C.decode_buffers(P, Vl, o[0], H, W, o[9], o[10], o[11])
            E_local = dec[3].view(Vl, 2, H * W)[:, 0].to(torch.int64).sum()
            R_cnt = torch.tensor(int(o[0]), device=dev, dtype=torch.int64)
            pair_counts = _C.count_blend_pairs(P, Vl, int(o[0]), H, W, o[9], o[10], o[11])   # hits, E', C, staged
            del o, dec
    else:
        E_local = torch.zeros((), device=dev, dtype=torch.int64)
        R_cnt = torch.zeros((), device=dev, dtype=torch.int64)
        pair_counts = torch.zeros(4, device=dev, dtype=torch.int64)
    R_true_local = int(R_cnt)   # in capacity mode `num_rendered` is the capacity, not the count
    local_counts = [int(x) for x in pair_counts.tolist()]   # this rank's: the roofline below is rank 0's kernel
    E_rank0 = int(E_local)
    cs = state.flat.view(torch.int32).to(torch.int64).sum()
    identical = None
    if world > 1:
        pack = torch.cat([torch.stack([E_local, R_cnt]), pair_counts])
        dist.all_reduce(pack)
        E_local, R_cnt, pair_counts = pack[0], pack[1], pack[2:]
        css = [torch.zeros_like(cs) for _ in range(world)]
        dist.all_gather(css, cs)
        identical = bool(all(int(c) == int(css[0]) for c in css))
    state.poll_status(wait=True)
    work = {"instances_R": int(R_cnt), "evaluated_pairs_E": int(E_local),
            "block_hits": int(pair_counts[0]), "exp_evaluated_pairs_Eprime": int(pair_counts[1]),
            "contributing_pairs_C": int(pair_counts[2]), "params_identical_across_ranks": identical,
            "binning": {"mode": "exact (host read of R per forward)" if args.exact_binning else
                        "capacity (no host round-trip; first step exact)", "overflowed_steps_rank0": state.overflows}}

    # ---- (2b) the exchange step alone (packed-gradient all-reduce), for the scaling analysis ----
    allreduce_ms = None
    if world > 1:
        zero = torch.zeros((), device=dev)

        def only_reduce():
            if state.fused is not None:
                return state.fused_exchange_step(zero, False)    # reduce + Adam + broadcast in one kernel (+ 2 barriers)
            return refine.reduce_gradients(state.grad, zero, group)
        allreduce_ms, _ = timed(only_reduce, 10, 3)

    # ---- (3) end to end: host buffers in, loss out ----
    e2e = None
    if not args.no_e2e:
        ms_e2e, _ = timed(step_e2e, args.steps, max(args.warmup, 3))
        e2e_state["blocking"] = True
        ms_e2e_blocking, _ = timed(step_e2e, args.steps, 2)
        e2e_state["blocking"] = False
        h2d = cam_host.numel() * 4 + (gt_host[:Vl].numel() * 4 if Vl else 0)
        e2e = {"value": ms_e2e, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4,
               "loss_readback": "every step: asynchronous 4-byte copy into pinned memory, consumed by the host one step later",
               "value_with_blocking_readback": ms_e2e_blocking}

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant stage ----
    peaks_file = ROOT / "MEASURED_PEAKS.json"
    if peaks_file.exists():
        peak = float(json.loads(peaks_file.read_text())["hbm_gbs"])
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    stage_ms = {n: (ms, c) for n, ms, c in stages if c > 0}
    VP = Vl * P
    Vpix = band_share if band_share is not None else Vl   # views' worth of pixels this rank blends
    R = R_true_local
    # tile-id sort: 8-bit digits, or 9-bit when that saves a pass (binning.cu radix_sort_pairs)
    tile_bits = max(1, (max(1, Vl) * ((W + 15) // 16) * ((H + 15) // 16) - 1).bit_length())
    sort_passes = min((tile_bits + 7) // 8, (tile_bits + 8) // 9)
    # algorithmic bytes per launch of each stage for THIS design
    alg = {
        "gauss_setup": P * (28 + 24 + 96),
        "project": VP * (12 + 4 + 192 + 24 + 96 + 32 + 64 + 4 + 4 + 4 + 1 + 12 + 4),
        "depth_sort": 5 * VP * (12 + 12) + 5 * VP * 8,
        "scan_rank": VP * (4 + 4 + 4 + 4) + VP * 8,
        "emit": VP * (4 + 8 + 32 + 4) + R * 12,
        "tile_sort": sort_passes * R * (8 + 8) + sort_passes * R * 4,
        "finalize_bins": R * (4 + 4 + 4 + 4),
        "blend_fwd": R * (4 + 32 + 64) + Vpix * N * 88,
        "blend_bwd": R * (4 + 4 + 32 + 4) + Vpix * N * (12 + 4 + 8) + R * 64,
        "gather_bwd": R * 64 + VP * (32 + 4 + 4 + 8 + 1 + 12 + 64),
        "finish_bwd": VP * 64 + P * (12 + 192 + 12 + 16 + 12 + 4 + 24 + 192 + 12 + 16),
        "adam": P * 59 * 4 * 7,
        "loss": Vpix * N * 3 * 4 * 3,
        "activate": P * 80,   # mean of the forward (64 B/Gaussian) and backward (96 B/Gaussian) launches
    }
    top = max(stage_ms.items(), key=lambda kv: kv[1][0])[0] if stage_ms else None
    roofline = None
    if top:
        tot_ms, cnt = stage_ms[top]
        per_launch_ms = tot_ms / cnt
        t_s = per_launch_ms / 1e3
        achieved = alg[top] / 1e9 / t_s
        tj = {}
        tf = ROOT / "profiles" / "traffic.json"   # per-launch numbers of the committed `ncu --set full` captures
        if tf.exists():
            tj = json.loads(tf.read_text())
        traffic = tj.get(top)
        roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": alg[top], "ms_per_launch": per_launch_ms}
        # The blend kernels are not HBM-shaped work: SURVEY 8(d) asks for their FP32, MUFU and shared-memory rates against
        # MEASURED pipe peaks (tools/micro/pipe_rates.cu -> profiles/r02_pipe_rates.json) next to the HBM fraction.  The work
        # is counted LIVE on this rank's views (igs_count_blend_pairs: E' = pairs whose exponent is evaluated, C = pairs
        # blended; E = sum of n_contrib); the per-pair constants are SURVEY 8(d)'s; `bound` names the largest fraction.
        pr_file = ROOT / "profiles" / "r02_pipe_rates.json"
        if top in ("blend_fwd", "blend_bwd") and pr_file.exists() and clocks and clocks.get("sm_mhz"):
            pr = json.loads(pr_file.read_text())
            sms = torch.cuda.get_device_properties(dev).multi_processor_count
            clk = clocks["sm_mhz"] * 1e6
            hits, Ep, C, _staged = local_counts
            E = E_rank0
            bwd = top == "blend_bwd"
            flop = 10.0 * E + (150.0 if bwd else 45.0) * C
            mufu = float(Ep + (C if bwd else 0))                      # ex2 per evaluated pair (+ one reciprocal per blended pair)
            wpp = tj.get("_shared_wavefronts_per_block_hit", {}).get(top)
            fr = {"hbm": achieved / peak}
            peak_flop = pr["ffma_warp_inst_per_clk_per_sm"] * 64.0 * sms * clk
            fr["fp32"] = flop / t_s / peak_flop
            peak_mufu = pr["mufu_ex2_warp_inst_per_clk_per_sm"] * 32.0 * sms * clk
            fr["mufu"] = mufu / t_s / peak_mufu
            detail = {"fp32": {"achieved_TFLOPs": flop / t_s / 1e12, "peak_TFLOPs": peak_flop / 1e12,
                               "model": f"10*E + {150 if bwd else 45}*C flop (SURVEY 8d)"},
                      "mufu": {"achieved_Gop_s": mufu / t_s / 1e9, "peak_Gop_s": peak_mufu / 1e9,
                               "model": "E' ex2" + (" + C rcp" if bwd else "")}}
            if wpp:
                peak_wf = pr.get("smem_wavefronts_per_clk_per_sm", 1.0) * sms * clk
                fr["smem"] = hits * wpp / t_s / peak_wf
                detail["smem"] = {"achieved_Gwavefronts_s": hits * wpp / t_s / 1e9, "peak_Gwavefronts_s": peak_wf / 1e9,
                                  "model": f"{wpp} shared wavefronts per (8x4 block, instance) hit (committed ncu capture, "
                                           f"profiles/traffic.json) x live hits"}
            ipp = tj.get("_warp_instructions_per_block_hit", {}).get(top)
            if ipp:   # a utilisation, not a work bound: issue slots used per second
                detail["issue"] = {"frac": hits * ipp / t_s / (4.0 * sms * clk),
                                   "model": f"{ipp} warp instructions per hit (committed ncu capture) x live hits / 4 per clk per SM"}
            roofline["fractions"] = {k: round(v, 4) for k, v in fr.items()}
            roofline["bound"] = max(fr, key=fr.get)
            roofline["pipes"] = detail
            roofline["counts"] = {"E": E, "E_prime": Ep, "C": C, "block_hits": hits}
            roofline["peaks_source"] = "profiles/r02_pipe_rates.json (measured on this pool's B200) at the SM clock sampled live"
            roofline["note"] = ("achieved/peak/frac/unit keep the contract's HBM form; the kernel's binding resource is "
                                "`bound`, its fraction fractions[bound]")
    stage_report = {n: {"ms_per_step": ms / args.steps, "launches_per_step": c / args.steps,
                        "alg_GBps": (alg[n] * (c / args.steps)) / 1e9 / (ms / args.steps / 1e3) if ms > 0 else None}
                    for n, (ms, c) in stage_ms.items()}

    # ---- CPU baseline (bounded sample) ----
    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            nv = max(2, args.cpu_sample_views)
            ids = cpu_sample_views(V, 0, nv)
            ms_view, adam_ms, threads = cpu_oracle_ms(cfg, ids)
            cpu_baseline = {"value": ms_view * V + adam_ms, "unit": UNIT, "cores": threads, "kind": "port",
                            "sample": f"views {ids} of {V}: CPU oracle (oracle/raster_oracle.c) fwd + torch-CPU L1+SSIM + "
                                      f"oracle bwd, scaled x{V}/{nv}, + one torch-CPU Adam step"}
        except Exception as ex:  # the oracle is a checker; its absence must not hide the GPU number
            cpu_baseline = {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": f"unavailable: {ex}"}

    # ---- reference CUDA rasterizer on the same views (extra evidence, not the `reference` arm) ----
    ref_cuda = None
    if world == 1 and not args.no_ref_cuda:
        try:
            ref_cuda = time_reference_cuda(cfg, cams, sc, gt_dev, dev, loss_impl="torch", steps=args.steps,
                                           warmup=max(args.warmup, 3))
        except Exception as ex:
            ref_cuda = {"unavailable": str(ex)[:200]}

    line = {
        "metric": METRIC, "value": ms_step, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(args, cfg), "e2e": e2e, "gpu_launches": int(launches),
        "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline,
        "value_without_unused_geometry_maps": ms_lean, "fwd_mpix_s": fwd_mpix, "fwd_ms": ms_fwd, "loss": loss_val, "loss_impl": loss_impl,
        "num_rendered_rank0": int(R_true_local), "binning_capacity_rank0": int(R_local), "allreduce_ms": allreduce_ms,
        "allreduce_bytes": int(state.grad.numel() * 4), "exchange": exchange,
        "allreduce_ms_note": "fused: the whole exchange step (reduce + Adam + broadcast + 2 cross-rank barriers); nccl: the all-reduce alone", "per_rank": per_rank, "view_assignment": assignment, "step_ms": step_stats.get("resident"), "first_attempt": first_attempt, "work": work, "stages": stage_report, "ref_cuda": ref_cuda,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def measure_view_costs(cfg, cams, sc, dev):
    """Instances R_v of each view (forward only, one view at a time) — the cost proxy for view balancing."""
    import torch
    from igs_b200 import scenes
    import igs_b200
    scales, rot, opac = scenes.activated(sc)
    t = [x.to(dev) for x in (sc.xyz, opac, scales, rot, sc.sh)]
    bg = torch.zeros(3, device=dev)
    e = torch.Tensor([])
    costs = []
    for v in range(cfg["V"]):
        out = igs_b200._C.rasterize_gaussians(bg, t[0], e, t[1], t[2], t[3], 1.0, e, cams.viewmatrix[v].to(dev).contiguous(),
                                              cams.projmatrix[v].to(dev).contiguous(), cams.tanfovx, cams.tanfovy, 0.0,
                                              cfg["H"], cfg["W"], t[4], 3, cams.campos[v].to(dev).contiguous(), False, False,
                                              False, False)
        costs.append(int(out[0]))
    return costs


def time_reference_cuda(cfg, cams, sc, gt_dev, dev, loss_impl="torch", steps=10, warmup=3):
    """The unmodified reference rasterizer (baseline/_ref) on the same V views: per view
    forward -> same loss -> backward, gradients accumulated over the views, one torch Adam step
    (our schedule; the reference's own schedule is one Adam step per view)."""
    import torch
    import torch.nn.functional as F
    sys.path.insert(0, str(ROOT / "tests"))
    import ref_loader
    from igs_b200 import refine
    ref = ref_loader.load_reference()
    if ref is None:
        return {"unavailable": "baseline/_ref not built on this box"}
    P, W, H, V = cfg["P"], cfg["W"], cfg["H"], cfg["V"]
    params = [sc.xyz.to(dev).requires_grad_(True), sc.sh.to(dev).requires_grad_(True),
              sc.log_scale.to(dev).requires_grad_(True), sc.rot.to(dev).requires_grad_(True),
              sc.opacity_logit.to(dev).requires_grad_(True)]
    lrs = refine.DEFAULT_LRS
    opt = torch.optim.Adam([{"params": [params[0]], "lr": lrs["xyz"]}, {"params": [params[1]], "lr": lrs["shs"]},
                            {"params": [params[2]], "lr": lrs["scaling"]}, {"params": [params[3]], "lr": lrs["rotation"]},
                            {"params": [params[4]], "lr": lrs["opacity"]}], lr=0.0, eps=1e-15)
    bg = torch.zeros(3, device=dev)
    vms = cams.viewmatrix.to(dev)
    pms = cams.projmatrix.to(dev)
    cps = cams.campos.to(dev)

    init = [q.detach().clone() for q in params]

    def one_step(fwd_only=False, per_view_adam=False):
        tot = 0.0
        with torch.no_grad():
            for q, q0 in zip(params, init):
                q.copy_(q0)   # same stationary workload as our arm
        for v in range(V):
            xyz, sh, ls, rot, ol = params
            stt = ref.GaussianRasterizationSettings(
                image_height=H, image_width=W, tanfovx=cams.tanfovx, tanfovy=cams.tanfovy, kernel_size=0.0, bg=bg,
                scale_modifier=1.0, viewmatrix=vms[v], projmatrix=pms[v], sh_degree=3, campos=cps[v], prefiltered=False,
                require_depth=True, require_coord=True, debug=False)
            m2 = torch.zeros_like(xyz, requires_grad=True)
            outs = ref.GaussianRasterizer(raster_settings=stt)(
                means3D=xyz, means2D=m2, opacities=torch.sigmoid(ol), shs=sh, scales=torch.exp(ls), rotations=F.normalize(rot))
            if fwd_only:
                continue
            loss = refine.refine_loss_torch(outs[0].unsqueeze(0), gt_dev[v:v + 1]).sum() / (1 if per_view_adam else V)
            loss.backward()
            tot = loss
            if per_view_adam:   # the reference's own schedule (infer_batch.py:279-324): one Adam step per view
                opt.step()
                opt.zero_grad(set_to_none=True)
        if not fwd_only and not per_view_adam:
            opt.step()
            opt.zero_grad(set_to_none=True)
        return tot

    # Same protocol as this library's arm: `warmup` untimed steps, `steps` timed steps with one event per step, and a
    # measurement whose slowest step exceeds 1.5x its median is repeated once (both attempts reported).
    def measure(fn, n):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
        ev[0].record()
        for i in range(n):
            fn()
            ev[i + 1].record()
        torch.cuda.synchronize()
        per = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(n))
        return ev[0].elapsed_time(ev[n]) / n, {"min": per[0], "median": per[n // 2], "max": per[-1]}

    def measure_with_outlier_rule(fn, n):
        ms, st = measure(fn, n)
        first = None
        if st["max"] > 1.5 * st["median"]:
            first = {"ms_per_step": ms, "step_ms": st}
            ms, st = measure(fn, n)
        return ms, st, first

    ms, st_ms, first = measure_with_outlier_rule(one_step, steps)
    with torch.no_grad():
        ms_f, _, _ = measure_with_outlier_rule(lambda: one_step(fwd_only=True), max(3, steps // 2))
    ms_1v, _, _ = measure_with_outlier_rule(lambda: one_step(per_view_adam=True), max(3, steps // 2))
    return {"ms_per_step": ms, "step_ms": st_ms, "first_attempt": first, "fwd_ms": ms_f,
            "fwd_mpix_s": V * W * H / 1e6 / (ms_f / 1e3), "steps": steps, "warmup": warmup,
            "schedule": "V views accumulated, one torch Adam step; torch L1+SSIM loss; reference rasterizer sm_100a build",
            "ms_per_V_single_view_iterations": ms_1v,
            "schedule_single_view": "the reference's own loop: V sequential 1-view iterations, each with its own "
                                    "torch Adam step (infer_batch.py:279-324)"}


if __name__ == "__main__":
    sys.exit(main())
