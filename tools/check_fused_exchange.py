#!/usr/bin/env python
"""N-rank check of the fused exchange kernel (csrc/exchange.cu) against the NCCL path it replaces.  Run on N GPUs of one box:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29531 \
      tools/check_fused_exchange.py [--P 300000] [--iters 4]

Both paths start from the same seeded state and run the same view-sharded refine iterations (igs_b200.refine):
  nccl   ncclAllReduce(packed gradients) + all-reduce(loss) + replicated Adam launch
  fused  one kernel per rank: multimem.ld_reduce of its slice, Adam, multimem.st of the updated slice
and must agree (losses; parameters to rounding — the switch sums in a different order than NCCL's ring), with parameters
BIT-IDENTICAL across ranks on both paths.  Prints one JSON line (rank 0) with the comparison and the exchange-step times.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--P", type=int, default=300_000)
    ap.add_argument("--iters", type=int, default=4)
    ap.add_argument("--views", type=int, default=0)
    ap.add_argument("--W", type=int, default=676)
    ap.add_argument("--H", type=int, default=507)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from igs_b200 import refine, scenes
    from igs_b200.multiview import MultiViewSettings

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    group = dist.group.WORLD
    V = args.views or 2 * world
    W, H, P = args.W, args.H, args.P
    sc = scenes.make_gaussians(P, 0)
    cams = scenes.make_cameras(V, W, H, 1)
    mine = refine.shard_views(V, rank, world)
    st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=dev), 1.0,
                           cams.viewmatrix[mine].to(dev).contiguous(), cams.projmatrix[mine].to(dev).contiguous(), 3,
                           cams.campos[mine].to(dev).contiguous())
    gt = torch.rand(V, 3, H, W, generator=torch.Generator().manual_seed(2))[mine].to(dev)

    def fresh():
        return refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, dev)

    def run(state, iters):
        return [float(refine.refine_iteration(state, st, gt, V, group=group).loss) for _ in range(iters)]

    a = fresh()
    la = run(a, args.iters)
    b = fresh()
    ok = b.enable_fused_exchange(group)
    out = {"world": world, "P": P, "views": V, "fused_available": bool(ok), "fused_error": getattr(b, "fused_error", None)}
    if ok:
        lb = run(b, args.iters)
        init = fresh().flat
        moved = float((a.flat - init).abs().mean())
        diff = float((a.flat - b.flat).abs().mean())
        cs = torch.stack([a.flat.view(torch.int32).to(torch.int64).sum(), b.flat.view(torch.int32).to(torch.int64).sum()])
        all_cs = [torch.zeros_like(cs) for _ in range(world)]
        dist.all_gather(all_cs, cs)
        out.update(losses_nccl=la, losses_fused=lb, max_rel_loss_diff=max(abs(x - y) / abs(x) for x, y in zip(la, lb)),
                   mean_abs_param_diff=diff, mean_abs_param_move=moved,
                   identical_across_ranks_nccl=bool(all(int(c[0]) == int(all_cs[0][0]) for c in all_cs)),
                   identical_across_ranks_fused=bool(all(int(c[1]) == int(all_cs[0][1]) for c in all_cs)),
                   adam_steps=[a.step, b.step], skipped=int(b.fused["skipped"].item()))

        # exchange step alone, both ways (CUDA events, max over ranks)
        def timed(fn, n=20):
            for _ in range(5):
                fn()
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                fn()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t)
        zero = torch.zeros((), device=dev)

        def nccl_step():
            refine.reduce_gradients(a.grad, zero, group)
            a.adam_step()
        out["exchange_ms_nccl_allreduce_plus_adam"] = timed(nccl_step)
        out["exchange_ms_fused_kernel_with_barriers"] = timed(lambda: b.fused_exchange_step(zero, False))
        out["bytes"] = int(a.grad.numel() * 4)
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
