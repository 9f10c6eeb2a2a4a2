#!/usr/bin/env python
"""Multi-rank check of the key-frame loop (igs_b200/stream.py) — launch with torchrun on N GPUs:
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/check_stream_dist.py
Every rank runs refine_keyframe with views_per_step = 2N (rank r renders every N-th view of the step, gradients and
densification statistics all-reduced); afterwards all ranks must hold bit-identical parameters, and they must agree
with a single-rank run of the same schedule within the gradient parity bar.  Prints one JSON line on rank 0."""
import json
import os
import random
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from igs_b200 import refine, scenes, stream  # noqa: E402
from igs_b200.multiview import MultiViewSettings  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    P, V, W, H = 30000, 4 * world, 320, 240
    sc = scenes.make_gaussians(P, 0, -1.8)
    cams = scenes.make_cameras(V, W, H, 1)
    st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=dev), 1.0,
                           cams.viewmatrix.to(dev).contiguous(), cams.projmatrix.to(dev).contiguous(), 3, cams.campos.to(dev).contiguous())
    gt = torch.rand(V, 3, H, W, generator=torch.Generator().manual_seed(5)).to(dev)

    def run(group, densify):
        state = refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, dev)
        cfg = stream.StreamConfig(refine_iterations=25, views_per_step=2 * world, use_densify=densify, max_num=P + 2000,
                                  radius=10.0, densify_grad_threshold=1e-7)
        res = stream.refine_keyframe(state, st, gt, cfg, random.Random(3), group=group,
                                     generator=torch.Generator(device=dev).manual_seed(9))
        return state, res

    out = {}
    for densify in (False, True):
        s_multi, r_multi = run(dist.group.WORLD, densify)
        cs = s_multi.flat.view(torch.int32).to(torch.int64).sum()
        all_cs = [torch.zeros_like(cs) for _ in range(world)]
        dist.all_gather(all_cs, cs)
        identical = all(int(c) == int(all_cs[0]) for c in all_cs)
        s_one, r_one = run(None, densify)
        same_P = s_one.P == s_multi.P
        rel = float((s_one.flat - s_multi.flat).abs().max() / s_one.flat.abs().max()) if same_P else None
        out["densify" if densify else "plain"] = {"identical_across_ranks": identical, "P": s_multi.P, "P_single_rank": s_one.P,
                                                  "max_rel_diff_vs_single_rank": rel, "densified_at": r_multi.densified_at,
                                                  "loss_first_last": [float(r_multi.losses[0]), float(r_multi.losses[-1])]}
    if rank == 0:
        print(json.dumps({"world": world, **out}), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
