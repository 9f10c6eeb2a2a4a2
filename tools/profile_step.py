"""Where does a refine step's wall time go?  torch.profiler over a few steps: CUDA kernel totals, runtime API
calls (cudaMalloc/cudaFree/cudaStreamSynchronize) and the step's CUDA-event time.
  python tools/profile_step.py --views 3 --steps 5
"""
import argparse
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from igs_b200 import scenes, refine  # noqa: E402
from igs_b200.multiview import MultiViewSettings  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--views", type=int, default=3)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--workload", default="C3")
    args = ap.parse_args()
    cfg = scenes.CONFIGS[args.workload]
    P, W, H, V = cfg["P"], cfg["W"], cfg["H"], args.views
    dev = torch.device("cuda:0")
    sc = scenes.make_gaussians(P, 0, cfg["log_scale_mean"])
    cams = scenes.make_cameras(cfg["V"], W, H, 1)
    st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=dev), 1.0,
                           cams.viewmatrix[:V].to(dev).contiguous(), cams.projmatrix[:V].to(dev).contiguous(), 3,
                           cams.campos[:V].to(dev).contiguous())
    state = refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, dev)
    gt = torch.rand(V, 3, H, W, device=dev)
    for _ in range(5):
        refine.refine_iteration(state, st, gt, V)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        refine.refine_iteration(state, st, gt, V)
    e1.record()
    torch.cuda.synchronize()
    print(f"step time (events): {e0.elapsed_time(e1) / args.steps:.3f} ms")
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for _ in range(args.steps):
            refine.refine_iteration(state, st, gt, V)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=30, max_name_column_width=60))
    print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=25, max_name_column_width=60))


if __name__ == "__main__":
    main()
