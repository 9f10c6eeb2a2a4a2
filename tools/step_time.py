"""Step time of a V-view refine iteration on this process's GPU (no collectives). CUDA_VISIBLE_DEVICES picks the GPU."""
import argparse, os, sys, time
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from igs_b200 import scenes, refine
from igs_b200.multiview import MultiViewSettings
ap = argparse.ArgumentParser(); ap.add_argument("--views", type=int, default=3); ap.add_argument("--steps", type=int, default=20)
a = ap.parse_args()
cfg = scenes.CONFIGS["C3"]; P, W, H, V = cfg["P"], cfg["W"], cfg["H"], a.views
dev = torch.device("cuda:0")
sc = scenes.make_gaussians(P, 0, cfg["log_scale_mean"]); cams = scenes.make_cameras(cfg["V"], W, H, 1)
st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=dev), 1.0, cams.viewmatrix[:V].to(dev).contiguous(),
                       cams.projmatrix[:V].to(dev).contiguous(), 3, cams.campos[:V].to(dev).contiguous())
state = refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, dev)
gt = torch.rand(V, 3, H, W, device=dev)
for _ in range(5): refine.refine_iteration(state, st, gt, V)
torch.cuda.synchronize()
# crude cross-process alignment: start at the next multiple of 5 s
time.sleep(5 - time.time() % 5)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
for _ in range(a.steps): refine.refine_iteration(state, st, gt, V)
e1.record(); torch.cuda.synchronize(); t1 = time.perf_counter()
print(f"gpu {os.environ.get('CUDA_VISIBLE_DEVICES')} cpus {os.cpu_count()} affinity {len(os.sched_getaffinity(0))} step {e0.elapsed_time(e1)/a.steps:.3f} ms wall {(t1-t0)*1e3/a.steps:.3f} ms", flush=True)
