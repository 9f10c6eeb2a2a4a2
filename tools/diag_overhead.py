import sys, subprocess, time
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from igs_b200 import scenes, refine, _C
from igs_b200.multiview import MultiViewSettings
cfg = scenes.CONFIGS["C3"]; P, W, H, V = cfg["P"], cfg["W"], cfg["H"], cfg["V"]
dev = torch.device("cuda:0")
sc = scenes.make_gaussians(P, 0, cfg["log_scale_mean"]); cams = scenes.make_cameras(V, W, H, 1)
st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=dev), 1.0, cams.viewmatrix.to(dev), cams.projmatrix.to(dev), 3, cams.campos.to(dev))
state = refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, dev)
gt = torch.rand(V, 3, H, W, device=dev)
def run(n=10):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): refine.refine_iteration(state, st, gt, V)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for _ in range(5): refine.refine_iteration(state, st, gt, V)
print("plain            ", run())
_C.stage_timing_enable(True); print("stage timers on  ", run()); _C.stage_timing_read(True); _C.stage_timing_enable(False)
print("plain            ", run())
p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader", "-lms", "100", "-i", "0"], stdout=subprocess.DEVNULL)
time.sleep(0.5); print("nvidia-smi 100ms ", run()); p.terminate(); p.wait()
p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader", "-lms", "500", "-i", "0"], stdout=subprocess.DEVNULL)
time.sleep(0.5); print("nvidia-smi 500ms ", run()); p.terminate(); p.wait()
print("plain            ", run())
