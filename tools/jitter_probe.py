"""Per-step device times of the C3 refine iteration over many steps, to find sources of step-time jitter.
Modes: plain | stages (library stage timers on) | nvml (50 ms NVML polling thread) | smi (nvidia-smi -lms 100 child)."""
import argparse, subprocess, sys, threading, time
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from igs_b200 import scenes, refine, _C
from igs_b200.multiview import MultiViewSettings
ap = argparse.ArgumentParser(); ap.add_argument("--steps", type=int, default=100); ap.add_argument("--mode", default="plain")
a = ap.parse_args()
cfg = scenes.CONFIGS["C3"]; P, W, H, V = cfg["P"], cfg["W"], cfg["H"], cfg["V"]
dev = torch.device("cuda:0")
sc = scenes.make_gaussians(P, 0, cfg["log_scale_mean"]); cams = scenes.make_cameras(V, W, H, 1)
st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=dev), 1.0, cams.viewmatrix.to(dev).contiguous(),
                       cams.projmatrix.to(dev).contiguous(), 3, cams.campos.to(dev).contiguous())
state = refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, dev)
snap = state.flat.clone()
gt = torch.rand(V, 3, H, W, device=dev)
def step():
    state.flat.copy_(snap)
    refine.refine_iteration(state, st, gt, V)
for _ in range(5): step()
torch.cuda.synchronize()
stop = threading.Event(); proc = None
if a.mode == "stages":
    _C.stage_timing_enable(True)
elif a.mode == "nvml":
    import pynvml
    pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
    def poll():
        while not stop.is_set():
            pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM); pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            pynvml.nvmlDeviceGetPowerUsage(h); stop.wait(0.05)
    threading.Thread(target=poll, daemon=True).start()
elif a.mode == "smi":
    proc = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader", "-lms", "100"],
                            stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps + 1)]
ev[0].record()
for i in range(a.steps):
    step(); ev[i + 1].record()
torch.cuda.synchronize()
stop.set()
if proc: proc.terminate()
ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(a.steps)]
s = sorted(ms)
print(f"mode {a.mode}: median {s[len(s)//2]:.3f} p90 {s[int(len(s)*0.9)]:.3f} max {s[-1]:.3f} mean {sum(ms)/len(ms):.3f}; outliers(>1.1x median): "
      f"{[(i, round(x,2)) for i, x in enumerate(ms) if x > 1.1 * s[len(s)//2]]}", flush=True)
