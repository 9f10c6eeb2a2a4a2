"""Under torchrun: profile a few sharded refine steps on rank 0 and list the GPU idle gaps (what ran before / after)."""
import os, sys
from pathlib import Path
import torch, torch.distributed as dist
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from igs_b200 import scenes, refine
from igs_b200.multiview import MultiViewSettings
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
group = None
if world > 1:
    dist.init_process_group("nccl", device_id=dev); group = dist.group.WORLD
cfg = scenes.CONFIGS["C3"]; P, W, H, V = cfg["P"], cfg["W"], cfg["H"], cfg["V"]
views = refine.shard_views(V, rank, world) if world > 1 else [0, 1, 2]
Vl = len(views)
sc = scenes.make_gaussians(P, 0, cfg["log_scale_mean"]); cams = scenes.make_cameras(V, W, H, 1)
st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=dev), 1.0, cams.viewmatrix[views].to(dev).contiguous(),
                       cams.projmatrix[views].to(dev).contiguous(), 3, cams.campos[views].to(dev).contiguous())
state = refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, dev)
gt = torch.rand(Vl, 3, H, W, device=dev)
for _ in range(5): refine.refine_iteration(state, st, gt, V, group=group)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(4): refine.refine_iteration(state, st, gt, V, group=group)
    torch.cuda.synchronize()
if rank == 0:
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    ev.sort(key=lambda e: e.time_range.start)
    t0 = ev[0].time_range.start
    print(f"{len(ev)} device events over {(ev[-1].time_range.end - t0)/1e3:.2f} ms (4 steps)")
    busy_end = ev[0].time_range.end; prev = ev[0]
    gaps = []
    for e in ev[1:]:
        if e.time_range.start > busy_end:
            gaps.append((e.time_range.start - busy_end, (busy_end - t0), prev.name[:60], e.name[:60]))
        if e.time_range.end > busy_end:
            busy_end = e.time_range.end; prev = e
    print("total idle %.3f ms" % (sum(g[0] for g in gaps) / 1e3))
    for g in sorted(gaps, key=lambda g: -g[0])[:25]:
        print(f"gap {g[0]:8.1f} us at t={g[1]/1e3:7.3f} ms  after [{g[2]}] before [{g[3]}]")
    big = [e for e in ev if (e.time_range.end - e.time_range.start) > 150]
    print("--- long device events")
    for e in big[:60]:
        print(f"t={(e.time_range.start - t0)/1e3:7.3f} dur {(e.time_range.end - e.time_range.start)/1e3:6.3f} ms {e.name[:70]}")
if world > 1:
    dist.barrier(); dist.destroy_process_group()
