#!/usr/bin/env python
"""RaDe-GS frame-0 training iteration (RADE/train.py:109-161) on one view: render with
kernel_size 0.1 and the 3D filter, L1 + D-SSIM, depth-normal regulariser (all seven cotangents live), backward.
Times this library (fused loss kernels) and, when baseline/_ref travelled, the UNMODIFIED reference rasterizer with the
reference's torch formulation of the losses.  Prints one JSON line.

  python tools/bench_frame0.py [--P 300000] [--W 1352] [--H 1014] [--steps 20]
"""
import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--P", type=int, default=300_000)
    ap.add_argument("--W", type=int, default=1352)
    ap.add_argument("--H", type=int, default=1014)
    ap.add_argument("--views", type=int, default=20)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--coord", action="store_true")
    args = ap.parse_args()
    import torch
    from igs_b200 import _C, frame0, scenes
    import ref_loader
    dev = torch.device("cuda:0")
    P, W, H, V = args.P, args.W, args.H, args.views
    sc = scenes.make_gaussians(P, 0)
    cams = scenes.make_cameras(V, W, H, 1)
    params = dict(xyz=sc.xyz, shs=sc.sh, scaling=sc.log_scale, rotation=sc.rot, opacity=sc.opacity_logit)
    params = {k: v.to(dev).clone().requires_grad_(True) for k, v in params.items()}
    f3d = frame0.compute_3D_filter(params["xyz"].detach(), cams.viewmatrix.to(dev), W, H, cams.tanfovx, cams.tanfovy)
    camlist = [dict(viewmatrix=cams.viewmatrix[v].to(dev), projmatrix=cams.projmatrix[v].to(dev), campos=cams.campos[v].to(dev),
                    tanfovx=cams.tanfovx, tanfovy=cams.tanfovy, W=W, H=H) for v in range(V)]
    gt = torch.rand(V, 3, H, W, generator=torch.Generator().manual_seed(2)).to(dev)
    bg = torch.zeros(3, device=dev)
    opt = frame0.Frame0Options(kernel_size=0.1, regularization_from_iter=0, use_coord_map=args.coord)

    def run(module, impl, steps):
        def it(i):
            for p in params.values():
                p.grad = None
            v = i % V
            out = frame0.training_loss(camlist[v], params, f3d, bg, gt[v], 15000, opt, impl=impl, module=module)
            out.loss.backward()
            return out
        for i in range(3):
            it(i)
        torch.cuda.synchronize()
        sampler = clocks.ClockSampler(0, period_s=0.01)
        sampler.start()
        sampler.begin()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            out = it(i)
        e1.record()
        torch.cuda.synchronize()
        sampler.end()
        clock_records.append(clocks.verdict(sampler.stop()))
        return e0.elapsed_time(e1) / steps, float(out.loss), float(out.depth_normal_loss)

    sys.path.insert(0, str(ROOT / "tools"))
    import clocks
    clock_records = []
    l0 = _C.kernel_launches()
    ours = run(None, "fused", args.steps)
    launches = (_C.kernel_launches() - l0) / (args.steps + 3)
    ours_torch_losses = run(None, "torch", args.steps)
    ref = ref_loader.load_reference()
    refr = run(ref, "torch", args.steps) if ref is not None else None
    print(json.dumps({
        "metric": "frame0_iter_ms", "workload": f"{P} Gaussians SH3, one {W}x{H} view per iteration, kernel_size 0.1, 3D filter, "
        f"L1 + D-SSIM + depth-normal regulariser ({'coord' if args.coord else 'depth'} maps), fwd + bwd (no optimizer step)",
        "ours_fused_ms": ours[0], "ours_rasterizer_with_torch_losses_ms": ours_torch_losses[0],
        "reference_cuda_ms": None if refr is None else refr[0], "igs_kernel_launches_per_iteration": launches,
        "loss": {"ours": ours[1], "ours_torch_losses": ours_torch_losses[1], "reference": None if refr is None else refr[1]},
        "depth_normal_loss": {"ours": ours[2], "reference": None if refr is None else refr[2]},
        "clocks": {"ours_fused": clock_records[0], "ours_torch_losses": clock_records[1],
                   "reference": clock_records[2] if len(clock_records) > 2 else None}}), flush=True)


if __name__ == "__main__":
    main()
