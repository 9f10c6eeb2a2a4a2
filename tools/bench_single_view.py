#!/usr/bin/env python
"""BASELINE.json configs[1] ("C2"): demo-shaped single view — 200k synthetic Gaussians, 512x512, forward + backward
through the drop-in Python API (`diff_gaussian_rasterization_rade.GaussianRasterizer`, torch autograd), this library
against the UNMODIFIED reference CUDA rasterizer (baseline/_ref) on the same inputs.  Prints one JSON line.

  python tools/bench_single_view.py [--workload C2] [--steps 50] [--all-cotangents]
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
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--all-cotangents", action="store_true", help="cotangents on all seven maps (default: colour only)")
    args = ap.parse_args()
    import torch
    import diff_gaussian_rasterization_rade as ours
    import ref_loader
    from igs_b200 import _C, scenes
    dev = torch.device("cuda:0")
    cfg = scenes.CONFIGS[args.workload]
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    sc = scenes.make_gaussians(P, 0, cfg["log_scale_mean"])
    cams = scenes.make_cameras(max(cfg["V"], 1), W, H, 1)
    scales, rot, opac = scenes.activated(sc)
    leaves = {k: v.to(dev).clone().requires_grad_(True) for k, v in
              dict(means3D=sc.xyz, shs=sc.sh, opacities=opac, scales=scales, rotations=rot).items()}
    cots = [c[0].to(dev) for c in scenes.make_cotangents(1, H, W, 3)]
    bg = torch.zeros(3, device=dev)

    def run(mod, steps):
        st = mod.GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=cams.tanfovx, tanfovy=cams.tanfovy, kernel_size=0.0, bg=bg, scale_modifier=1.0,
            viewmatrix=cams.viewmatrix[0].to(dev), projmatrix=cams.projmatrix[0].to(dev), sh_degree=3, campos=cams.campos[0].to(dev),
            prefiltered=False, require_depth=True, require_coord=True, debug=False)
        rast = mod.GaussianRasterizer(raster_settings=st)

        def fwd():
            return rast(means2D=torch.zeros_like(leaves["means3D"], requires_grad=True), **leaves)

        def fwd_bwd():
            for v in leaves.values():
                v.grad = None
            out = fwd()
            maps = [out[0], out[2], out[3], out[4], out[5], out[6], out[7]]
            loss = sum((m * c).sum() for m, c in zip(maps, cots)) if args.all_cotangents else (out[0] * cots[0]).sum()
            loss.backward()
        for _ in range(5):
            fwd_bwd()
        torch.cuda.synchronize()
        sampler = clocks.ClockSampler(0, period_s=0.01)
        sampler.start()
        sampler.begin()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        for _ in range(steps):
            fwd_bwd()
        e[1].record()
        with torch.no_grad():
            for _ in range(steps):
                fwd()
        e[2].record()
        torch.cuda.synchronize()
        sampler.end()
        f = e[1].elapsed_time(e[2]) / steps
        return {"fwd_bwd_ms": e[0].elapsed_time(e[1]) / steps, "fwd_ms": f, "fwd_mpix_s": W * H / 1e6 / (f / 1e3),
                "clocks": clocks.verdict(sampler.stop())}

    sys.path.insert(0, str(ROOT / "tools"))
    import clocks
    l0 = _C.kernel_launches()
    a = run(ours, args.steps)
    a["igs_kernel_launches_per_fwd_bwd_plus_fwd"] = (_C.kernel_launches() - l0) / (args.steps + 5 + args.steps)
    ref = ref_loader.load_reference()
    b = run(ref, args.steps) if ref is not None else {"unavailable": "baseline/_ref not built on this box"}
    print(json.dumps({"metric": "single_view_fwd_bwd_ms", "workload": f"{args.workload}: {P} Gaussians SH3, one {W}x{H} view, "
                      f"drop-in Python API + torch autograd, cotangents on {'all seven maps' if args.all_cotangents else 'the colour map'}",
                      "ours": a, "reference_cuda": b}), flush=True)


if __name__ == "__main__":
    main()
