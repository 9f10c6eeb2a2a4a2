#!/usr/bin/env python
"""BASELINE.json configs[3] ("C4"): MeetingRoom-shaped stream — 150k Gaussians, 13 views 1280x720, 10 key frames x 50
refine iterations, ONE view per iteration (the reference's own schedule, infer_batch.py:279-324), end to end: every
key frame's 13 target images come from pinned HOST memory, the key frame's output view is rendered after the loop.

  python tools/bench_stream.py [--keyframes 10] [--iters 50] [--densify] [--views-per-step 1] [--no-ref]

Prints one JSON line: ours (igs_b200.stream.run_stream) and, when baseline/_ref travelled to the box, the same loop on
the UNMODIFIED reference CUDA rasterizer with the torch loss and torch.optim.Adam a reference user runs.
"""
from __future__ import annotations

import argparse
import json
import random
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C4")
    ap.add_argument("--keyframes", type=int, default=10)
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--views-per-step", type=int, default=1)
    ap.add_argument("--densify", action="store_true")
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--ref-keyframes", type=int, default=2)
    ap.add_argument("--mode", default="async", choices=["exact", "async", "graph"],
                    help="exact: host read of the instance count per forward (reference behaviour); async: capacity mode; "
                         "graph: one CUDA-graph replay per iteration (falls back to async with --densify)")
    args = ap.parse_args()

    import torch
    import torch.nn.functional as F
    from igs_b200 import _C, refine, scenes, stream
    from igs_b200.multiview import MultiViewSettings

    dev = torch.device("cuda:0")
    cfg = scenes.CONFIGS[args.workload]
    P, V, W, H = cfg["P"], cfg["V"], cfg["W"], cfg["H"]
    sc = scenes.make_gaussians(P, 0, cfg["log_scale_mean"])
    cams = scenes.make_cameras(V, W, H, 1)
    st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, torch.zeros(3, device=dev), 1.0,
                           cams.viewmatrix.to(dev).contiguous(), cams.projmatrix.to(dev).contiguous(), 3,
                           cams.campos.to(dev).contiguous())
    # Targets: renders of the seeded scene itself, so the loop is a real fit (the Gaussians start perturbed) and the
    # cost per iteration stays that of the scene instead of collapsing as it does against noise images.
    state0 = refine.RefineState(sc.xyz, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh, dev)
    with torch.no_grad():
        tgt = refine.forward_views(state0, st)[0].clamp(0, 1).cpu()
    frames = [tgt.pin_memory() for _ in range(min(args.keyframes, 2))]
    frames = [frames[k % len(frames)] for k in range(args.keyframes)]
    g = torch.Generator().manual_seed(4)
    sh_noise = 0.2 * torch.randn(P, 16, 3, generator=g)
    xyz_noise = 0.01 * torch.randn(P, 3, generator=g)

    def fresh_state():
        return refine.RefineState(sc.xyz + xyz_noise, sc.log_scale, sc.rot, sc.opacity_logit, sc.sh + sh_noise, dev)

    mode = "async" if (args.densify and args.mode == "graph") else args.mode
    scfg = stream.StreamConfig(refine_iterations=args.iters, use_densify=args.densify, max_num=cfg["P"], radius=10.0,
                               views_per_step=args.views_per_step, mode=mode)
    sys.path.insert(0, str(ROOT / "tools"))
    import clocks
    # warm-up: one short key frame (allocator, lazy module loads)
    stream.run_stream(fresh_state(), st, frames[:1], scfg._replace(refine_iterations=5), seed=1)
    state = fresh_state()
    torch.cuda.synchronize()
    l0 = _C.kernel_launches()
    sampler = clocks.ClockSampler(0, period_s=0.02)
    sampler.start()
    sampler.begin()
    t0 = time.perf_counter()
    res = stream.run_stream(state, st, frames, scfg, seed=0)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    sampler.end()
    clk = clocks.verdict(sampler.stop())
    launches = _C.kernel_launches() - l0
    total_iters = args.iters * args.keyframes
    # where one iteration's time goes: the library's own stage timers over one more key frame (outside the timed run)
    _C.stage_timing_enable(True)
    stream.run_stream(fresh_state(), st, frames[:1], scfg._replace(mode="async" if mode == "graph" else mode), seed=2)
    torch.cuda.synchronize()
    stages = {n: round(ms / args.iters, 4) for n, ms, c in _C.stage_timing_read(True) if c > 0}
    _C.stage_timing_enable(False)
    ours = {"total_s": wall, "stage_ms_per_iteration": stages, "stage_ms_sum": round(sum(stages.values()), 4), "device_s": sum(res.keyframe_ms) / 1e3, "ms_per_iteration": wall * 1e3 / total_iters,
            "ms_per_keyframe": wall * 1e3 / args.keyframes, "psnr_first": res.psnrs[0], "psnr_last": res.psnrs[-1],
            "points_last": res.points_num[-1], "gpu_launches": int(launches),
            "h2d_bytes_per_keyframe": int(frames[0].numel() * 4), "mode": mode, "clocks": clk}

    ref_line = None
    if not args.no_ref:
        import ref_loader
        ref = ref_loader.load_reference()
        if ref is None:
            ref_line = {"unavailable": "baseline/_ref not built on this box"}
        else:
            p0 = dict(xyz=sc.xyz + xyz_noise, shs=sc.sh + sh_noise, scaling=sc.log_scale, rotation=sc.rot,
                      opacity=sc.opacity_logit)
            params = {k: v.to(dev).clone().requires_grad_(True) for k, v in p0.items()}
            lrs = refine.DEFAULT_LRS
            bg = torch.zeros(3, device=dev)
            rng = random.Random(0)

            def keyframe(host_imgs, iters):
                opt = torch.optim.Adam([{"params": [params[k]], "lr": lrs[k]} for k in params], lr=0.0, eps=1e-15)
                for (v,) in stream.view_schedule(V, iters, rng):
                    gt = host_imgs[v].to(dev)            # the reference uploads the picked image every iteration
                    rs = ref.GaussianRasterizationSettings(
                        image_height=H, image_width=W, tanfovx=cams.tanfovx, tanfovy=cams.tanfovy, kernel_size=0.0, bg=bg,
                        scale_modifier=1.0, viewmatrix=st.viewmatrices[v], projmatrix=st.projmatrices[v], sh_degree=3,
                        campos=st.campos[v], prefiltered=False, require_depth=True, require_coord=True, debug=False)
                    m2 = torch.zeros_like(params["xyz"], requires_grad=True)
                    out = ref.GaussianRasterizer(raster_settings=rs)(
                        means3D=params["xyz"], means2D=m2, opacities=torch.sigmoid(params["opacity"]), shs=params["shs"],
                        scales=torch.exp(params["scaling"]), rotations=F.normalize(params["rotation"]))
                    loss = refine.refine_loss_torch(out[0].unsqueeze(0), gt.unsqueeze(0)).sum()
                    loss.backward()
                    opt.step()
                    opt.zero_grad(set_to_none=True)
                with torch.no_grad():
                    rs = ref.GaussianRasterizationSettings(
                        image_height=H, image_width=W, tanfovx=cams.tanfovx, tanfovy=cams.tanfovy, kernel_size=0.0, bg=bg,
                        scale_modifier=1.0, viewmatrix=st.viewmatrices[0], projmatrix=st.projmatrices[0], sh_degree=3,
                        campos=st.campos[0], prefiltered=False, require_depth=True, require_coord=True, debug=False)
                    img = ref.GaussianRasterizer(raster_settings=rs)(
                        means3D=params["xyz"], means2D=torch.zeros_like(params["xyz"]),
                        opacities=torch.sigmoid(params["opacity"]), shs=params["shs"], scales=torch.exp(params["scaling"]),
                        rotations=F.normalize(params["rotation"]))[0]
                    return float(stream.psnr(img, host_imgs[0].to(dev)))

            keyframe(frames[0], 5)
            with torch.no_grad():
                for k, v in p0.items():
                    params[k].copy_(v.to(dev))
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ps = [keyframe(frames[k], args.iters) for k in range(args.ref_keyframes)]
            torch.cuda.synchronize()
            rw = time.perf_counter() - t0
            ref_line = {"keyframes_timed": args.ref_keyframes, "ms_per_iteration": rw * 1e3 / (args.iters * args.ref_keyframes),
                        "ms_per_keyframe": rw * 1e3 / args.ref_keyframes, "total_s_extrapolated": rw / args.ref_keyframes * args.keyframes,
                        "psnr_first": ps[0], "psnr_last": ps[-1],
                        "what": "unmodified reference CUDA rasterizer (sm_100a build) + torch L1/SSIM + torch Adam, densify off"}

    print(json.dumps({"metric": "stream_total_s", "workload": f"{args.workload}: {P} Gaussians, {V} views {W}x{H}, "
                      f"{args.keyframes} key frames x {args.iters} iterations, {args.views_per_step} view(s)/step, "
                      f"densify {'on' if args.densify else 'off'}", "ours": ours, "reference_cuda": ref_line,
                      "data": "synthetic (targets = renders of the seeded scene; Gaussians start perturbed)"}), flush=True)


if __name__ == "__main__":
    main()
