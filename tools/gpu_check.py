"""GPU bring-up diagnostics: this library vs the reference CUDA rasterizer (baseline/_ref) on the same
seeded inputs, stage by stage.  Prints one line per compared quantity; writes gpurun_out/check_<cfg>.json.

  python tools/gpu_check.py --cfg C1 [--views 1] [--time]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from igs_b200 import scenes  # noqa: E402
import igs_b200  # noqa: E402
from igs_b200 import _C  # noqa: E402
import ref_loader  # noqa: E402

OUT = ROOT / "gpurun_out"
OUT.mkdir(exist_ok=True)
report = {}


def cmp_exact(name, a, b):
    a = a.reshape(-1)
    b = b.reshape(-1)
    if a.numel() != b.numel():
        print(f"[EXACT] {name:28s} SIZE MISMATCH {a.numel()} vs {b.numel()}")
        report[name] = {"size_mismatch": [a.numel(), b.numel()]}
        return False
    if a.dtype.is_floating_point:
        neq = (a.view(torch.int32) != b.view(torch.int32))
    else:
        neq = a != b
    n = int(neq.sum())
    print(f"[EXACT] {name:28s} mismatches {n} / {a.numel()}")
    report[name] = {"mismatch": n, "n": a.numel()}
    if n and n < 10_000_000:
        idx = neq.nonzero().reshape(-1)[:5].tolist()
        print("         first idx", idx, "ours", a[idx].tolist(), "ref", b[idx].tolist())
    return n == 0


def cmp_tol(name, a, b, atol=None):
    a = a.reshape(-1).double()
    b = b.reshape(-1).double()
    d = (a - b).abs()
    mx = float(d.max()) if d.numel() else 0.0
    rel = float((a - b).norm() / (b.norm() + 1e-30))
    nan_a, nan_b = int(torch.isnan(a).sum()), int(torch.isnan(b).sum())
    scale = float(b.abs().max()) if b.numel() else 0.0
    print(f"[TOL]   {name:28s} max-abs {mx:.3e}  rel-l2 {rel:.3e}  ref-max {scale:.3e}  nan {nan_a}/{nan_b}")
    report[name] = {"max_abs": mx, "rel_l2": rel, "ref_max": scale, "nan": [nan_a, nan_b]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", default="C1")
    ap.add_argument("--P", type=int, default=0)
    ap.add_argument("--view", type=int, default=0, help="which camera of the config's arc to use")
    ap.add_argument("--time", action="store_true")
    ap.add_argument("--geo", action="store_true", help="full (7-map) cotangents instead of colour only")
    args = ap.parse_args()
    cfg = dict(scenes.CONFIGS[args.cfg])
    if args.P:
        cfg["P"] = args.P
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    dev = torch.device("cuda:0")
    sc = scenes.make_gaussians(P, 0, cfg["log_scale_mean"])
    cams = scenes.make_cameras(cfg["V"], W, H, 1)
    vi = args.view
    scales, rot, opac = scenes.activated(sc)
    means = sc.xyz.to(dev)
    scales = scales.to(dev)
    rot = rot.to(dev)
    opac = opac.to(dev)
    sh = sc.sh.to(dev)
    bg = torch.zeros(3, device=dev)
    vm = cams.viewmatrix[vi].to(dev).contiguous()
    pm = cams.projmatrix[vi].to(dev).contiguous()
    cp = cams.campos[vi].to(dev).contiguous()
    cot = scenes.make_cotangents(1, H, W, 3)
    cot = [c[0].to(dev) for c in cot]

    ref = ref_loader.load_reference()
    print("reference available:", ref is not None, "| P", P, "WxH", W, H, "| device", torch.cuda.get_device_name(0))

    def run(mod_settings, mod_rast, leaf=True):
        m = means.clone().requires_grad_(leaf)
        s_ = scales.clone().requires_grad_(leaf)
        r_ = rot.clone().requires_grad_(leaf)
        o_ = opac.clone().requires_grad_(leaf)
        h_ = sh.clone().requires_grad_(leaf)
        m2 = torch.zeros_like(m, requires_grad=True)
        st = mod_settings(image_height=H, image_width=W, tanfovx=cams.tanfovx, tanfovy=cams.tanfovy, kernel_size=0.0,
                          bg=bg, scale_modifier=1.0, viewmatrix=vm, projmatrix=pm, sh_degree=3, campos=cp,
                          prefiltered=False, require_depth=True, require_coord=True, debug=False)
        outs = mod_rast(raster_settings=st)(means3D=m, means2D=m2, opacities=o_, shs=h_, scales=s_, rotations=r_)
        return (m, m2, h_, o_, s_, r_), outs

    # ---------------- forward ----------------
    ours_in, ours = run(igs_b200.GaussianRasterizationSettings, igs_b200.GaussianRasterizer)
    torch.cuda.synchronize()
    names = ["color", "radii", "coord", "mcoord", "depth", "mdepth", "alpha", "normal"]
    for n, t in zip(names, ours):
        print(f"ours {n:8s} shape {tuple(t.shape)} finite {bool(torch.isfinite(t.float()).all())} "
              f"absmax {float(t.float().abs().max()):.4g}")
    fn = ours[0].grad_fn
    print("grad_fn:", fn)

    # decode our buffers via a direct native call (same inputs)
    e = torch.Tensor([])
    nat = _C.rasterize_gaussians(bg, means, e, opac, scales, rot, 1.0, e, vm, pm, cams.tanfovx, cams.tanfovy, 0.0, H, W,
                                 sh, 3, cp, False, True, True, False)
    R_ours = nat[0]
    dec = _C.decode_buffers(P, 1, R_ours, H, W, nat[9], nat[10], nat[11])
    (keys, plist, ranges, ncontrib, tiles_t, depths, means2D, conic, rgb, cov3D, vpts, cpl, rpl, ts, nrm) = dec
    print("num_rendered ours:", R_ours, "| visible:", int((nat[8] > 0).sum()), "| tiles_touched sum", int(tiles_t.sum()))
    report["R_ours"] = R_ours

    if ref is not None:
        ref_in, refo = run(ref.GaussianRasterizationSettings, ref.GaussianRasterizer)
        torch.cuda.synchronize()
        rnat = ref._C.rasterize_gaussians(bg, means, e, opac, scales, rot, 1.0, e, vm, pm, cams.tanfovx, cams.tanfovy,
                                          0.0, H, W, sh, 3, cp, False, True, True, False)
        R_ref = rnat[0]
        print("num_rendered ref :", R_ref)
        report["R_ref"] = R_ref
        g = ref_loader.decode_ref_geom(rnat[9], P)
        rradii = rnat[8]
        vis = rradii > 0
        cmp_exact("radii", nat[8], rradii)
        cmp_exact("tiles_touched", tiles_t.reshape(-1), g["tiles_touched"])
        cmp_exact("cov3D(all)", cov3D, g["cov3D"]) if True else None
        cmp_exact("depths(vis)", depths.reshape(-1)[vis], g["depths"][vis])
        cmp_exact("means2D(vis)", means2D.reshape(P, 2)[vis], g["means2D"][vis])
        cmp_exact("conic_opacity(vis)", conic.reshape(P, 4)[vis], g["conic_opacity"][vis])
        cmp_tol("rgb(vis)", rgb.reshape(P, 3)[vis], g["rgb"][vis])
        cmp_tol("view_points(vis)", vpts.reshape(P, 3)[vis], g["view_points"][vis])
        cmp_tol("camera_planes(vis)", cpl.reshape(P, 6)[vis], g["camera_planes"][vis])
        cmp_tol("ray_planes(vis)", rpl.reshape(P, 2)[vis], g["ray_planes"][vis])
        cmp_tol("ts(vis)", ts.reshape(-1)[vis], g["ts"][vis])
        cmp_tol("normals(vis)", nrm.reshape(P, 3)[vis], g["normals"][vis])
        if R_ref == R_ours and R_ref > 0:
            b = ref_loader.decode_ref_binning(rnat[10], R_ref)
            cmp_exact("point_list_keys", keys, b["point_list_keys"])
            cmp_exact("point_list", plist, b["point_list"])
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        im = ref_loader.decode_ref_image(rnat[11], W * H, tiles)
        cmp_exact("ranges", ranges.reshape(-1, 2), im["ranges"])
        cmp_exact("n_contrib", ncontrib.reshape(2, -1), im["n_contrib"])
        rnames = ["color", "radii", "coord", "mcoord", "depth", "mdepth", "alpha", "normal"]
        for n, a, b_ in zip(rnames, ours, refo):
            if n == "radii":
                continue
            cmp_tol("map:" + n, a, b_)

        # ---------------- backward ----------------
        def loss_of(outs):
            if args.geo:
                idx = [0, 2, 3, 4, 5, 6, 7]
                return sum((outs[i] * c).sum() for i, c in zip(idx, cot))
            return (outs[0] * cot[0]).sum()

        loss_of(ours).backward()
        torch.cuda.synchronize()
        # reference gradients: mean of 3 runs (float atomics are order-dependent)
        acc = None
        for _ in range(3):
            ri, ro = run(ref.GaussianRasterizationSettings, ref.GaussianRasterizer)
            loss_of(ro).backward()
            gs = [t.grad.detach().clone() for t in ri]
            acc = gs if acc is None else [a + b_ for a, b_ in zip(acc, gs)]
        acc = [a / 3 for a in acc]
        gnames = ["means3D", "means2D", "sh", "opacity", "scales", "rotations"]
        for n, t, rg in zip(gnames, ours_in, acc):
            if t.grad is None:
                print("grad", n, "is None!")
                continue
            cmp_tol("grad:" + n, t.grad, rg)
    else:
        (ours[0] * cot[0]).sum().backward()
        torch.cuda.synchronize()
        for n, t in zip(["means3D", "means2D", "sh", "opacity", "scales", "rotations"], ours_in):
            print("grad", n, None if t.grad is None else (float(t.grad.abs().max()), bool(torch.isfinite(t.grad).all())))

    # ---------------- timing ----------------
    if args.time:
        def bench(mod_settings, mod_rast, iters=20):
            for _ in range(3):
                i_, o_ = run(mod_settings, mod_rast)
                (o_[0] * cot[0]).sum().backward()
            torch.cuda.synchronize()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            tf = tb = 0.0
            for _ in range(iters):
                ev[0].record()
                i_, o_ = run(mod_settings, mod_rast)
                ev[1].record()
                (o_[0] * cot[0]).sum().backward()
                ev[2].record()
                torch.cuda.synchronize()
                tf += ev[0].elapsed_time(ev[1])
                tb += ev[1].elapsed_time(ev[2])
            return tf / iters, tb / iters
        f, b = bench(igs_b200.GaussianRasterizationSettings, igs_b200.GaussianRasterizer)
        print(f"[TIME] ours fwd {f:.3f} ms  bwd {b:.3f} ms")
        report["time_ours"] = [f, b]
        if ref is not None:
            f, b = bench(ref.GaussianRasterizationSettings, ref.GaussianRasterizer)
            print(f"[TIME] ref  fwd {f:.3f} ms  bwd {b:.3f} ms")
            report["time_ref"] = [f, b]

    with open(OUT / f"check_{args.cfg}{'_geo' if args.geo else ''}.json", "w") as f:
        json.dump(report, f, indent=1)


if __name__ == "__main__":
    main()
