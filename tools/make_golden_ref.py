"""Generates tests/golden/ref_<case>.npz from the UNMODIFIED reference CUDA rasterizer (baseline/_ref,
built by baseline/build_ref.sh) on a B200.  Run on the GPU box:

    python tools/make_golden_ref.py --out gpurun_out/golden        # then copy into tests/golden/

Each file holds the case inputs (in_*), the reference's decoded intermediates (radii, tiles_touched,
sorted keys, point_list, ranges, n_contrib, per-Gaussian preprocess arrays), its seven output maps,
and its gradients for two cotangent sets: colour only (what the refine loop produces) and all seven
maps.  Reference gradients are float-atomics ordered, so the mean of 3 runs is stored.
With --compare the library under test is run on the same inputs and differences are printed.
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import cases  # noqa: E402
import ref_loader  # noqa: E402


def run_module(mod, case, dev, cot=None, which="color"):
    settings_kw, tensors = cases.to_torch_args(case, dev)
    st = mod.GaussianRasterizationSettings(**settings_kw)
    leaves = {}
    kw = {}
    for k, v in tensors.items():
        if v is None:
            continue
        v = v.clone().requires_grad_(True)
        leaves[k] = v
        kw[k] = v
    m2 = torch.zeros_like(leaves["means3D"], requires_grad=True)
    leaves["means2D"] = m2
    outs = mod.GaussianRasterizer(raster_settings=st)(means2D=m2, **kw)
    color, radii, coord, mcoord, depth, mdepth, alpha, normal = outs
    maps = [color, coord, mcoord, depth, mdepth, alpha, normal]
    grads = None
    if cot is not None:
        if which == "color":
            loss = (maps[0] * cot[0]).sum()
        else:
            loss = sum((m * c).sum() for m, c in zip(maps, cot))
        loss.backward()
        grads = {}
        key = dict(means3D="means3D", means2D="means2D", sh="shs", colors_precomp="colors_precomp", opacities="opacities",
                   scales="scales", rotations="rotations", cov3D_precomp="cov3D_precomp")
        for gname, lname in key.items():
            if lname in leaves and leaves[lname].grad is not None:
                grads[gname] = leaves[lname].grad.detach().clone()
    return maps, radii, grads


def native_ref(ref, case, dev):
    settings_kw, t = cases.to_torch_args(case, dev)
    e = torch.Tensor([])
    o = lambda x: x if x is not None else e  # noqa: E731
    out = ref._C.rasterize_gaussians(settings_kw["bg"], t["means3D"], o(t["colors_precomp"]), t["opacities"], o(t["scales"]),
                                     o(t["rotations"]), settings_kw["scale_modifier"], o(t["cov3D_precomp"]),
                                     settings_kw["viewmatrix"], settings_kw["projmatrix"], settings_kw["tanfovx"],
                                     settings_kw["tanfovy"], settings_kw["kernel_size"], case["H"], case["W"], o(t["shs"]),
                                     case["sh_degree"], settings_kw["campos"], False, case["require_coord"],
                                     case["require_depth"], False)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=str(ROOT / "gpurun_out" / "golden"))
    ap.add_argument("--cases", default="all")
    ap.add_argument("--compare", action="store_true")
    args = ap.parse_args()
    out = Path(args.out)
    out.mkdir(parents=True, exist_ok=True)
    dev = torch.device("cuda:0")
    ref = ref_loader.load_reference()
    if ref is None:
        raise SystemExit("baseline/_ref is not built (run baseline/build_ref.sh in the build container)")
    names = list(cases.CASES) if args.cases == "all" else args.cases.split(",")
    ours = None
    if args.compare:
        import igs_b200
        ours = igs_b200.rasterizer
    for name in names:
        case = cases.CASES[name]()
        P, W, H = case["P"], case["W"], case["H"]
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        cot = [torch.from_numpy(c).to(dev) for c in cases.cotangents(case)]
        rec = {}
        for k, v in case.items():
            rec["in_" + k] = np.asarray(v)
        nat = native_ref(ref, case, dev)
        R = int(nat[0])
        rec["R"] = np.int64(R)
        g = ref_loader.decode_ref_geom(nat[9], P)
        radii = nat[8]
        vis = (radii > 0)
        rec["radii"] = radii.cpu().numpy()
        rec["tiles_touched"] = (g["tiles_touched"] * vis).cpu().numpy()  # scratch is uninitialised where culled
        for k in ("depths", "means2D", "conic_opacity", "rgb", "view_points", "camera_planes", "ray_planes", "ts", "normals"):
            a = g[k].clone()
            a[~vis] = 0
            rec["geom_" + k] = a.cpu().numpy()
        rec["geom_cov3D"] = g["cov3D"].cpu().numpy() if "cov3D_precomp" not in case else case["cov3D_precomp"]
        if R > 0:
            b = ref_loader.decode_ref_binning(nat[10], R)
            rec["point_list_keys"] = b["point_list_keys"].cpu().numpy()
            rec["point_list"] = b["point_list"].cpu().numpy()
        else:
            rec["point_list_keys"] = np.zeros(0, np.int64)
            rec["point_list"] = np.zeros(0, np.int32)
        im = ref_loader.decode_ref_image(nat[11], W * H, tiles)
        rec["ranges"] = im["ranges"].cpu().numpy() if R > 0 else np.zeros((tiles, 2), np.int32)
        rec["n_contrib"] = im["n_contrib"].cpu().numpy()
        maps, _, _ = run_module(ref, case, dev)
        for n, m in zip(cases.MAP_NAMES, maps):
            rec["map_" + n] = m.detach().cpu().numpy()
        for which in ("color", "all"):
            acc = None
            for _ in range(3):
                _, _, gr = run_module(ref, case, dev, cot, which)
                acc = gr if acc is None else {k: acc[k] + gr[k] for k in gr}
            for k, v in acc.items():
                rec[f"grad_{which}_{k}"] = (v / 3).cpu().numpy()
        np.savez_compressed(out / f"ref_{name}.npz", **rec)
        print(f"[golden] {name}: P={P} {W}x{H} R={R} visible={int(vis.sum())} -> {out / f'ref_{name}.npz'}")

        if ours is not None:
            try:
                omaps, oradii, _ = run_module(ours, case, dev)
                print(f"   radii mismatches {int((oradii != radii).sum())}")
                for n, a, b_ in zip(cases.MAP_NAMES, omaps, maps):
                    d = (a - b_).abs()
                    print(f"   map {n:7s} max-abs {float(d.max()):.3e} (nan ours {int(torch.isnan(a).sum())} ref {int(torch.isnan(b_).sum())})")
                for which in ("color", "all"):
                    _, _, og = run_module(ours, case, dev, cot, which)
                    for k, v in og.items():
                        r = torch.from_numpy(rec[f"grad_{which}_{k}"]).to(dev)
                        den = float(r.abs().max()) + 1e-30
                        print(f"   grad[{which}] {k:14s} max-abs {float((v - r).abs().max()):.3e} ref-max {den:.3e} "
                              f"rel {float((v - r).abs().max()) / den:.3e} nan {int(torch.isnan(v).sum())}/{int(torch.isnan(r).sum())}")
            except Exception as ex:  # keep generating goldens even if the library under test fails
                print("   COMPARE FAILED:", repr(ex))


if __name__ == "__main__":
    main()
