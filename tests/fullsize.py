"""Full-size parity cases (BASELINE.json configs[1] / configs[2]) shared by tools/make_golden_fullsize.py (run on the
GPU box against the UNMODIFIED reference CUDA build, baseline/_ref) and tests/test_parity_fullsize_gpu.py.

  C2    200k Gaussians, one 512x512 view                     (configs[1])
  C3v7  300k Gaussians, view 7 of the 20 1352x1014 views     (one view of configs[2])
  C3x3  the same Gaussians, views 3 / 11 / 18 in ONE batched call: R ~ 5 M instances, i.e. above the size where
        the radix sorts leave the single cooperative launch for the multi-launch path (binning.cu)

What is compared (north_star): bit-exact radii, tiles_touched, sorted 64-bit keys, point_list, tile ranges,
n_contrib; the seven maps <= 1e-4 max-abs; the gradients of the colour-only cotangent (what the refine loop produces)
<= 1e-3 max-normalised AND <= 1e-4 relative L2 against the mean of three reference runs (float atomics).

The reference's outputs are far too large to commit, so the goldens hold sha256 digests of the integer outputs and
seeded PROBES of the float outputs (1024 pixels per map, 1024 Gaussians per gradient, plus each reference gradient's
max and L2 norm): with them the test asserts on a box where baseline/_ref is absent instead of skipping.
Test infrastructure only.
"""
from __future__ import annotations

import hashlib
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

from igs_b200 import scenes  # noqa: E402

GOLDEN_DIR = Path(__file__).resolve().parent / "golden"
CASES = {
    "C2": dict(cfg="C2", views=[0]),
    "C3v7": dict(cfg="C3", views=[7]),
    "C3x3": dict(cfg="C3", views=[3, 11, 18]),
}
INT_NAMES = ["radii", "tiles_touched", "keys", "point_list", "ranges", "n_contrib"]
MAP_NAMES = ["color", "coord", "mcoord", "depth", "mdepth", "alpha", "normal"]
GRAD_NAMES = ["means3D", "means2D", "sh", "opacities", "scales", "rotations"]
N_PROBE = 1024


def _activated_reproducible(sc):
    """scenes.activated (exp, normalize, sigmoid) evaluated in float64 and rounded ONCE to float32.  torch's float32 CPU
    exp runs vectorised per thread chunk and was seen to differ in the last bit between processes (5 of 45 runs at
    P = 300k: tools/determinism_probe.sh), which moved the radii of ~10 Gaussians and with them every committed digest; the
    float64 value rounds to the same float32 whatever code path computed it."""
    scales = torch.from_numpy(np.exp(sc.log_scale.double().numpy()).astype(np.float32))
    opac = torch.from_numpy((1.0 / (1.0 + np.exp(-sc.opacity_logit.double().numpy()))).astype(np.float32))
    r = sc.rot.double()
    rot = (r / r.norm(dim=1, keepdim=True).clamp_min(1e-12)).float()
    return scales, rot, opac


def inputs(name: str, dev):
    c = CASES[name]
    cfg = scenes.CONFIGS[c["cfg"]]
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    sc = scenes.make_gaussians(P, 0, cfg["log_scale_mean"])
    cams = scenes.make_cameras(cfg["V"], W, H, 1)
    scales, rot, opac = _activated_reproducible(sc)
    vs = c["views"]
    g = torch.Generator().manual_seed(3 + 17 * len(vs))
    cot = torch.randn(len(vs), 3, H, W, generator=g)
    return dict(P=P, W=W, H=H, views=vs, tanfovx=cams.tanfovx, tanfovy=cams.tanfovy,
                means3D=sc.xyz.to(dev), shs=sc.sh.to(dev), opacities=opac.to(dev), scales=scales.to(dev), rotations=rot.to(dev),
                scales_np=scales.numpy(), bg=torch.zeros(3, device=dev),
                vm=cams.viewmatrix[vs].to(dev).contiguous(), pm=cams.projmatrix[vs].to(dev).contiguous(),
                cp=cams.campos[vs].to(dev).contiguous(), cot=cot.to(dev))


def _leaves(inp):
    return {k: inp[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}


def run_ours(inp, native):
    """One batched call of this library over the case's views (V == 1: the drop-in single-view package).
    Returns dict(ints={name: [per-view tensors]}, maps={name: [V,...]}, grads={name: tensor})."""
    import igs_b200
    from igs_b200.multiview import MultiViewSettings, rasterize_views
    P, W, H, V = inp["P"], inp["W"], inp["H"], len(inp["views"])
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    e = torch.Tensor([])
    if V == 1:
        nat = native.rasterize_gaussians(inp["bg"], inp["means3D"], e, inp["opacities"], inp["scales"], inp["rotations"], 1.0, e,
                                         inp["vm"][0], inp["pm"][0], inp["tanfovx"], inp["tanfovy"], 0.0, H, W, inp["shs"], 3,
                                         inp["cp"][0], False, True, True, False)
    else:
        nat = native.rasterize_views(inp["bg"], inp["means3D"], e, inp["opacities"], inp["scales"], inp["rotations"], 1.0, e,
                                     inp["vm"], inp["pm"], inp["tanfovx"], inp["tanfovy"], 0.0, H, W, inp["shs"], 3, inp["cp"],
                                     False, True, True, False)
    R = int(nat[0])
    keys, plist, ranges, ncontrib, tiles_t = native.decode_buffers(P, V, R, H, W, nat[9], nat[10], nat[11])[:5]
    radii = nat[8].reshape(V, P)
    tiles_t = tiles_t.reshape(V, P)
    Rv = tiles_t.to(torch.int64).sum(dim=1)
    off = torch.cumsum(Rv, 0) - Rv
    ints = {n: [] for n in INT_NAMES}
    ranges = ranges.reshape(V, tiles, 2).to(torch.int64)
    for v in range(V):
        a, b = int(off[v]), int(off[v] + Rv[v])
        ints["radii"].append(radii[v].to(torch.int32))
        ints["tiles_touched"].append(tiles_t[v].to(torch.int32))
        ints["keys"].append(keys[a:b].to(torch.int64))
        ints["point_list"].append(plist[a:b].to(torch.int32))
        r = ranges[v]
        nonempty = (r[:, 1] > r[:, 0]).unsqueeze(1)
        ints["ranges"].append(torch.where(nonempty, r - a, torch.zeros_like(r)).to(torch.int32))
        ints["n_contrib"].append(ncontrib.reshape(V, 2, H * W)[v].to(torch.int32))
    del nat, keys, plist
    # maps + gradients through the public autograd entry points
    lv = _leaves(inp)
    if V == 1:
        st = igs_b200.GaussianRasterizationSettings(H, W, inp["tanfovx"], inp["tanfovy"], 0.0, inp["bg"], 1.0, inp["vm"][0],
                                                    inp["pm"][0], 3, inp["cp"][0], False, True, True, False)
        m2 = torch.zeros(P, 3, device=inp["means3D"].device, requires_grad=True)
        o = igs_b200.GaussianRasterizer(st)(means3D=lv["means3D"], means2D=m2, shs=lv["shs"], opacities=lv["opacities"],
                                            scales=lv["scales"], rotations=lv["rotations"])
        (o[0] * inp["cot"][0]).sum().backward()
        maps = {n: t.detach().unsqueeze(0) for n, t in zip(MAP_NAMES, (o[0], o[2], o[3], o[4], o[5], o[6], o[7]))}
        g2d = m2.grad.unsqueeze(0)
    else:
        st = MultiViewSettings(H, W, inp["tanfovx"], inp["tanfovy"], 0.0, inp["bg"], 1.0, inp["vm"], inp["pm"], 3, inp["cp"])
        o = rasterize_views(lv["means3D"], lv["shs"], lv["opacities"], lv["scales"], lv["rotations"], st)
        (o[0] * inp["cot"]).sum().backward()
        maps = {n: t.detach() for n, t in zip(MAP_NAMES, (o[0], o[2], o[3], o[4], o[5], o[6], o[7]))}
        g2d = o[0].grad_fn.grad_means2D
    grads = dict(means3D=lv["means3D"].grad, means2D=g2d, sh=lv["shs"].grad, opacities=lv["opacities"].grad,
                 scales=lv["scales"].grad, rotations=lv["rotations"].grad)
    return dict(ints=ints, maps=maps, grads=grads, R=R)


def run_ref(ref, inp, grad_runs: int = 3):
    """The unmodified reference build, view by view through its own Python API; gradients are the mean of
    `grad_runs` runs of the per-view-accumulated backward."""
    import ref_loader
    P, W, H, V = inp["P"], inp["W"], inp["H"], len(inp["views"])
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    e = torch.Tensor([])
    ints = {n: [] for n in INT_NAMES}
    Rtot = 0
    for v in range(V):
        rnat = ref._C.rasterize_gaussians(inp["bg"], inp["means3D"], e, inp["opacities"], inp["scales"], inp["rotations"], 1.0, e,
                                          inp["vm"][v], inp["pm"][v], inp["tanfovx"], inp["tanfovy"], 0.0, H, W, inp["shs"], 3,
                                          inp["cp"][v], False, True, True, False)
        R = int(rnat[0])
        Rtot += R
        g = ref_loader.decode_ref_geom(rnat[9], P)
        b = ref_loader.decode_ref_binning(rnat[10], R)
        im = ref_loader.decode_ref_image(rnat[11], W * H, tiles)
        ints["radii"].append(rnat[8].to(torch.int32).clone())
        ints["tiles_touched"].append(g["tiles_touched"].to(torch.int32).clone())
        ints["keys"].append(b["point_list_keys"].to(torch.int64).clone())
        ints["point_list"].append(b["point_list"].to(torch.int32).clone())
        ints["ranges"].append(im["ranges"].to(torch.int32).clone())
        ints["n_contrib"].append(im["n_contrib"].to(torch.int32).clone())
        del rnat
    maps, acc, g2d_acc = None, None, None
    for _ in range(grad_runs):
        lv = _leaves(inp)
        per_view_maps, g2d = [], []
        for v in range(V):
            st = ref.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=inp["tanfovx"], tanfovy=inp["tanfovy"],
                                                   kernel_size=0.0, bg=inp["bg"], scale_modifier=1.0, viewmatrix=inp["vm"][v],
                                                   projmatrix=inp["pm"][v], sh_degree=3, campos=inp["cp"][v], prefiltered=False,
                                                   require_depth=True, require_coord=True, debug=False)
            m2 = torch.zeros(P, 3, device=inp["means3D"].device, requires_grad=True)
            o = ref.GaussianRasterizer(raster_settings=st)(means3D=lv["means3D"], means2D=m2, shs=lv["shs"],
                                                           opacities=lv["opacities"], scales=lv["scales"],
                                                           rotations=lv["rotations"])
            (o[0] * inp["cot"][v]).sum().backward()
            g2d.append(m2.grad)
            if maps is None:
                per_view_maps.append([t.detach() for t in (o[0], o[2], o[3], o[4], o[5], o[6], o[7])])
        if maps is None:
            maps = {n: torch.stack([pv[i] for pv in per_view_maps]) for i, n in enumerate(MAP_NAMES)}
        gs = dict(means3D=lv["means3D"].grad, means2D=torch.stack(g2d), sh=lv["shs"].grad, opacities=lv["opacities"].grad,
                  scales=lv["scales"].grad, rotations=lv["rotations"].grad)
        acc = gs if acc is None else {k: acc[k] + gs[k] for k in acc}
    grads = {k: t / grad_runs for k, t in acc.items()}
    return dict(ints=ints, maps=maps, grads=grads, R=Rtot)


def digest(t: torch.Tensor) -> str:
    return hashlib.sha256(t.contiguous().cpu().numpy().tobytes()).hexdigest()


def probe_index(n: int, seed: int, count: int = N_PROBE) -> np.ndarray:
    rng = np.random.RandomState(seed)
    return np.sort(rng.choice(n, size=min(count, n), replace=False)).astype(np.int64)


def rotation_mask(inp):
    sys.path.insert(0, str(Path(__file__).resolve().parent))
    from tolerances import well_conditioned_rotation_mask
    return well_conditioned_rotation_mask(inp["scales_np"])


def golden_path(name: str) -> Path:
    return GOLDEN_DIR / f"ref_fullsize_{name}.npz"


def make_golden(ref_out, inp) -> dict:
    """The committed summary of one reference run (see module docstring)."""
    P, V = inp["P"], len(inp["views"])
    g = {"R": np.int64(ref_out["R"])}
    for n in INT_NAMES:
        g["sha_" + n] = np.array([digest(t) for t in ref_out["ints"][n]])
    pix = probe_index(inp["W"] * inp["H"], 101)
    g["probe_pix"] = pix
    for n in MAP_NAMES:
        m = ref_out["maps"][n]
        g["map_" + n] = m.reshape(V, m.shape[1], -1)[:, :, torch.from_numpy(pix).to(m.device)].cpu().numpy()
    gi = probe_index(P, 202)
    g["probe_gauss"] = gi
    gidx = torch.from_numpy(gi).to(inp["means3D"].device)
    rm = rotation_mask(inp)
    for n in GRAD_NAMES:
        t = ref_out["grads"][n]
        rows = t[:, gidx] if n == "means2D" else t[gidx]
        g["grad_" + n] = rows.cpu().numpy()
        full = t.double()
        if n == "rotations" and not isinstance(rm, slice):
            full = full[torch.from_numpy(rm).to(full.device)]
        g["gradmax_" + n] = np.float64(full.abs().max().item())
        g["gradl2_" + n] = np.float64(full.norm().item())
    return g
