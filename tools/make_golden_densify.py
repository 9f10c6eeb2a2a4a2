"""Generates tests/golden/ref_densify_<case>.npz by running the reference's OWN GaussianModel
(/root/reference/igs/models/gaussian_model.py, unmodified) through the call path of the refine loop:

    GaussianModel(3) -> load_fromstream(stream, training_lr, refine_item)      infer_batch.py:266-272, gaussian_model.py:265-348
    -> optimizer.step() x2 (so that the Adam moments are not trivial)          infer_batch.py:323
    -> max_radii2D / add_densification_stats x3                                infer_batch.py:311-313, gaussian_model.py:865-868
    -> densify_and_prune(thr, 0.005, extent, None, control_max)                infer_batch.py:314-320, gaussian_model.py:640-666

in the build container (CPU; the literal device="cuda" of the reference is mapped to the CPU by tools/ref_import.py).
The files hold the inputs (stream Gaussians, the gradients fed to Adam and to the statistics, the RNG seed) and the
reference's results: every parameter and both Adam moments after the call, and the statistics before it.

Cases: `refine` (the refine loop as shipped: percent_dense stays 0 -> no clones, every selected Gaussian is split),
`bounded` (max_num forces the top-k branch, :648-653), `cloned` (percent_dense set by hand as training_setup would:
the clone branch :619-638).  Run here:  python tools/make_golden_densify.py
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tools"))
import ref_import  # noqa: E402

LRS = dict(position_lr_init=0.0016, rotation_lr=0.01, feature_lr=0.0025, opacity_lr=0.05, scaling_lr=0.005)
GROUPS = ("xyz", "rotation", "shs", "opacity", "scaling")
CASES = {
    "refine": dict(P=400, seed=11, max_num=100000, percent_dense=None, thr=0.00015, extent=4.0, control_max=True),
    "bounded": dict(P=400, seed=12, max_num=460, percent_dense=None, thr=0.00015, extent=4.0, control_max=True),
    "cloned": dict(P=360, seed=13, max_num=100000, percent_dense=0.01, thr=0.00015, extent=6.0, control_max=True),
}


def make_inputs(P, seed):
    g = torch.Generator().manual_seed(seed)
    inp = dict(
        xyz=4.0 * torch.rand(P, 3, generator=g) - 2.0,
        rotation=torch.randn(P, 4, generator=g),
        shs=0.3 * torch.randn(P, 16, 3, generator=g),
        opacity=2.5 * torch.randn(P, 1, generator=g) - 1.0,          # logits: a share falls below sigmoid^-1(0.005)
        scaling=float(np.log(0.05)) + 0.8 * torch.randn(P, 3, generator=g),
    )
    adam_grads = [{k: 1e-2 * torch.randn(v.shape, generator=g) for k, v in inp.items()} for _ in range(2)]
    stat_grads = [1.2e-4 * torch.randn(P, 3, generator=g) for _ in range(3)]
    stat_vis = [torch.rand(P, generator=g) < 0.7 for _ in range(3)]
    stat_radii = [torch.randint(1, 40, (P,), generator=g).float() for _ in range(3)]
    return inp, adam_grads, stat_grads, stat_vis, stat_radii


def run_reference(gm, c):
    inp, adam_grads, stat_grads, stat_vis, stat_radii = make_inputs(c["P"], c["seed"])
    out = {f"in_{k}": v.numpy().copy() for k, v in inp.items()}
    with ref_import.cuda_as_cpu():
        gs = gm.GaussianModel(3)
        stream = ref_import.AttrDict({k: v.clone() for k, v in inp.items()}, mask=None)
        item = ref_import.AttrDict(use_new_shs=False, use_mask=False, no_shs=False, no_opacity=False, no_scaling=False)
        gs.load_fromstream(stream, LRS, item)
        gs.max_num = c["max_num"]
        if c["percent_dense"] is not None:
            gs.percent_dense = c["percent_dense"]
        params = dict(xyz=gs._xyz, rotation=gs._rotation, shs=gs._shs, opacity=gs._opacity, scaling=gs._scaling)
        for i, gr in enumerate(adam_grads):
            for k, p in params.items():
                p.grad = gr[k].clone()
                out[f"adamgrad{i}_{k}"] = gr[k].numpy().copy()
            gs.optimizer.step()
            gs.optimizer.zero_grad(set_to_none=True)
        for k, p in params.items():
            st = gs.optimizer.state[p]
            out[f"pre_{k}"] = p.detach().numpy().copy()
            out[f"pre_m_{k}"] = st["exp_avg"].numpy().copy()
            out[f"pre_v_{k}"] = st["exp_avg_sq"].numpy().copy()
        with torch.no_grad():
            for i in range(3):
                vsp = torch.zeros(c["P"], 3, requires_grad=True)
                vsp.grad = stat_grads[i].clone()
                vis = stat_vis[i]
                gs.max_radii2D[vis] = torch.max(gs.max_radii2D[vis], stat_radii[i][vis])      # infer_batch.py:312
                gs.add_densification_stats(vsp, vis)
                out[f"stat{i}_grad"] = stat_grads[i].numpy().copy()
                out[f"stat{i}_vis"] = vis.numpy().copy()
                out[f"stat{i}_radii"] = stat_radii[i].numpy().copy()
            out["grad_accum"] = gs.xyz_gradient_accum.numpy().copy()
            out["denom"] = gs.denom.numpy().copy()
            out["max_radii2D"] = gs.max_radii2D.numpy().copy()
            torch.manual_seed(1000 + c["seed"])
            gs.densify_and_prune(c["thr"], 0.005, c["extent"], None, control_max=c["control_max"])
        after = dict(xyz=gs._xyz, rotation=gs._rotation, shs=gs._shs, opacity=gs._opacity, scaling=gs._scaling)
        for k, p in after.items():
            st = gs.optimizer.state.get(p, None)
            out[f"post_{k}"] = p.detach().numpy().copy()
            out[f"post_m_{k}"] = st["exp_avg"].numpy().copy()
            out[f"post_v_{k}"] = st["exp_avg_sq"].numpy().copy()
    for k in ("max_num", "thr", "extent", "seed", "P"):
        out[f"cfg_{k}"] = np.float64(c[k])
    out["cfg_percent_dense"] = np.float64(0.0 if c["percent_dense"] is None else c["percent_dense"])
    out["cfg_rng_seed"] = np.int64(1000 + c["seed"])
    return out


def main():
    gm = ref_import.load_gaussian_model()
    dst = ROOT / "tests" / "golden"
    for name, c in CASES.items():
        out = run_reference(gm, c)
        keep = {k: v for k, v in out.items() if not k.startswith("adamgrad")}   # Adam inputs are not needed by the test
        np.savez_compressed(dst / f"ref_densify_{name}.npz", **keep)
        print(name, "P", c["P"], "->", out["post_xyz"].shape[0], "selected>=thr",
              int((np.nan_to_num(out["grad_accum"] / out["denom"]) >= c["thr"]).sum()))


if __name__ == "__main__":
    main()
