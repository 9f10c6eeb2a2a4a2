"""Generates tests/golden/ref_fullsize_<case>.npz from the UNMODIFIED reference CUDA rasterizer (baseline/_ref) on a
B200, for the full-size parity cases of tests/fullsize.py.  Run on the GPU box:

    python tools/make_golden_fullsize.py --out gpurun_out/golden [--compare]     # then copy into tests/golden/

Each file holds sha256 digests of the reference's integer outputs per view (radii, tiles_touched, sorted keys,
point_list, ranges, n_contrib) and seeded probes of its float outputs (1024 pixels of each of the seven maps, 1024
Gaussians of each gradient = mean of three runs, plus every gradient's max and L2 norm).  With --compare the library
under test is run on the same inputs and a table of differences (both gradient criteria) is printed.
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

import fullsize  # noqa: E402
import ref_loader  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=str(ROOT / "gpurun_out" / "golden"))
    ap.add_argument("--cases", default="all")
    ap.add_argument("--compare", action="store_true")
    args = ap.parse_args()
    out = Path(args.out)
    out.mkdir(parents=True, exist_ok=True)
    ref = ref_loader.load_reference()
    if ref is None:
        raise SystemExit("baseline/_ref is not built on this box")
    dev = torch.device("cuda:0")
    names = list(fullsize.CASES) if args.cases == "all" else args.cases.split(",")
    for name in names:
        inp = fullsize.inputs(name, dev)
        r = fullsize.run_ref(ref, inp)
        g = fullsize.make_golden(r, inp)
        np.savez_compressed(out / f"ref_fullsize_{name}.npz", **g)
        print(f"[{name}] R={r['R']} written {out / f'ref_fullsize_{name}.npz'}", flush=True)
        if not args.compare:
            continue
        from igs_b200 import _C
        o = fullsize.run_ours(inp, _C)
        print(f"[{name}] R ours {o['R']} ref {r['R']}")
        for n in fullsize.INT_NAMES:
            for v, (a, b) in enumerate(zip(o["ints"][n], r["ints"][n])):
                same = a.shape == b.shape and bool(torch.equal(a, b))
                print(f"[{name}] EXACT {n:14s} view {v}: {'ok' if same else 'MISMATCH'}")
        for n in fullsize.MAP_NAMES:
            print(f"[{name}] MAP   {n:14s} max-abs {float((o['maps'][n] - r['maps'][n]).abs().max()):.3e}")
        rm = fullsize.rotation_mask(inp)
        for n in fullsize.GRAD_NAMES:
            a, b = o["grads"][n].double(), r["grads"][n].double()
            if n == "rotations" and not isinstance(rm, slice):
                m = torch.from_numpy(rm).to(a.device)
                print(f"[{name}] GRAD  {n:14s} (unmasked) max-norm {float((a - b).abs().max() / b.abs().max()):.3e} "
                      f"rel-l2 {float((a - b).norm() / b.norm()):.3e}")
                a, b = a[m], b[m]
            print(f"[{name}] GRAD  {n:14s} max-norm {float((a - b).abs().max() / b.abs().max()):.3e} "
                  f"rel-l2 {float((a - b).norm() / b.norm()):.3e}  ref-max {float(b.abs().max()):.3e}")
        del o, r
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
