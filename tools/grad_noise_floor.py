"""Noise floor of the full-size gradient comparison (GPU box; needs baseline/_ref).

The reference's backward accumulates with atomics, so two runs of the UNMODIFIED reference differ; this tool measures
by how much, in the two criteria of tests/test_parity_fullsize_gpu.py (max-normalised and relative L2), next to the
distance between this library and the reference:

    python tools/grad_noise_floor.py --case C3v7 [--save ours.pt] [--against other.pt]

  ref vs ref   : single run A vs single run B, and 3-run mean A vs 3-run mean B
  ours vs ref  : this library (deterministic) vs the 3-run mean
  ours vs file : this library vs gradients saved by another invocation (e.g. another IGS_BWD2_CFG) with --save
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import fullsize  # noqa: E402
import ref_loader  # noqa: E402


def errs(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / (b.abs().max() + 1e-30)), float((a - b).norm() / (b.norm() + 1e-30))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="C3v7")
    ap.add_argument("--save", default=None)
    ap.add_argument("--against", default=None)
    ap.add_argument("--no-ref", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    inp = fullsize.inputs(args.case, dev)
    from igs_b200 import _C as native
    ours = fullsize.run_ours(inp, native)["grads"]
    rm = fullsize.rotation_mask(inp)
    rm_t = None if isinstance(rm, slice) else torch.from_numpy(rm).to(dev)

    def masked(n, t):
        return t[rm_t] if (n == "rotations" and rm_t is not None) else t

    if args.save:
        torch.save({k: v.cpu() for k, v in ours.items()}, args.save)
    if args.against:
        other = torch.load(args.against)
        for n in fullsize.GRAD_NAMES:
            e = errs(masked(n, ours[n]), masked(n, other[n].to(dev)))
            print(f"[{args.case}] ours vs {Path(args.against).name:12s} {n:10s} max-norm {e[0]:.3e} rel-l2 {e[1]:.3e}")
    if args.no_ref:
        return
    ref = ref_loader.load_reference()
    if ref is None:
        raise SystemExit("baseline/_ref is not built on this box")
    r1a = fullsize.run_ref(ref, inp, grad_runs=1)["grads"]
    r1b = fullsize.run_ref(ref, inp, grad_runs=1)["grads"]
    r3a = fullsize.run_ref(ref, inp, grad_runs=3)["grads"]
    r3b = fullsize.run_ref(ref, inp, grad_runs=3)["grads"]
    for n in fullsize.GRAD_NAMES:
        rows = [("ref1 vs ref1", errs(masked(n, r1a[n]), masked(n, r1b[n]))),
                ("ref3 vs ref3", errs(masked(n, r3a[n]), masked(n, r3b[n]))),
                ("ours vs ref3", errs(masked(n, ours[n]), masked(n, r3a[n]))),
                ("ours vs ref3'", errs(masked(n, ours[n]), masked(n, r3b[n])))]
        for tag, e in rows:
            print(f"[{args.case}] {tag:14s} {n:10s} max-norm {e[0]:.3e} rel-l2 {e[1]:.3e}")


if __name__ == "__main__":
    main()
