"""Generates tests/golden/refmath_loss.npz by IMPORTING the reference's own torch-side loss
(/root/reference/igs/utils/loss_utils.py: l1_loss, ssim) in the build container (CPU) and evaluating the refine
loop's loss expression (infer_batch.py:301-305) and its autograd gradient on small seeded images.
Run here (needs /root/reference):  python tools/make_golden_refmath.py
"""
import importlib.util
import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference/igs/utils/loss_utils.py")


def main():
    sys.modules.setdefault("icecream", types.SimpleNamespace(ic=lambda *a, **k: None))  # loss_utils.py:15 imports it
    spec = importlib.util.spec_from_file_location("ref_loss_utils", REF)
    lu = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(lu)
    g = torch.Generator().manual_seed(123)
    out = {}
    for name, (H, W) in dict(a=(37, 53), b=(64, 96)).items():
        pred = torch.rand(3, H, W, generator=g).requires_grad_(True)
        gt = torch.rand(3, H, W, generator=g)
        if name == "b":
            gt = (pred.detach() + 0.05 * torch.randn(3, H, W, generator=g)).clamp(0, 1)  # a near-converged pair
        lam = 0.8
        loss = lam * lu.l1_loss(pred, gt) + (1 - lam) * (1.0 - lu.ssim(pred, gt.unsqueeze(0), size_average=False))
        loss = loss.sum()
        loss.backward()
        out[f"{name}_pred"] = pred.detach().numpy()
        out[f"{name}_gt"] = gt.numpy()
        out[f"{name}_loss"] = np.float32(loss.item())
        out[f"{name}_dpred"] = pred.grad.numpy()
    (ROOT / "tests" / "golden").mkdir(parents=True, exist_ok=True)
    np.savez_compressed(ROOT / "tests" / "golden" / "refmath_loss.npz", **out)
    print({k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items()})


if __name__ == "__main__":
    main()
