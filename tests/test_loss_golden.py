"""CPU: the torch restatement of the refine loss (igs_b200.refine.refine_loss_torch) against values produced by the
reference's own loss_utils.py (tests/golden/refmath_loss.npz, generator tools/make_golden_refmath.py)."""
import numpy as np
import torch

import cases


def test_torch_loss_matches_reference_loss_utils(native):
    from igs_b200 import refine
    g = dict(np.load(cases.GOLDEN_DIR / "refmath_loss.npz"))
    for name in ("a", "b"):
        pred = torch.from_numpy(g[f"{name}_pred"]).unsqueeze(0).requires_grad_(True)
        gt = torch.from_numpy(g[f"{name}_gt"]).unsqueeze(0)
        loss = refine.refine_loss_torch(pred, gt, 0.8).sum()
        loss.backward()
        assert abs(float(loss) - float(g[f"{name}_loss"])) <= 1e-6
        assert np.abs(pred.grad[0].numpy() - g[f"{name}_dpred"]).max() <= 1e-7
