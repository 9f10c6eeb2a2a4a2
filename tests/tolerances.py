"""Parity tolerances (BASELINE.json north_star): bit-exact integer/index outputs, max-abs 1e-4 on the
seven fp32 maps, 1e-3 relative on gradients.  Reference gradients are sums of float atomics in arbitrary order
(SURVEY T24), so element-wise relative error is not defined by the reference itself; two per-tensor criteria are
used together:
  * GRAD_RTOL  max|a-b| / max|ref|      (no element may be off by more than 1e-3 of the tensor's scale), and
  * GRAD_RL2   ||a-b||_2 / ||ref||_2    (the error ENERGY is bounded too: a max-normalised bar alone lets every
               element 100x smaller than the maximum be 10 % wrong).  Against the mean of three runs of the reference
               build the observed values are 2e-7 .. 3e-5 (1.8e-4 for rotations before the conditioning mask)."""
import numpy as np

MAP_ATOL = 1e-4
GRAD_RTOL = 1e-3
GRAD_RL2 = 1e-4


def grad_rel_l2(a, ref):
    a = np.asarray(a, np.float64).reshape(np.shape(ref))
    ref = np.asarray(ref, np.float64)
    return float(np.linalg.norm(a - ref) / (np.linalg.norm(ref) + 1e-30))


def grad_err(a, ref):
    a = np.asarray(a, np.float64).reshape(ref.shape)
    ref = np.asarray(ref, np.float64)
    den = np.abs(ref).max() + 1e-30
    return float(np.abs(a - ref).max() / den)


def well_conditioned_rotation_mask(scales, min_ratio=1.2, scale_modifier=1.0):
    """Rotation gradients are a difference of nearly equal terms when two axes of a Gaussian have nearly the
    same scale (exactly 0 for equal axes), and the plane/normal branch flips when the smallest eigenvalue
    s_min^2 sits at the reference's 1e-8 `well_conditioned` threshold (forward.cu:142): in both situations the
    reference's own fp32 value is rounding noise.  Such Gaussians are excluded from the
    ROTATION-gradient comparison only; they are still compared through every other gradient."""
    if scales is None:
        return slice(None)
    s = np.sort(np.asarray(scales, np.float64) * scale_modifier, axis=1)
    ratios = np.minimum(s[:, 1] / np.maximum(s[:, 0], 1e-30), s[:, 2] / np.maximum(s[:, 1], 1e-30))
    near_threshold = (s[:, 0] ** 2 < 1e-7)
    return (ratios >= min_ratio) & ~near_threshold
