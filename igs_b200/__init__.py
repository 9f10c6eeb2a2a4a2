"""igs_b200 — B200-native differentiable Gaussian-splat rasterizer for IGS's refine loop.

Only what the hot path needs lives here: `csrc/` (hand-written sm_100a kernels + C-ABI),
`rasterizer.py` (the reference's Python API), `multiview.py` (batched V-view call),
`refine.py` (the refine-iteration driver used by bench.py), `densify.py` (densify / prune on the packed record),
`scenes.py` (seeded synthetic inputs).

API names are resolved lazily (PEP 562) so that `python -m igs_b200.build` and `igs_b200.scenes`
work before the CUDA extension exists; touching any rasterizer name without a built `igs_b200._C`
raises ImportError — there is deliberately no CPU or PyTorch fallback.
"""
import importlib

_API = {
    "GaussianRasterizationSettings": "rasterizer", "GaussianRasterizer": "rasterizer",
    "GaussianRasterizerClamp": "rasterizer", "rasterize_gaussians": "rasterizer",
    "MultiViewSettings": "multiview", "rasterize_views": "multiview",
    "RefineState": "refine", "refine_iteration": "refine",
    "DensifyStats": "densify", "densify_and_prune": "densify",
}
__all__ = sorted(_API)


def __getattr__(name):
    if name in _API:
        return getattr(importlib.import_module("." + _API[name], __name__), name)
    if name == "_C":
        import torch  # noqa: F401  libtorch must be loaded by torch itself before the extension that links it
    if name in ("_C", "rasterizer", "multiview", "refine", "densify", "scenes", "build"):
        return importlib.import_module("." + name, __name__)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
