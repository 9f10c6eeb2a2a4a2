"""Loads modules of the UNMODIFIED reference tree (/root/reference) in the build container, where neither a GPU nor
the reference's heavier dependencies exist, so that golden vectors can be produced by the reference's own code
(tools/make_golden_densify.py, tools/make_golden_renderer.py).  Test infrastructure; needs /root/reference.

Two aids, neither of which touches the reference's sources:
  * `load(...)` imports single files by path under their real dotted names, with `sys.modules` stubs for what the
    container lacks (plyfile, simple_knn, omegaconf-typed packages, ...): the stubs only satisfy `import` statements,
    the code under test never calls them;
  * `cuda_as_cpu()` maps the literal device="cuda" that the reference hard-codes in tensor factories
    (e.g. igs/models/gaussian_model.py:589,597,633; igs/utils/general_utils.py:66,84,103) to the CPU while it is
    active, so the same statements run here.
"""
from __future__ import annotations

import contextlib
import importlib.util
import sys
import types
from pathlib import Path

import torch

REF = Path("/root/reference")


def _stub(name: str, **attrs):
    m = sys.modules.get(name)
    if m is None:
        m = types.ModuleType(name)
        m.__path__ = []  # behaves as a package for `import a.b`
        sys.modules[name] = m
    for k, v in attrs.items():
        setattr(m, k, v)
    return m


def _stub_plyfile():
    """`plyfile` is absent from the image: tools/plyfile_min.py provides the few calls the reference makes."""
    sys.path.insert(0, str(Path(__file__).resolve().parent))
    import plyfile_min
    _stub("plyfile", PlyData=plyfile_min.PlyData, PlyElement=plyfile_min.PlyElement)


def load_file(dotted: str, rel: str):
    """Executes /root/reference/<rel> as module `dotted` (parents are stubbed as empty packages)."""
    parts = dotted.split(".")
    for i in range(1, len(parts)):
        _stub(".".join(parts[:i]))
    spec = importlib.util.spec_from_file_location(dotted, REF / rel)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[dotted] = mod
    spec.loader.exec_module(mod)
    parent = sys.modules.get(".".join(parts[:-1]))
    if parent is not None:
        setattr(parent, parts[-1], mod)
    return mod


def load_gaussian_model():
    """igs/models/gaussian_model.py (GaussianModel of the refine loop) with its helper modules loaded for real and
    the absent third-party / network-side imports stubbed."""
    _stub("icecream", ic=lambda *a, **k: None)
    _stub_plyfile()
    _stub("simple_knn")
    _stub("simple_knn._C", distCUDA2=None)
    load_file("igs.utils.general_utils", "igs/utils/general_utils.py")
    load_file("igs.utils.system_utils", "igs/utils/system_utils.py")
    load_file("igs.utils.graphics_utils", "igs/utils/graphics_utils.py")
    _stub("igs.models")
    if not hasattr(sys.modules.get("igs.models.gs"), "GS3DRenderer"):   # (keep the real module if load_gs_module ran)
        _stub("igs.models.gs", GaussianModel=object)   # only imported as a type name (gaussian_model.py:22)
    return load_file("igs.models.gaussian_model", "igs/models/gaussian_model.py")


_FACTORIES = ("zeros", "ones", "empty", "full", "tensor", "arange", "rand", "randn", "zeros_like", "ones_like", "eye",
              "linspace", "as_tensor")


@contextlib.contextmanager
def cuda_as_cpu():
    saved = {n: getattr(torch, n) for n in _FACTORIES}
    saved_to = torch.Tensor.to
    saved_cuda = torch.Tensor.cuda

    def wrap(fn):
        def f(*a, **k):
            d = k.get("device")
            if d is not None and str(d).startswith("cuda"):
                k["device"] = "cpu"
            return fn(*a, **k)
        return f

    def to(self, *a, **k):
        a = tuple("cpu" if (isinstance(x, str) and x.startswith("cuda")) else x for x in a)
        if "device" in k and str(k["device"]).startswith("cuda"):
            k["device"] = "cpu"
        return saved_to(self, *a, **k)

    try:
        for n, fn in saved.items():
            setattr(torch, n, wrap(fn))
        torch.Tensor.to = to
        torch.Tensor.cuda = lambda self, *a, **k: self
        yield
    finally:
        for n, fn in saved.items():
            setattr(torch, n, fn)
        torch.Tensor.to = saved_to
        torch.Tensor.cuda = saved_cuda


class AttrDict(dict):
    """Stands in for the OmegaConf node `refine_item` (attribute and item access, `.keys()`)."""
    __getattr__ = dict.__getitem__


class _Subscriptable:
    """Stands in for jaxtyping's Float[Tensor, "..."] etc. in annotations."""
    def __class_getitem__(cls, item):
        return cls


class RecordingRasterizer:
    """Fake `diff_gaussian_rasterization_rade_clamp`: records what the reference renderer hands to the rasterizer
    (settings + tensors) and returns blank maps of the right shapes, so the renderer-side code runs without a GPU."""
    calls = []

    class Settings(dict):
        def __getattr__(self, k):
            return self[k]

    @staticmethod
    def GaussianRasterizationSettings(**kw):
        return RecordingRasterizer.Settings(kw)

    class GaussianRasterizer:
        def __init__(self, raster_settings):
            self.s = raster_settings

        def __call__(self, **kw):
            RecordingRasterizer.calls.append((self.s, kw))
            H, W, P = self.s["image_height"], self.s["image_width"], kw["means3D"].shape[0]
            z = lambda c: torch.zeros(c, H, W)  # noqa: E731
            return z(3), torch.zeros(P, dtype=torch.int32), z(3), z(3), z(1), z(1), z(1), z(3)


def load_gs_module():
    """igs/models/gs.py (Camera, stream GaussianModel, GS3DRenderer, load_ply / save_ply) with the network-side and
    third-party imports stubbed and the clamp rasterizer replaced by `RecordingRasterizer`."""
    import dataclasses
    import typing
    _stub("icecream", ic=lambda *a, **k: None)
    _stub("fpsample")
    _stub("open3d")
    _stub("torch_cluster", fps=None, knn=None)
    _stub_plyfile()
    _stub("igs.models")
    _stub("igs.models.networks", MLP=object)
    ty = _stub("igs.utils.typing")
    for n in ("Any", "Callable", "Dict", "Iterable", "List", "Literal", "NamedTuple", "NewType", "Optional", "Sized", "Tuple",
              "Type", "TypeVar", "Union"):
        setattr(ty, n, getattr(typing, n))
    for n in ("Bool", "Complex", "Float", "Inexact", "Int", "Integer", "Num", "Shaped", "UInt"):
        setattr(ty, n, _Subscriptable)
    ty.Tensor = torch.Tensor
    ty.DictConfig = dict
    ty.__all__ = [n for n in vars(ty) if not n.startswith("_")]

    class BaseModule(torch.nn.Module):
        @dataclasses.dataclass
        class Config:
            pass
    _stub("igs.utils.base", BaseModule=BaseModule)
    _stub("igs.utils.ops", scale_tensor_batch=None, scale_tensor=None, select_points_bbox=None, position_embedding=None)
    load_file("igs.utils.general_utils", "igs/utils/general_utils.py")
    load_file("igs.utils.system_utils", "igs/utils/system_utils.py")
    load_file("igs.utils.graphics_utils", "igs/utils/graphics_utils.py")
    load_file("igs.utils.sh_utils", "igs/utils/sh_utils.py")
    _stub("diff_gaussian_rasterization_rade_clamp",
          GaussianRasterizationSettings=RecordingRasterizer.GaussianRasterizationSettings,
          GaussianRasterizer=RecordingRasterizer.GaussianRasterizer)
    return load_file("igs.models.gs", "igs/models/gs.py")
