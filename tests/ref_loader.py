"""Loads the UNMODIFIED reference rasterizer built by baseline/build_ref.sh (baseline/_ref/, git-ignored,
shipped to the GPU box by gpurun) under an alias module name, and decodes its three opaque scratch
buffers with the reference's own carve order (DGR/cuda_rasterizer/rasterizer_impl.cu:190-250, 128-B
alignment per sub-array, rasterizer_impl.h:22-27).  Test infrastructure only.
"""
from __future__ import annotations

import importlib.util
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
REF_DIR = ROOT / "baseline" / "_ref" / "diff_gaussian_rasterization_rade"


def have_reference() -> bool:
    return (REF_DIR / "__init__.py").exists() and any(REF_DIR.glob("_C*.so"))


def load_reference():
    """Returns the reference python package (module object) or None."""
    if not have_reference():
        return None
    name = "ref_dgr_rade"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, REF_DIR / "__init__.py",
                                                  submodule_search_locations=[str(REF_DIR)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


class _Carver:
    def __init__(self, buf: torch.Tensor):
        self.buf = buf
        self.base = buf.data_ptr()
        self.off = 0

    def take(self, count: int, dtype: torch.dtype, itemsize: int):
        addr = (self.base + self.off + 127) & ~127
        start = addr - self.base
        nbytes = count * itemsize
        out = self.buf[start:start + nbytes].view(dtype)
        self.off = start + nbytes
        return out


def decode_ref_geom(geom: torch.Tensor, P: int) -> dict:
    c = _Carver(geom)
    d = {}
    d["depths"] = c.take(P, torch.float32, 4)
    d["camera_planes"] = c.take(P * 6, torch.float32, 4).view(P, 6)
    d["ray_planes"] = c.take(P * 2, torch.float32, 4).view(P, 2)
    d["ts"] = c.take(P, torch.float32, 4)
    d["normals"] = c.take(P * 3, torch.float32, 4).view(P, 3)
    d["clamped"] = c.take(P * 3, torch.uint8, 1).view(P, 3)
    d["internal_radii"] = c.take(P, torch.int32, 4)
    d["means2D"] = c.take(P * 2, torch.float32, 4).view(P, 2)
    d["view_points"] = c.take(P * 3, torch.float32, 4).view(P, 3)
    d["cov3D"] = c.take(P * 6, torch.float32, 4).view(P, 6)
    d["conic_opacity"] = c.take(P * 4, torch.float32, 4).view(P, 4)
    d["rgb"] = c.take(P * 3, torch.float32, 4).view(P, 3)
    d["tiles_touched"] = c.take(P, torch.int32, 4)
    return d


def decode_ref_binning(binning: torch.Tensor, R: int) -> dict:
    c = _Carver(binning)
    d = {}
    d["point_list"] = c.take(R, torch.int32, 4)
    d["point_list_unsorted"] = c.take(R, torch.int32, 4)
    d["point_list_keys"] = c.take(R, torch.int64, 8)
    d["point_list_keys_unsorted"] = c.take(R, torch.int64, 8)
    return d


def decode_ref_image(img: torch.Tensor, N: int, tiles: int) -> dict:
    c = _Carver(img)
    d = {}
    d["n_contrib"] = c.take(N * 2, torch.int32, 4).view(2, N)
    d["ranges"] = c.take(N * 2, torch.int32, 4).view(N, 2)[:tiles]
    c.take(N * 2, torch.int32, 4)  # point_ranges
    d["accum_coord"] = c.take(N * 3, torch.float32, 4).view(3, N)
    d["accum_depth"] = c.take(N, torch.float32, 4)
    d["normal_length"] = c.take(N, torch.float32, 4)
    return d
