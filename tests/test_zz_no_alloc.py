"""GPU: igs_depth_normal_loss_noalloc (static __device__ scratch, no allocation; no counterpart in the reference)
against the existing igs_depth_normal_loss, through the C-ABI with raw device
pointers, on two calls in a row (the static array keeps the previous call's contents).  Collected last on purpose
(file name)."""
import ctypes as C
import math
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parent.parent
ALLOC = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)
MAX_TILES = 65536   # IGS_NOALLOC_MAX_TILES (include/igs_raster.h)


def _lib():
    lib = C.CDLL(str(ROOT / "igs_b200" / "lib" / "libigs_raster.so"))
    lib.igs_last_error.restype = C.c_char_p
    return lib


def _inputs(dev, V, H, W, coord, seed):
    g = torch.Generator().manual_seed(seed)
    c = 3 if coord else 1
    m0 = (2.0 + torch.rand(V, c, H, W, generator=g)).to(dev).contiguous()
    m1 = (2.0 + torch.rand(V, c, H, W, generator=g)).to(dev).contiguous()
    n = torch.nn.functional.normalize(torch.randn(V, 3, H, W, generator=g), dim=1).to(dev).contiguous()
    return m0, m1, n


def _call(lib, which, coord, m0, m1, n, keep):
    V, _, H, W = n.shape
    dev = n.device
    loss = torch.full((1,), float("nan"), device=dev)
    g0, g1, gn = torch.full_like(m0, float("nan")), torch.full_like(m1, float("nan")), torch.full_like(n, float("nan"))
    ptr = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    tfx, tfy = math.tan(0.5), math.tan(0.4)
    common = (C.c_int(1 if coord else 0), ptr(m0), ptr(m1), ptr(n), C.c_int(V), C.c_int(H), C.c_int(W), C.c_float(tfx),
              C.c_float(tfy), C.c_float(0.6), C.c_float(0.5), ptr(loss), ptr(g0), ptr(g1), ptr(gn))
    if which == "noalloc":
        rc = lib.igs_depth_normal_loss_noalloc(*common, stream)
    else:
        def alloc(user, nbytes):
            keep["t"] = torch.empty(int(nbytes) + 256, dtype=torch.uint8, device=dev)
            return (keep["t"].data_ptr() + 255) & ~255
        keep["cb"] = ALLOC(alloc)
        rc = lib.igs_depth_normal_loss(*common, keep["cb"], None, stream)
    torch.cuda.synchronize()
    return rc, loss, g0, g1, gn


def test_noalloc_entry_point_matches_the_allocating_one_twice_in_a_row(cuda_dev, native):
    lib = _lib()
    assert hasattr(lib, "igs_depth_normal_loss_noalloc")
    # two calls in a row with DIFFERENT sizes and inputs: the second starts from the first one's leftovers
    for (V, H, W, coord, seed) in ((3, 200, 333, False, 1), (2, 97, 130, True, 2)):
        m0, m1, n = _inputs(cuda_dev, V, H, W, coord, seed)
        keep = {}
        rc_a, la, a0, a1, an = _call(lib, "alloc", coord, m0, m1, n, keep)
        rc_b, lb, b0, b1, bn = _call(lib, "noalloc", coord, m0, m1, n, keep)
        assert rc_a == 0 and rc_b == 0, lib.igs_last_error()
        assert bool(torch.isfinite(lb).all()) and bool(torch.isfinite(b0).all()) and bool(torch.isfinite(bn).all())
        # same kernels, same arithmetic, same reduction order: the tolerance of the suite's depth-normal tests is met
        # with room to spare (the results are in fact identical)
        assert abs(float(la) - float(lb)) <= 1e-5 * abs(float(la))
        for x, y in ((a0, b0), (a1, b1), (an, bn)):
            assert float((x - y).abs().max()) <= 1e-3 * float(x.abs().max())
        assert torch.equal(la, lb) and torch.equal(a0, b0) and torch.equal(a1, b1) and torch.equal(an, bn)


def test_noalloc_entry_point_rejects_sizes_beyond_its_static_scratch(cuda_dev, native):
    lib = _lib()
    # V * ceil(W/32) * ceil(H/8) = 65 * 32 * 32 = 66560 > 65536: rejected before anything is touched
    V, H, W = 65, 256, 1024
    assert V * ((W + 31) // 32) * ((H + 7) // 8) > MAX_TILES
    t = torch.zeros(8, device=cuda_dev)
    p = C.c_void_p(t.data_ptr())
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.igs_depth_normal_loss_noalloc(C.c_int(0), p, p, p, C.c_int(V), C.c_int(H), C.c_int(W), C.c_float(0.5),
                                           C.c_float(0.5), C.c_float(0.6), C.c_float(1.0), p, p, p, p, stream)
    assert rc == -3 and b"IGS_NOALLOC_MAX_TILES" in lib.igs_last_error()     # IGS_ERR_UNSUPPORTED
    rc = lib.igs_depth_normal_loss_noalloc(C.c_int(0), None, p, p, C.c_int(1), C.c_int(8), C.c_int(32), C.c_float(0.5),
                                           C.c_float(0.5), C.c_float(0.6), C.c_float(1.0), p, p, p, p, stream)
    assert rc == -1                                                           # IGS_ERR_INVALID: null pointer
    torch.cuda.synchronize()
