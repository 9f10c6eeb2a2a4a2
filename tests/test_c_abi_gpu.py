"""GPU: the C-ABI called directly through ctypes with raw device pointers (torch only owns the memory),
as a cgo/JNI/ctypes binding of the reference would: forward + backward on the `small` golden case."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest
import torch

import cases
from tolerances import GRAD_RTOL, MAP_ATOL, grad_err

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
ALLOC = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)


class Fwd(C.Structure):
    _fields_ = ([(n, C.c_int) for n in ("P", "D", "M", "V", "width", "height")] +
                [(n, C.c_void_p) for n in ("background", "means3D", "shs", "colors_precomp", "opacities", "scales")] +
                [("scale_modifier", C.c_float)] +
                [(n, C.c_void_p) for n in ("rotations", "cov3D_precomp", "viewmatrix", "projmatrix", "cam_pos")] +
                [(n, C.c_float) for n in ("tan_fovx", "tan_fovy", "kernel_size")] +
                [(n, C.c_int) for n in ("prefiltered", "require_coord", "require_depth", "debug")] +
                [(n, C.c_void_p) for n in ("out_color", "out_coord", "out_mcoord", "out_depth", "out_mdepth", "out_alpha",
                                           "out_normal", "radii", "tile_rows")] +
                [("instance_capacity", C.c_int), ("status", C.c_void_p)])


class Bwd(C.Structure):
    _fields_ = ([(n, C.c_int) for n in ("P", "D", "M", "V", "R", "width", "height")] +
                [(n, C.c_void_p) for n in ("background", "means3D", "shs", "colors_precomp", "scales")] +
                [("scale_modifier", C.c_float)] +
                [(n, C.c_void_p) for n in ("rotations", "cov3D_precomp", "viewmatrix", "projmatrix", "cam_pos")] +
                [(n, C.c_float) for n in ("tan_fovx", "tan_fovy", "kernel_size")] +
                [(n, C.c_void_p) for n in ("radii", "alphas", "normalmap", "geom_buffer", "binning_buffer", "image_buffer",
                                           "dL_dout_color", "dL_dout_coord", "dL_dout_mcoord", "dL_dout_depth", "dL_dout_mdepth",
                                           "dL_dout_alpha", "dL_dout_normal", "dL_dmeans2D", "dL_dcolors", "dL_dopacity",
                                           "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations")] +
                [(n, C.c_int) for n in ("require_coord", "require_depth", "debug")])


def test_forward_backward_through_ctypes(cuda_dev, native):
    lib = C.CDLL(str(ROOT / "igs_b200" / "lib" / "libigs_raster.so"))
    lib.igs_last_error.restype = C.c_char_p
    g = cases.load_golden("small")
    case = cases.case_from_golden(g)
    dev = cuda_dev
    P, W, H = case["P"], case["W"], case["H"]
    T = {k: torch.from_numpy(np.ascontiguousarray(case[k])).to(dev) for k in
         ("bg", "means3D", "shs", "opacities", "scales", "rotations", "viewmatrix", "projmatrix", "campos")}
    outs = {n: torch.zeros(c, H, W, device=dev) for n, c in zip(cases.MAP_NAMES, cases.MAP_CH)}
    radii = torch.zeros(P, dtype=torch.int32, device=dev)
    keep = {}

    def make_alloc(name):
        def alloc(user, nbytes):
            t = torch.empty(int(nbytes) + 256, dtype=torch.uint8, device=dev)
            keep[name] = t
            return (t.data_ptr() + 255) & ~255
        return ALLOC(alloc)

    a_geom, a_bin, a_img, a_tmp = make_alloc("geom"), make_alloc("bin"), make_alloc("img"), make_alloc("tmp")
    f = Fwd(P=P, D=3, M=16, V=1, width=W, height=H, background=T["bg"].data_ptr(), means3D=T["means3D"].data_ptr(),
            shs=T["shs"].data_ptr(), colors_precomp=None, opacities=T["opacities"].data_ptr(), scales=T["scales"].data_ptr(),
            scale_modifier=1.0, rotations=T["rotations"].data_ptr(), cov3D_precomp=None, viewmatrix=T["viewmatrix"].data_ptr(),
            projmatrix=T["projmatrix"].data_ptr(), cam_pos=T["campos"].data_ptr(), tan_fovx=case["tanfovx"],
            tan_fovy=case["tanfovy"], kernel_size=0.0, prefiltered=0, require_coord=1, require_depth=1, debug=1,
            out_color=outs["color"].data_ptr(), out_coord=outs["coord"].data_ptr(), out_mcoord=outs["mcoord"].data_ptr(),
            out_depth=outs["depth"].data_ptr(), out_mdepth=outs["mdepth"].data_ptr(), out_alpha=outs["alpha"].data_ptr(),
            out_normal=outs["normal"].data_ptr(), radii=radii.data_ptr())
    R = C.c_int(0)
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.igs_rasterize_forward(C.byref(f), a_geom, None, a_bin, None, a_img, None, C.c_void_p(stream), C.byref(R))
    assert rc == 0, lib.igs_last_error()
    torch.cuda.synchronize()
    assert R.value == int(g["R"])
    np.testing.assert_array_equal(radii.cpu().numpy(), g["radii"])
    for n in cases.MAP_NAMES:
        assert np.abs(outs[n].cpu().numpy() - g["map_" + n]).max() <= MAP_ATOL, n
    geom, binning, img = keep["geom"], keep["bin"], keep["img"]
    al = lambda t: (t.data_ptr() + 255) & ~255  # noqa: E731
    cot = torch.from_numpy(cases.cotangents(case)[0]).to(dev)
    G = dict(m2=torch.empty(P, 3, device=dev), col=torch.empty(P, 3, device=dev), op=torch.empty(P, device=dev),
             m3=torch.empty(P, 3, device=dev), cov=torch.empty(P, 6, device=dev), sh=torch.empty(P, 16, 3, device=dev),
             sc=torch.empty(P, 3, device=dev), rot=torch.empty(P, 4, device=dev))
    b = Bwd(P=P, D=3, M=16, V=1, R=R.value, width=W, height=H, background=T["bg"].data_ptr(), means3D=T["means3D"].data_ptr(),
            shs=T["shs"].data_ptr(), colors_precomp=None, scales=T["scales"].data_ptr(), scale_modifier=1.0,
            rotations=T["rotations"].data_ptr(), cov3D_precomp=None, viewmatrix=T["viewmatrix"].data_ptr(),
            projmatrix=T["projmatrix"].data_ptr(), cam_pos=T["campos"].data_ptr(), tan_fovx=case["tanfovx"],
            tan_fovy=case["tanfovy"], kernel_size=0.0, radii=radii.data_ptr(), alphas=outs["alpha"].data_ptr(),
            normalmap=outs["normal"].data_ptr(), geom_buffer=al(geom), binning_buffer=al(binning), image_buffer=al(img),
            dL_dout_color=cot.data_ptr(), dL_dmeans2D=G["m2"].data_ptr(), dL_dcolors=G["col"].data_ptr(),
            dL_dopacity=G["op"].data_ptr(), dL_dmeans3D=G["m3"].data_ptr(), dL_dcov3D=G["cov"].data_ptr(),
            dL_dsh=G["sh"].data_ptr(), dL_dscales=G["sc"].data_ptr(), dL_drotations=G["rot"].data_ptr(), require_coord=1,
            require_depth=1, debug=1)
    rc = lib.igs_rasterize_backward(C.byref(b), a_tmp, None, C.c_void_p(stream))
    assert rc == 0, lib.igs_last_error()
    torch.cuda.synchronize()
    for k, t in (("means3D", G["m3"]), ("means2D", G["m2"]), ("sh", G["sh"]), ("opacities", G["op"]), ("scales", G["sc"])):
        ref = g["grad_color_" + k]
        assert grad_err(t.cpu().numpy().reshape(ref.shape), ref) <= GRAD_RTOL, k
    assert lib.igs_kernel_launches() > 0
    # mismatching buffers are rejected, not read through
    b.R = R.value + 1
    assert lib.igs_rasterize_backward(C.byref(b), a_tmp, None, C.c_void_p(stream)) == -1
    b.R = R.value
    b.binning_buffer, b.image_buffer = al(img), al(binning)   # swapped
    assert lib.igs_rasterize_backward(C.byref(b), a_tmp, None, C.c_void_p(stream)) == -1
    assert b"does not belong" in lib.igs_last_error()


def test_depth_normal_loss_through_ctypes(cuda_dev, native):
    """igs_depth_normal_loss with raw device pointers against the golden produced by the reference's own helpers."""
    import math
    lib = C.CDLL(str(ROOT / "igs_b200" / "lib" / "libigs_raster.so"))
    lib.igs_last_error.restype = C.c_char_p
    g = dict(np.load(cases.GOLDEN_DIR / "refmath_frame0.npz"))
    dev = cuda_dev
    d1 = torch.from_numpy(g["a_depth1"]).to(dev).contiguous()
    d2 = torch.from_numpy(g["a_depth2"]).to(dev).contiguous()
    nrm = torch.from_numpy(g["a_normal"]).to(dev).contiguous()
    H, W = nrm.shape[1:]
    tfx, tfy = (math.tan(float(f) / 2) for f in g["a_fov"])
    loss = torch.zeros(1, device=dev)
    g1, g2, gn = torch.empty_like(d1), torch.empty_like(d2), torch.empty_like(nrm)
    keep = {}

    def alloc(user, nbytes):
        keep["t"] = torch.empty(int(nbytes) + 256, dtype=torch.uint8, device=dev)
        return (keep["t"].data_ptr() + 255) & ~255
    ptr = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.igs_depth_normal_loss(C.c_int(0), ptr(d1), ptr(d2), ptr(nrm), C.c_int(1), C.c_int(H), C.c_int(W), C.c_float(tfx),
                                   C.c_float(tfy), C.c_float(0.6), C.c_float(1.0), ptr(loss), ptr(g1), ptr(g2), ptr(gn),
                                   ALLOC(alloc), None, stream)
    assert rc == 0, lib.igs_last_error()
    torch.cuda.synchronize()
    assert abs(float(loss) - float(g["a_loss"])) <= 2e-6
    for got, k in ((g1, "a_d_depth1"), (g2, "a_d_depth2"), (gn, "a_d_normal")):
        assert np.abs(got.cpu().numpy() - g[k]).max() <= 1e-3 * np.abs(g[k]).max(), k
    # argument checks: no allocator / null pointers -> IGS_ERR_INVALID, nothing launched
    assert lib.igs_depth_normal_loss(C.c_int(0), ptr(d1), ptr(d2), ptr(nrm), C.c_int(1), C.c_int(H), C.c_int(W), C.c_float(tfx),
                                     C.c_float(tfy), C.c_float(0.6), C.c_float(1.0), ptr(loss), ptr(g1), ptr(g2), ptr(gn),
                                     None, None, stream) == -1
    assert lib.igs_depth_normal_loss(C.c_int(0), None, ptr(d2), ptr(nrm), C.c_int(1), C.c_int(H), C.c_int(W), C.c_float(tfx),
                                     C.c_float(tfy), C.c_float(0.6), C.c_float(1.0), ptr(loss), ptr(g1), ptr(g2), ptr(gn),
                                     ALLOC(alloc), None, stream) == -1
