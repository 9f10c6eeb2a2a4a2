"""ctypes front-end of the CPU oracle (oracle/raster_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Nothing under igs_b200/ may import this module.

forward()/backward() orchestrate the C stages exactly as the reference's Rasterizer::forward /
Rasterizer::backward do (DGR/cuda_rasterizer/rasterizer_impl.cu:254-425, 429-571) for one view.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
_LIB = None


def build(force: bool = False) -> Path:
    so = HERE / "liboracle.so"
    src = HERE / "raster_oracle.c"
    if force or not so.exists() or so.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(HERE), "-B"], check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(str(build()))
        _LIB.oracle_binning.restype = C.c_int64
        _LIB.oracle_num_threads.restype = C.c_int
    return _LIB


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def set_num_threads(n: int) -> int:
    """OpenMP threads for the following oracle calls (returns the count in effect; 1 without OpenMP)."""
    lib().oracle_set_num_threads.restype = C.c_int
    return int(lib().oracle_set_num_threads(C.c_int(int(n))))


class _Geom(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("depths", "means2D", "conic_opacity", "rgb", "cov3D", "view_points",
                                          "camera_planes", "ray_planes", "ts", "normals", "clamped", "radii",
                                          "tiles_touched")]


class _Image(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("color", "coord", "mcoord", "depth", "mdepth", "alpha", "normal", "n_contrib",
                                          "accum_coord", "accum_depth", "normal_length")]


class _Grads(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("dL_dmean2D", "dL_dconic", "dL_dopacity", "dL_dcolors", "dL_dview_points",
                                          "dL_dts", "dL_dcamera_planes", "dL_dray_planes", "dL_dnormals")]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def forward(means3D, scales, rotations, opacities, shs, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy,
            sh_degree=3, bg=None, scale_modifier=1.0, kernel_size=0.0, colors_precomp=None, cov3D_precomp=None,
            require_coord=True, require_depth=True):
    L = lib()
    means3D, scales, rotations, shs = _f32(means3D), _f32(scales), _f32(rotations), _f32(shs)
    opacities = _f32(opacities).reshape(-1)
    colors_precomp, cov3D_precomp = _f32(colors_precomp), _f32(cov3D_precomp)
    viewmatrix, projmatrix, campos = _f32(viewmatrix).reshape(-1), _f32(projmatrix).reshape(-1), _f32(campos)
    bg = np.zeros(3, np.float32) if bg is None else _f32(bg)
    P = means3D.shape[0]
    M = 0 if shs is None else shs.shape[1]
    N = W * H
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    g = dict(depths=np.zeros(P, np.float32), means2D=np.zeros((P, 2), np.float32), conic_opacity=np.zeros((P, 4), np.float32),
             rgb=np.zeros((P, 3), np.float32), cov3D=np.zeros((P, 6), np.float32), view_points=np.zeros((P, 3), np.float32),
             camera_planes=np.zeros((P, 6), np.float32), ray_planes=np.zeros((P, 2), np.float32), ts=np.zeros(P, np.float32),
             normals=np.zeros((P, 3), np.float32), clamped=np.zeros((P, 3), np.uint8), radii=np.zeros(P, np.int32),
             tiles_touched=np.zeros(P, np.uint32))
    gs = _Geom(**{k: _p(v) for k, v in g.items()})
    L.oracle_preprocess(C.c_int(P), C.c_int(sh_degree), C.c_int(M), _p(means3D), _p(scales), C.c_float(scale_modifier),
                        _p(rotations), _p(opacities), _p(shs), _p(cov3D_precomp), _p(colors_precomp), _p(viewmatrix),
                        _p(projmatrix), _p(campos), C.c_int(W), C.c_int(H), C.c_float(tanfovx), C.c_float(tanfovy),
                        C.c_float(kernel_size), C.byref(gs))
    Rmax = int(g["tiles_touched"].sum())
    keys = np.zeros(max(Rmax, 1), np.uint64)
    plist = np.zeros(max(Rmax, 1), np.uint32)
    ranges = np.zeros((tiles, 2), np.uint32)
    R = int(L.oracle_binning(C.c_int(P), C.c_int(W), C.c_int(H), C.byref(gs), _p(keys), _p(plist), _p(ranges)))
    assert R == Rmax
    im = dict(color=np.zeros((3, H, W), np.float32), coord=np.zeros((3, H, W), np.float32), mcoord=np.zeros((3, H, W), np.float32),
              depth=np.zeros((1, H, W), np.float32), mdepth=np.zeros((1, H, W), np.float32), alpha=np.zeros((1, H, W), np.float32),
              normal=np.zeros((3, H, W), np.float32), n_contrib=np.zeros((2, N), np.uint32),
              accum_coord=np.zeros((3, N), np.float32), accum_depth=np.zeros(N, np.float32), normal_length=np.zeros(N, np.float32))
    ims = _Image(**{k: _p(v) for k, v in im.items()})
    colors = colors_precomp if colors_precomp is not None else g["rgb"]
    L.oracle_blend_forward(C.c_int(W), C.c_int(H), C.byref(gs), _p(colors), _p(plist), _p(ranges), _p(bg),
                           C.c_float(tanfovx), C.c_float(tanfovy), C.c_int(int(require_coord)), C.c_int(int(require_depth)),
                           C.byref(ims))
    return dict(P=P, M=M, W=W, H=H, R=R, geom=g, image=im, keys=keys[:R], point_list=plist[:R], ranges=ranges, colors=colors,
                inputs=dict(means3D=means3D, scales=scales, rotations=rotations, opacities=opacities, shs=shs,
                            viewmatrix=viewmatrix, projmatrix=projmatrix, campos=campos, bg=bg, tanfovx=tanfovx,
                            tanfovy=tanfovy, sh_degree=sh_degree, scale_modifier=scale_modifier, kernel_size=kernel_size,
                            colors_precomp=colors_precomp, cov3D_precomp=cov3D_precomp, require_coord=require_coord,
                            require_depth=require_depth),
                _keep=(gs, ims))


def backward(fw, dL_color=None, dL_coord=None, dL_mcoord=None, dL_depth=None, dL_mdepth=None, dL_alpha=None,
             dL_normal=None, ref_bug=True):
    """Gradients in the reference's names (rasterize_points.cu:245): means2D [P,3], colors [P,3], opacity [P,1],
    means3D [P,3], cov3D [P,6], sh [P,M,3], scales [P,3], rotations [P,4]."""
    L = lib()
    i = fw["inputs"]
    P, W, H, M = fw["P"], fw["W"], fw["H"], fw["M"]
    g, im = fw["geom"], fw["image"]
    gs = _Geom(**{k: _p(v) for k, v in g.items()})
    ims = _Image(**{k: _p(v) for k, v in im.items()})
    cots = [_f32(x) for x in (dL_color, dL_coord, dL_mcoord, dL_depth, dL_mdepth, dL_alpha, dL_normal)]
    if cots[0] is None:
        cots[0] = np.zeros((3, H, W), np.float32)
    sg = dict(dL_dmean2D=np.zeros((P, 3)), dL_dconic=np.zeros((P, 4)), dL_dopacity=np.zeros(P), dL_dcolors=np.zeros((P, 3)),
              dL_dview_points=np.zeros((P, 3)), dL_dts=np.zeros(P), dL_dcamera_planes=np.zeros((P, 6)),
              dL_dray_planes=np.zeros((P, 2)), dL_dnormals=np.zeros((P, 3)))
    sgs = _Grads(**{k: _p(v) for k, v in sg.items()})
    L.oracle_blend_backward(C.c_int(W), C.c_int(H), C.byref(gs), _p(fw["colors"]), _p(fw["point_list"] if fw["R"] else np.zeros(1, np.uint32)),
                            _p(fw["ranges"]), _p(i["bg"]), C.c_float(i["tanfovx"]), C.c_float(i["tanfovy"]),
                            C.c_int(int(i["require_coord"])), C.c_int(int(i["require_depth"])), C.byref(ims),
                            *[_p(c) for c in cots], C.byref(sgs))
    f = {k: v.astype(np.float32) for k, v in sg.items()}
    cov3D = i["cov3D_precomp"] if i["cov3D_precomp"] is not None else g["cov3D"]
    dL_dmeans = np.zeros((P, 3), np.float32)
    dL_dcov = np.zeros((P, 6), np.float32)
    dL_dopacity = f["dL_dopacity"].copy()
    L.oracle_cov2d_backward(C.c_int(P), _p(i["means3D"]), _p(g["radii"]), _p(cov3D), C.c_int(W), C.c_int(H),
                            C.c_float(i["tanfovx"]), C.c_float(i["tanfovy"]), C.c_float(i["kernel_size"]), _p(i["viewmatrix"]),
                            _p(f["dL_dconic"]), _p(f["dL_dcamera_planes"]), _p(f["dL_dray_planes"]), _p(f["dL_dnormals"]),
                            _p(g["conic_opacity"]), C.c_int(int(ref_bug)), _p(dL_dmeans), _p(dL_dcov), _p(dL_dopacity))
    dL_dsh = np.zeros((P, max(M, 1), 3), np.float32)
    dL_dscale = np.zeros((P, 3), np.float32)
    dL_drot = np.zeros((P, 4), np.float32)
    L.oracle_preprocess_backward(C.c_int(P), C.c_int(i["sh_degree"]), C.c_int(M), _p(i["means3D"]), _p(g["radii"]), _p(i["shs"]),
                                 _p(g["clamped"]), _p(i["scales"]), _p(i["rotations"]), C.c_float(i["scale_modifier"]),
                                 _p(i["viewmatrix"]), _p(i["projmatrix"]), _p(i["campos"]), _p(f["dL_dmean2D"]),
                                 _p(f["dL_dview_points"]), _p(f["dL_dts"]), _p(f["dL_dcolors"]), _p(dL_dcov), _p(dL_dmeans),
                                 _p(dL_dsh), _p(dL_dscale), _p(dL_drot))
    return dict(means2D=f["dL_dmean2D"], colors_precomp=f["dL_dcolors"], opacities=dL_dopacity.reshape(P, 1),
                means3D=dL_dmeans, cov3D_precomp=dL_dcov, sh=dL_dsh[:, :M] if M else dL_dsh[:, :0], scales=dL_dscale,
                rotations=dL_drot)


def forward_case(case: dict):
    """Runs forward() on a tests/cases.py-style dict."""
    return forward(case["means3D"], case.get("scales"), case.get("rotations"), case["opacities"], case.get("shs"),
                   case["viewmatrix"], case["projmatrix"], case["campos"], case["W"], case["H"], case["tanfovx"],
                   case["tanfovy"], sh_degree=case["sh_degree"], bg=case["bg"], scale_modifier=case["scale_modifier"],
                   kernel_size=case["kernel_size"], colors_precomp=case.get("colors_precomp"),
                   cov3D_precomp=case.get("cov3D_precomp"), require_coord=case["require_coord"],
                   require_depth=case["require_depth"])
