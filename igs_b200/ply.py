"""`.ply` I/O of the streamed Gaussians: the on-disk format either side of the refine path.

Reference: igs/models/gaussian_model.py:163-212 (`construct_list_of_attributes`, `save_ply`: binary little-endian PLY,
one `vertex` element of float32 properties x y z nx ny nz f_dc_0..2 f_rest_0..44 opacity scale_0..2 rot_0..3, values
PRE-activation, SH stored channel-major: f_rest_k = sh[:, 1 + k % 15, k // 15]) and igs/models/gs.py:400-460
(`load_ply`: the frame-0 RaDe-GS model additionally carries `filter_3D`, which is folded into scale / opacity with
`get_scaling_n_opacity_with_3D_filter` so that the streamed model renders with kernel_size 0).

The reference goes through the `plyfile` package (not in this image); the format itself is the public PLY layout, read
and written here with numpy structured arrays (one fromfile / tofile, no per-property Python loop over points).
"""
from __future__ import annotations

from pathlib import Path
from typing import Dict, Optional

import numpy as np
import torch


def attribute_names(n_sh: int = 16, with_filter: bool = False):
    names = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] + [f"f_rest_{i}" for i in range(3 * (n_sh - 1))]
    names += ["opacity"] + [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)]
    if with_filter:
        names.append("filter_3D")
    return names


def save_ply(path, xyz, shs, opacity_logit, log_scale, rot, filter_3D: Optional[torch.Tensor] = None,
             normalize_rotation: bool = False) -> None:
    """xyz [P,3], shs [P,K,3], opacity_logit [P,1], log_scale [P,3], rot [P,4] (any device), written pre-activation.

    The reference has two writers for this layout.  The stream model's `save_ply` (igs/models/gs.py:317-342) writes
    the stored fields as they are: that is the default here.  The refine model's `save_ply(save_type='all')`
    (igs/models/gaussian_model.py:198-212) writes `get_rotation`, i.e. the NORMALISED quaternion (and opacity / scale
    through inverse-activation(activation(.)), equal up to rounding): `normalize_rotation=True` gives that, so files
    are value-compatible with reference-written ones for tools that assume unit quaternions.  Both are pinned by
    tests/test_renderer_ref_host.py against files the reference's own code wrote."""
    if normalize_rotation:
        rot = torch.nn.functional.normalize(rot.detach().float())
    t = lambda a: a.detach().cpu().float().numpy()  # noqa: E731
    xyz, shs = t(xyz), t(shs)
    P, K = shs.shape[0], shs.shape[1]
    f_dc = shs[:, 0:1, :].transpose(0, 2, 1).reshape(P, -1)      # [:,0:1,:].transpose(1,2).flatten(1)
    f_rest = shs[:, 1:, :].transpose(0, 2, 1).reshape(P, -1)
    cols = [xyz, np.zeros_like(xyz), f_dc, f_rest, t(opacity_logit).reshape(P, 1), t(log_scale), t(rot)]
    if filter_3D is not None:
        cols.append(t(filter_3D).reshape(P, 1))
    names = attribute_names(K, filter_3D is not None)
    data = np.ascontiguousarray(np.concatenate(cols, axis=1), dtype="<f4")
    assert data.shape[1] == len(names)
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P
    header += "".join(f"property float {n}\n" for n in names) + "end_header\n"
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        data.tofile(f)


_PLY_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1",
              "char": "i1", "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2",
              "int": "<i4", "int32": "<i4", "uint": "<u4", "uint32": "<u4"}


def read_vertex_table(path) -> Dict[str, np.ndarray]:
    """Named columns of the first element of a binary little-endian (or ascii) PLY file."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("not a PLY file")
        fmt, count, props, in_first, n_elements = None, None, [], False, 0
        while True:
            line = f.readline()
            if not line:
                raise ValueError("PLY header is not terminated")
            tok = line.decode("ascii").split()
            if not tok or tok[0] == "comment":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                n_elements += 1
                in_first = n_elements == 1
                if in_first:
                    count = int(tok[2])
            elif tok[0] == "property" and in_first:
                if tok[1] == "list":
                    raise ValueError("list properties are not supported in the vertex element")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt == "binary_little_endian":
            table = np.fromfile(f, dtype=np.dtype(props), count=count)
        elif fmt == "ascii":
            raw = np.loadtxt(f, max_rows=count, ndmin=2)
            table = np.empty(count, dtype=np.dtype(props))
            for i, (n, _) in enumerate(props):
                table[n] = raw[:, i]
        else:
            raise ValueError(f"unsupported PLY format {fmt}")
    if table.shape[0] != count:
        raise ValueError("PLY file is truncated")
    return {n: table[n] for n, _ in props}


def _numbered(cols, prefix):
    names = sorted((n for n in cols if n.startswith(prefix)), key=lambda x: int(x.split("_")[-1]))
    return names


def load_ply(path, max_sh_degree: int = 3, apply_filter_3D: bool = True):
    """gs.py:400-460.  Returns dict(xyz [P,3], shs [P,K,3], opacity [P,1] (logit), scaling [P,3] (log), rotation [P,4],
    filter_3D [P,1] or None).  With `apply_filter_3D` and a `filter_3D` column, scale and opacity are replaced by
    log / inverse-sigmoid of their 3D-filtered activations, exactly as the reference's loader does."""
    c = read_vertex_table(path)
    P = c["x"].shape[0]
    xyz = np.stack([c["x"], c["y"], c["z"]], axis=1)
    f_dc = np.stack([c["f_dc_0"], c["f_dc_1"], c["f_dc_2"]], axis=1).reshape(P, 3, 1)
    extra = _numbered(c, "f_rest_")
    K = (max_sh_degree + 1) ** 2
    assert len(extra) == 3 * K - 3, (len(extra), K)
    f_rest = np.stack([c[n] for n in extra], axis=1).reshape(P, 3, K - 1) if extra else np.zeros((P, 3, 0))
    scales = np.stack([c[n] for n in _numbered(c, "scale_")], axis=1)
    rots = np.stack([c[n] for n in _numbered(c, "rot")], axis=1)
    shs = torch.cat([torch.tensor(f_dc, dtype=torch.float).transpose(1, 2), torch.tensor(f_rest, dtype=torch.float).transpose(1, 2)],
                    dim=1).contiguous()
    opacity = torch.tensor(c["opacity"][..., None], dtype=torch.float)
    scaling = torch.tensor(scales, dtype=torch.float)
    filt = torch.tensor(c["filter_3D"][..., None], dtype=torch.float) if "filter_3D" in c else None
    if filt is not None and apply_filter_3D:
        from .frame0 import scaling_n_opacity_with_3D_filter
        s_act, o_act = scaling_n_opacity_with_3D_filter(scaling, opacity, filt)
        opacity = torch.log(o_act / (1 - o_act))    # inverse_sigmoid
        scaling = torch.log(s_act)
    return dict(xyz=torch.tensor(xyz, dtype=torch.float), shs=shs, opacity=opacity, scaling=scaling,
                rotation=torch.tensor(rots, dtype=torch.float), filter_3D=filt)
