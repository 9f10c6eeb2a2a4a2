"""Minimal stand-in for the `plyfile` package (dranjan/python-plyfile; the reference imports `PlyData, PlyElement`,
igs/models/gs.py:25, igs/models/gaussian_model.py:18 — not installed in this image), covering exactly the calls the
reference's save_ply / load_ply make:

    el = PlyElement.describe(structured_array, 'vertex');  PlyData([el]).write(path)
    plydata = PlyData.read(path);  plydata.elements[0]["x"];  [p.name for p in plydata.elements[0].properties]

The container is the public PLY format (binary_little_endian 1.0, one element of scalar properties), which is what
plyfile writes for a little-endian structured array of 'f4' fields.  Test infrastructure for the golden generators
(tools/make_golden_renderer.py); the product's own reader / writer is igs_b200/ply.py.
"""
from __future__ import annotations

import numpy as np

_TO_PLY = {"f4": "float", "f8": "double", "u1": "uchar", "i1": "char", "i2": "short", "u2": "ushort", "i4": "int", "u4": "uint"}
_FROM_PLY = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1", "char": "i1",
             "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2", "int": "<i4", "int32": "<i4",
             "uint": "<u4", "uint32": "<u4"}


class PlyProperty:
    def __init__(self, name, dtype):
        self.name, self.dtype = name, dtype


class PlyElement:
    def __init__(self, name, data):
        self.name, self.data = name, data
        self.properties = tuple(PlyProperty(n, data.dtype[n]) for n in data.dtype.names)

    @staticmethod
    def describe(data, name):
        return PlyElement(name, np.asarray(data))

    def __getitem__(self, key):
        return self.data[key]


class PlyData:
    def __init__(self, elements):
        self.elements = list(elements)

    def write(self, path):
        with open(path, "wb") as f:
            hdr = "ply\nformat binary_little_endian 1.0\n"
            for el in self.elements:
                hdr += f"element {el.name} {el.data.shape[0]}\n"
                for n in el.data.dtype.names:
                    dt = el.data.dtype[n]
                    hdr += f"property {_TO_PLY[dt.str.lstrip('<=|')]} {n}\n"
            hdr += "end_header\n"
            f.write(hdr.encode("ascii"))
            for el in self.elements:
                le = el.data.astype(el.data.dtype.newbyteorder("<"), copy=False)
                f.write(np.ascontiguousarray(le).tobytes())

    @staticmethod
    def read(path):
        with open(path, "rb") as f:
            assert f.readline().strip() == b"ply"
            elements, fmt = [], None
            while True:
                tok = f.readline().decode("ascii").split()
                if not tok or tok[0] == "comment":
                    continue
                if tok[0] == "format":
                    fmt = tok[1]
                elif tok[0] == "element":
                    elements.append([tok[1], int(tok[2]), []])
                elif tok[0] == "property":
                    elements[-1][2].append((tok[2], _FROM_PLY[tok[1]]))
                elif tok[0] == "end_header":
                    break
            assert fmt == "binary_little_endian"
            out = []
            for name, count, props in elements:
                out.append(PlyElement(name, np.fromfile(f, dtype=np.dtype(props), count=count)))
        return PlyData(out)
