"""CPU: `.ply` I/O (igs_b200/ply.py) against a known-answer file laid out by hand with struct.pack in the reference's
property order (igs/models/gaussian_model.py:163-175, :207-212) and the reference loader's conventions
(igs/models/gs.py:400-460): channel-major f_rest, filter_3D folded into scale / opacity."""
import struct

import numpy as np
import pytest

torch = pytest.importorskip("torch")


def _mod():
    try:
        from igs_b200 import ply
    except ImportError as ex:
        pytest.skip(f"igs_b200 not importable: {ex}")
    return ply


def _hand_written(tmp_path, with_filter):
    """Two points; value of property k of point i is 100*i + k, so every column is identifiable."""
    names = ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"] + [f"f_rest_{i}" for i in range(45)]
    names += ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"]
    if with_filter:
        names.append("filter_3D")
    hdr = "ply\nformat binary_little_endian 1.0\nelement vertex 2\n" + "".join(f"property float {n}\n" for n in names) + "end_header\n"
    body = b""
    for i in range(2):
        vals = [100.0 * i + k for k in range(len(names))]
        if with_filter:
            vals[names.index("opacity")] = 0.3 - i          # logits
            for a in range(3):
                vals[names.index(f"scale_{a}")] = -2.0 + 0.1 * a + i
            vals[-1] = 0.05
        body += struct.pack("<%df" % len(names), *vals)
    p = tmp_path / ("kat_f.ply" if with_filter else "kat.ply")
    p.write_bytes(hdr.encode("ascii") + body)
    return p, names


def test_load_known_answer_file(tmp_path):
    ply = _mod()
    p, names = _hand_written(tmp_path, False)
    g = ply.load_ply(p)
    col = {n: k for k, n in enumerate(names)}
    for i in range(2):
        b = 100.0 * i
        assert g["xyz"][i].tolist() == [b + col["x"], b + col["y"], b + col["z"]]
        assert g["shs"].shape == (2, 16, 3)
        assert g["shs"][i, 0].tolist() == [b + col["f_dc_0"], b + col["f_dc_1"], b + col["f_dc_2"]]
        # f_rest is channel-major: f_rest_k -> channel k // 15, coefficient 1 + k % 15
        for k in (0, 1, 14, 15, 29, 30, 44):
            assert float(g["shs"][i, 1 + k % 15, k // 15]) == b + col[f"f_rest_{k}"]
        assert float(g["opacity"][i, 0]) == b + col["opacity"]
        assert g["scaling"][i].tolist() == [b + col[f"scale_{a}"] for a in range(3)]
        assert g["rotation"][i].tolist() == [b + col[f"rot_{a}"] for a in range(4)]
    assert g["filter_3D"] is None


def test_filter_3d_is_folded_like_the_reference_loader(tmp_path):
    ply = _mod()
    p, names = _hand_written(tmp_path, True)
    g = ply.load_ply(p)
    raw = ply.load_ply(p, apply_filter_3D=False)
    # gs.py:452-457 restated
    s2 = torch.exp(raw["scaling"]) ** 2
    s2f = s2 + raw["filter_3D"] ** 2
    coef = torch.sqrt(s2.prod(dim=1) / s2f.prod(dim=1))
    o = torch.sigmoid(raw["opacity"]) * coef[:, None]
    assert torch.allclose(g["scaling"], torch.log(torch.sqrt(s2f)), atol=1e-6)
    assert torch.allclose(g["opacity"], torch.log(o / (1 - o)), atol=1e-5)


def test_save_load_round_trip_is_bit_exact(tmp_path):
    ply = _mod()
    g = torch.Generator().manual_seed(0)
    P = 257
    xyz, shs = torch.randn(P, 3, generator=g), torch.randn(P, 16, 3, generator=g)
    op, sc, rot = torch.randn(P, 1, generator=g), torch.randn(P, 3, generator=g), torch.randn(P, 4, generator=g)
    ply.save_ply(tmp_path / "sub" / "a.ply", xyz, shs, op, sc, rot)
    got = ply.load_ply(tmp_path / "sub" / "a.ply")
    for a, b in ((xyz, got["xyz"]), (shs, got["shs"]), (op, got["opacity"]), (sc, got["scaling"]), (rot, got["rotation"])):
        assert torch.equal(a, b)
    # header is the reference's attribute list, in order
    head = (tmp_path / "sub" / "a.ply").read_bytes().split(b"end_header\n")[0].decode()
    props = [l.split()[-1] for l in head.splitlines() if l.startswith("property")]
    assert props == ply.attribute_names(16) and props[:9] == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"]
    assert len(props) == 62 and props[54] == "opacity"


def test_truncated_and_foreign_files_are_rejected(tmp_path):
    ply = _mod()
    p, _ = _hand_written(tmp_path, False)
    data = p.read_bytes()
    (tmp_path / "t.ply").write_bytes(data[:-8])
    with pytest.raises(ValueError):
        ply.load_ply(tmp_path / "t.ply")
    (tmp_path / "x.ply").write_bytes(b"solid\n")
    with pytest.raises(ValueError):
        ply.load_ply(tmp_path / "x.ply")
