"""Gaussian-scene PLY: byte layout as scene/gaussian_model.py:167-246 writes it with plyfile
(header text, property order, channel-major SH), round trip, and the reference's load-side checks."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rain_b200 import ply  # noqa: E402


def leaves(P=23, M=16, seed=0):
    g = torch.Generator().manual_seed(seed)
    return dict(xyz=torch.randn(P, 3, generator=g), f_dc=torch.randn(P, 1, 3, generator=g), f_rest=torch.randn(P, M - 1, 3, generator=g),
                opacity=torch.randn(P, 1, generator=g), scaling=torch.randn(P, 3, generator=g), rotation=torch.randn(P, 4, generator=g))


def test_header_and_byte_layout_match_the_reference_writer(tmp_path):
    L = leaves()
    path = str(tmp_path / "point_cloud" / "iteration_7" / "point_cloud.ply")
    ply.save_ply(path, **L)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().splitlines()
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 23"]
    props = [l.split()[-1] for l in lines[3:]]
    assert all(l.startswith("property float ") for l in lines[3:])
    assert props == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"] + [f"f_rest_{i}" for i in range(45)] + ["opacity", "scale_0", "scale_1", "scale_2",
                                                                                                "rot_0", "rot_1", "rot_2", "rot_3"]
    rows = np.frombuffer(body, dtype="<f4").reshape(23, 62)
    assert np.array_equal(rows[:, 0:3], L["xyz"].numpy()) and not rows[:, 3:6].any()
    # channel-major SH: f_rest_k holds channel k // 15, coefficient 1 + k % 15 (transpose(1, 2).flatten, :187-188)
    assert np.array_equal(rows[:, 9 + 0], L["f_rest"][:, 0, 0].numpy()) and np.array_equal(rows[:, 9 + 1], L["f_rest"][:, 1, 0].numpy())
    assert np.array_equal(rows[:, 9 + 15], L["f_rest"][:, 0, 1].numpy()) and np.array_equal(rows[:, 9 + 44], L["f_rest"][:, 14, 2].numpy())
    assert np.array_equal(rows[:, 54], L["opacity"][:, 0].numpy()) and np.array_equal(rows[:, 58:62], L["rotation"].numpy())


@pytest.mark.parametrize("P,deg", [(23, 3), (1, 3), (0, 3), (5, 0), (9, 1)])
def test_round_trip(tmp_path, P, deg):
    L = leaves(P, (deg + 1) ** 2, seed=P)
    path = str(tmp_path / "s.ply")
    ply.save_ply(path, **L)
    R = ply.load_ply(path, max_sh_degree=deg)
    for k in L:
        assert R[k].dtype == torch.float32 and R[k].shape == L[k].shape and torch.equal(R[k], L[k]), k


def test_reader_accepts_shuffled_property_order_ascii_and_rejects_bad_files(tmp_path):
    L = leaves(4, 16, 3)
    path = str(tmp_path / "a.ply")
    ply.save_ply(path, **L)
    v = ply.read_vertex_table(path)
    # rewrite with shuffled property order and a comment, as ascii
    names = list(v.dtype.names)[::-1]
    with open(str(tmp_path / "b.ply"), "w") as f:
        f.write("ply\nformat ascii 1.0\ncomment shuffled\nelement vertex 4\n" + "".join(f"property float {n}\n" for n in names) + "end_header\n")
        for row in v:
            f.write(" ".join(repr(float(row[n])) for n in names) + "\n")
    R = ply.load_ply(str(tmp_path / "b.ply"))
    assert all(torch.equal(R[k], L[k]) for k in L)
    with pytest.raises(ValueError):
        ply.load_ply(path, max_sh_degree=2)          # 45 f_rest properties but degree 2 expects 24 (the reference asserts, :218)
    open(str(tmp_path / "c.ply"), "wb").write(open(path, "rb").read()[:-10])
    with pytest.raises(ValueError):
        ply.load_ply(str(tmp_path / "c.ply"))
    open(str(tmp_path / "d.ply"), "w").write("plx\n")
    with pytest.raises(ValueError):
        ply.load_ply(str(tmp_path / "d.ply"))
