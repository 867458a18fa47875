"""On-disk format of a Gaussian scene: the binary little-endian PLY the reference writes and reads
with the `plyfile` package (scene/gaussian_model.py:167-246) — `plyfile` is not part of this image, and the format needs
no general PLY parser: one `vertex` element of float32 properties
    x y z nx ny nz f_dc_0..2 f_rest_0..(3(M-1)-1) opacity scale_0..2 rot_0..3
with the SH coefficients stored CHANNEL-MAJOR (save_ply transposes [P,M-1,3] -> [P,3,M-1] before flattening, :187-188;
load_ply reshapes to (P, 3, M-1) and transposes back, :222,241).  Host-side numpy only; tensors come back on the CPU."""
import os

import numpy as np
import torch

_PLY_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1", "char": "i1", "int8": "i1",
              "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2", "int": "<i4", "int32": "<i4", "uint": "<u4", "uint32": "<u4"}


def attribute_names(M):
    """construct_list_of_attributes (scene/gaussian_model.py:167-179) for M SH coefficients per channel."""
    return (["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] + [f"f_rest_{i}" for i in range(3 * (M - 1))] + ["opacity"] +
            [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)])


def save_ply(path, xyz, f_dc, f_rest, opacity, scaling, rotation):
    """GaussianModel.save_ply (:181-199): raw (pre-activation) leaves, normals written as zeros."""
    t = lambda a: a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    xyz, f_dc, f_rest, opacity, scaling, rotation = (t(a).astype(np.float32) for a in (xyz, f_dc, f_rest, opacity, scaling, rotation))
    P = xyz.shape[0]
    M = 1 + f_rest.shape[1]
    cols = np.concatenate([xyz, np.zeros_like(xyz), f_dc.transpose(0, 2, 1).reshape(P, 3), f_rest.transpose(0, 2, 1).reshape(P, 3 * (M - 1)),
                           opacity.reshape(P, 1), scaling, rotation], axis=1).astype("<f4")
    names = attribute_names(M)
    assert cols.shape[1] == len(names)
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P + "".join(f"property float {n}\n" for n in names) + "end_header\n"
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)                 # mkdir_p (:182)
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(np.ascontiguousarray(cols).tobytes())


def read_vertex_table(path):
    """The `vertex` element of a binary little-endian (or ascii) PLY as a numpy structured array."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, elements, cur = None, [], None
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: header has no end_header")
            tok = line.decode("ascii").split()
            if not tok or tok[0] in ("comment", "obj_info"):
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                cur = dict(name=tok[1], count=int(tok[2]), props=[])
                elements.append(cur)
            elif tok[0] == "property":
                if tok[1] == "list":
                    raise ValueError(f"{path}: list properties are not part of the Gaussian-scene format")
                if tok[1] not in _PLY_TYPES:
                    raise ValueError(f"{path}: unknown property type {tok[1]}")
                cur["props"].append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if not elements or elements[0]["name"] != "vertex":
            raise ValueError(f"{path}: the first element must be `vertex` (plydata.elements[0], :203)")
        el = elements[0]
        if fmt == "binary_little_endian":
            dt = np.dtype(el["props"])
            data = np.frombuffer(f.read(dt.itemsize * el["count"]), dtype=dt, count=el["count"])
            if data.shape[0] != el["count"]:
                raise ValueError(f"{path}: truncated vertex data")
            return data
        if fmt == "ascii":
            dt = np.dtype([(n, t.lstrip("<")) for n, t in el["props"]])
            rows = [tuple(f.readline().split()) for _ in range(el["count"])]
            return np.array(rows, dtype=dt)
        raise ValueError(f"{path}: unsupported PLY format {fmt}")


def load_ply(path, max_sh_degree=3):
    """GaussianModel.load_ply (:206-246): returns dict(xyz[P,3], f_dc[P,1,3], f_rest[P,M-1,3], opacity[P,1], scaling[P,3],
    rotation[P,4]) float32 CPU tensors (raw leaves)."""
    v = read_vertex_table(path)
    names = v.dtype.names
    P = v.shape[0]
    col = lambda n: np.asarray(v[n], dtype=np.float64)
    xyz = np.stack((col("x"), col("y"), col("z")), axis=1)
    opacities = col("opacity")[..., np.newaxis]
    features_dc = np.zeros((P, 3, 1))
    for c in range(3):
        features_dc[:, c, 0] = col(f"f_dc_{c}")
    by_index = lambda prefix: sorted([n for n in names if n.startswith(prefix)], key=lambda x: int(x.split("_")[-1]))
    extra = by_index("f_rest_")
    if len(extra) != 3 * (max_sh_degree + 1) ** 2 - 3:                                   # the reference asserts (:218)
        raise ValueError(f"{path}: {len(extra)} f_rest properties, expected {3 * (max_sh_degree + 1) ** 2 - 3} for SH degree {max_sh_degree}")
    features_extra = np.stack([col(n) for n in extra], axis=1).reshape(P, 3, (max_sh_degree + 1) ** 2 - 1) if extra else np.zeros((P, 3, 0))
    scales = np.stack([col(n) for n in by_index("scale_")], axis=1)
    rots = np.stack([col(n) for n in by_index("rot")], axis=1)
    f32 = lambda a: torch.tensor(a, dtype=torch.float)
    return dict(xyz=f32(xyz), f_dc=f32(features_dc).transpose(1, 2).contiguous(), f_rest=f32(features_extra).transpose(1, 2).contiguous(),
                opacity=f32(opacities), scaling=f32(scales), rotation=f32(rots))
