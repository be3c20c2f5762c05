"""3DGS point-cloud PLY files: the on-disk format on either side of the hot path.

Mirrors hugs/models/scene.py:229-308 (construct_list_of_attributes, save_ply, load_ply), which go
through the `plyfile` package; here the same file is produced and parsed directly with numpy:
  ply / format binary_little_endian 1.0 / element vertex N / property float <name> ... / end_header
followed by N records of 62 float32 (SH degree 3): x y z nx ny nz f_dc_0..2 f_rest_0..44 opacity
scale_0..2 rot_0..3.  f_dc / f_rest are stored channel-major, i.e. features[:, k, c] sits at column
c * K + k (the reference's transpose(1, 2).flatten(start_dim=1), scene.py:247-248).
Values are the raw parameters (log-scales, logit-opacities, unnormalised quaternions).
`load_ply` also returns `shs` [P,(D+1)^2,3], the layout GaussianRasterizer takes.
"""
import os

import numpy as np


def construct_list_of_attributes(n_dc=3, n_rest=45, n_scale=3, n_rot=4):
    """hugs/models/scene.py:229-242."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += ["f_dc_%d" % i for i in range(n_dc)]
    names += ["f_rest_%d" % i for i in range(n_rest)]
    names.append("opacity")
    names += ["scale_%d" % i for i in range(n_scale)]
    names += ["rot_%d" % i for i in range(n_rot)]
    return names


def _np(t):
    if hasattr(t, "detach"):
        t = t.detach().cpu().numpy()
    return np.asarray(t, dtype=np.float32)


def save_ply(path, xyz, features_dc, features_rest, opacity, scaling, rotation):
    """hugs/models/scene.py:244-262.  features_dc [P,1,3], features_rest [P,K-1,3] (the model's layout),
    opacity [P,1], scaling [P,3], rotation [P,4]."""
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    xyz = _np(xyz)
    P = xyz.shape[0]
    f_dc = _np(features_dc).reshape(P, -1, 3).transpose(0, 2, 1).reshape(P, -1)
    f_rest = _np(features_rest).reshape(P, -1, 3).transpose(0, 2, 1).reshape(P, -1)
    cols = [xyz, np.zeros_like(xyz), f_dc, f_rest, _np(opacity).reshape(P, 1), _np(scaling).reshape(P, -1),
            _np(rotation).reshape(P, -1)]
    table = np.ascontiguousarray(np.concatenate(cols, axis=1).astype("<f4"))
    names = construct_list_of_attributes(f_dc.shape[1], f_rest.shape[1], cols[5].shape[1], cols[6].shape[1])
    assert table.shape[1] == len(names)
    header = ["ply", "format binary_little_endian 1.0", "element vertex %d" % P]
    header += ["property float %s" % n for n in names]
    header.append("end_header")
    with open(path, "wb") as f:
        f.write(("\n".join(header) + "\n").encode("ascii"))
        f.write(table.tobytes())


_PLY_TYPES = {"float": "f4", "float32": "f4", "double": "f8", "float64": "f8", "uchar": "u1", "uint8": "u1",
              "char": "i1", "int8": "i1", "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2",
              "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4"}


def read_vertex_table(path):
    """Parses the vertex element of a binary or ascii PLY into {property name: 1-D array}."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("%s is not a PLY file" % path)
        fmt, props, count, in_vertex, seen_element = None, [], None, False, False
        while True:
            line = f.readline()
            if not line:
                raise ValueError("PLY header without end_header")
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] == "comment" or tok[0] == "obj_info":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                if seen_element and in_vertex:
                    in_vertex = False            # a later element: stop collecting properties
                elif tok[1] == "vertex" and not seen_element:
                    in_vertex, count = True, int(tok[2])
                seen_element = True
            elif tok[0] == "property" and in_vertex:
                if tok[1] == "list":
                    raise ValueError("list properties in the vertex element are not supported")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if count is None or fmt is None:
            raise ValueError("PLY header lacks a format line or a leading vertex element")
        if fmt == "ascii":
            rows = np.loadtxt(f, max_rows=count, ndmin=2, dtype=np.float64)
            return {n: rows[:, i].astype(t) for i, (n, t) in enumerate(props)}
        order = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(n, order + t) for n, t in props])
        rec = np.frombuffer(f.read(dt.itemsize * count), dtype=dt, count=count)
        return {n: np.ascontiguousarray(rec[n]) for n, _ in props}


def load_ply(path, max_sh_degree=3):
    """hugs/models/scene.py:269-308.  Returns float32 arrays in the model's layouts: xyz [P,3],
    features_dc [P,1,3], features_rest [P,K-1,3], opacity [P,1], scaling [P,3], rotation [P,4], plus
    shs [P,K,3] = cat(features_dc, features_rest) -- what GaussianRasterizer's `shs` argument takes."""
    v = read_vertex_table(path)
    P = len(v["x"])
    xyz = np.stack([v["x"], v["y"], v["z"]], axis=1).astype(np.float32)
    opacity = np.asarray(v["opacity"], dtype=np.float32)[:, None]
    dc = np.stack([v["f_dc_0"], v["f_dc_1"], v["f_dc_2"]], axis=1).astype(np.float32)[:, :, None]   # [P,3,1]
    idx = lambda n: int(n.split("_")[-1])  # noqa: E731
    rest_names = sorted([n for n in v if n.startswith("f_rest_")], key=idx)
    K = (max_sh_degree + 1) ** 2
    if len(rest_names) != 3 * K - 3:
        raise ValueError("%d f_rest_* properties, SH degree %d needs %d" % (len(rest_names), max_sh_degree, 3 * K - 3))
    rest = np.stack([v[n] for n in rest_names], axis=1).astype(np.float32).reshape(P, 3, K - 1) if K > 1 else \
        np.zeros((P, 3, 0), np.float32)
    scale_names = sorted([n for n in v if n.startswith("scale_")], key=idx)
    rot_names = sorted([n for n in v if n.startswith("rot")], key=idx)
    scaling = np.stack([v[n] for n in scale_names], axis=1).astype(np.float32)
    rotation = np.stack([v[n] for n in rot_names], axis=1).astype(np.float32)
    features_dc = np.ascontiguousarray(dc.transpose(0, 2, 1))        # [P,1,3]
    features_rest = np.ascontiguousarray(rest.transpose(0, 2, 1))    # [P,K-1,3]
    return dict(xyz=xyz, features_dc=features_dc, features_rest=features_rest, opacity=opacity, scaling=scaling,
                rotation=rotation, shs=np.concatenate([features_dc, features_rest], axis=1))
