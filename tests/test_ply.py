"""3DGS PLY format (hugs/models/scene.py:229-308).  `plyfile` is not installable here, so the writer is
pinned on the format itself (header text, column order, channel-major SH columns computed with the
reference's own tensor expressions) and on round trips through the reader."""
import numpy as np
import torch

import refutil  # noqa: F401
from hugs_b200 import ply


def _model(P=37, K=16, seed=0):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)  # noqa: E731
    return dict(xyz=r(P, 3), features_dc=r(P, 1, 3), features_rest=r(P, K - 1, 3), opacity=r(P, 1), scaling=r(P, 3),
                rotation=r(P, 4))


def test_attribute_list_matches_reference_order():
    names = ply.construct_list_of_attributes()
    assert len(names) == 62 and names[:6] == ["x", "y", "z", "nx", "ny", "nz"]
    assert names[6:9] == ["f_dc_0", "f_dc_1", "f_dc_2"] and names[9] == "f_rest_0" and names[53] == "f_rest_44"
    assert names[54] == "opacity" and names[55:58] == ["scale_0", "scale_1", "scale_2"]
    assert names[58:] == ["rot_0", "rot_1", "rot_2", "rot_3"]


def test_save_writes_the_plyfile_layout(tmp_path):
    m = _model()
    path = str(tmp_path / "sub" / "point_cloud.ply")
    ply.save_ply(path, **m)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().split("\n")
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 37"]
    assert lines[3:-1] == ["property float %s" % n for n in ply.construct_list_of_attributes()]
    table = np.frombuffer(body, dtype="<f4").reshape(37, 62)
    # the reference's own expressions (scene.py:246-258)
    f_dc = m["features_dc"].transpose(1, 2).flatten(start_dim=1).contiguous().numpy()
    f_rest = m["features_rest"].transpose(1, 2).flatten(start_dim=1).contiguous().numpy()
    want = np.concatenate((m["xyz"].numpy(), np.zeros((37, 3), np.float32), f_dc, f_rest, m["opacity"].numpy(),
                           m["scaling"].numpy(), m["rotation"].numpy()), axis=1)
    assert np.array_equal(table, want)


def test_round_trip_and_rasterizer_layout(tmp_path):
    for K, deg in ((16, 3), (4, 1), (1, 0)):
        m = _model(P=11, K=K, seed=K)
        path = str(tmp_path / ("k%d.ply" % K))
        ply.save_ply(path, **m)
        got = ply.load_ply(path, max_sh_degree=deg)
        for k in m:
            assert np.array_equal(got[k], m[k].numpy()), k
        assert got["shs"].shape == (11, K, 3)
        assert np.array_equal(got["shs"], torch.cat([m["features_dc"], m["features_rest"]], dim=1).numpy())


def test_reader_handles_ascii_big_endian_and_trailing_elements(tmp_path):
    names = ply.construct_list_of_attributes(3, 0, 3, 4)
    rows = np.arange(2 * len(names), dtype=np.float32).reshape(2, -1) / 7
    a = tmp_path / "a.ply"
    a.write_text("ply\nformat ascii 1.0\ncomment made by hand\nelement vertex 2\n" +
                 "".join("property float %s\n" % n for n in names) + "element face 0\nproperty list uchar int vertex_indices\n"
                 "end_header\n" + "\n".join(" ".join(repr(float(x)) for x in r) for r in rows) + "\n")
    got = ply.load_ply(str(a), max_sh_degree=0)
    assert np.allclose(got["xyz"], rows[:, :3]) and np.allclose(got["rotation"], rows[:, -4:])
    b = tmp_path / "b.ply"
    with open(b, "wb") as f:
        f.write(("ply\nformat binary_big_endian 1.0\nelement vertex 2\n" + "".join("property float %s\n" % n for n in names) +
                 "end_header\n").encode())
        f.write(rows.astype(">f4").tobytes())
    got = ply.load_ply(str(b), max_sh_degree=0)
    assert np.array_equal(got["scaling"], rows[:, -7:-4]) and np.array_equal(got["opacity"][:, 0], rows[:, 9])
