"""Scene checkpoint / PLY -> scene_gs_out loaders and the human PLY export:
the activations of hugs/models/scene.py:147-160 and the column conventions of hugs/utils/vis.py:40-60."""
import numpy as np
import pytest
import torch

import refutil  # noqa: F401
from hugs_b200 import checkpoint, ply


def _raw(n=50, seed=0):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)  # noqa: E731
    return dict(xyz=r(n, 3), features_dc=r(n, 1, 3), features_rest=r(n, 15, 3), scaling=r(n, 3) - 3,
                rotation=r(n, 4), opacity=r(n, 1))


def test_state_dict_keys_and_round_trip(tmp_path):
    raw = _raw()
    state = checkpoint.make_scene_state(**raw, active_sh_degree=2)
    assert tuple(state) == checkpoint.SCENE_KEYS                      # scene.py:70-85, same order
    p = str(tmp_path / "ckpt" / "scene_final.pth")
    checkpoint.save_scene_checkpoint(p, state)
    out, back = checkpoint.load_scene_checkpoint(p, device="cpu")
    assert back["active_sh_degree"] == 2 and out["active_sh_degree"] == 2
    assert torch.equal(out["xyz"], raw["xyz"])
    assert torch.equal(out["scales"], torch.exp(raw["scaling"]))
    assert torch.equal(out["opacity"], torch.sigmoid(raw["opacity"]))
    assert torch.allclose(out["rotq"].norm(dim=1), torch.ones(50), atol=1e-6)
    assert torch.equal(out["rotq"], torch.nn.functional.normalize(raw["rotation"]))
    assert out["shs"].shape == (50, 16, 3)
    assert torch.equal(out["shs"][:, 0], raw["features_dc"][:, 0]) and torch.equal(out["shs"][:, 1:], raw["features_rest"])
    with pytest.raises(KeyError, match="rotation"):
        checkpoint.scene_gs_out_from_state({k: v for k, v in state.items() if k != "rotation"}, device="cpu")


def test_ply_to_scene_gs_out(tmp_path):
    raw = _raw(31, seed=3)
    p = str(tmp_path / "scene_splat.ply")
    ply.save_ply(p, raw["xyz"], raw["features_dc"], raw["features_rest"], raw["opacity"], raw["scaling"],
                 raw["rotation"])
    out = checkpoint.scene_gs_out_from_ply(p, device="cpu")
    ref = checkpoint.scene_gs_out_from_state(raw, device="cpu")
    for k in ("xyz", "scales", "rotq", "shs", "opacity"):
        assert torch.equal(out[k], ref[k]), k
    assert out["active_sh_degree"] == 3


def test_human_ply_export_columns(tmp_path):
    g = torch.Generator().manual_seed(5)
    n = 20
    human = dict(xyz_canon=torch.randn(n, 3, generator=g), shs=torch.randn(n, 16, 3, generator=g),
                 opacity=torch.rand(n, 1, generator=g) * 0.98 + 0.01, scales_canon=torch.rand(n, 3, generator=g) + 0.01,
                 rotq_canon=torch.nn.functional.normalize(torch.randn(n, 4, generator=g)))
    p = str(tmp_path / "meshes" / "human_final_splat.ply")
    checkpoint.save_human_ply(human, p)
    v = ply.read_vertex_table(p)
    assert list(v) == ply.construct_list_of_attributes()
    assert np.array_equal(v["x"], human["xyz_canon"][:, 0].numpy())
    assert np.array_equal(v["opacity"], torch.log(human["opacity"] / (1 - human["opacity"]))[:, 0].numpy())
    assert np.array_equal(v["scale_1"], torch.log(human["scales_canon"])[:, 1].numpy())
    assert np.array_equal(v["rot_3"], human["rotq_canon"][:, 3].numpy())
    assert np.array_equal(v["f_dc_2"], human["shs"][:, 0, 2].numpy())
    # f_rest is channel-major: column c*15 + k holds shs[:, 1+k, c]  (vis.py:48)
    assert np.array_equal(v["f_rest_17"], human["shs"][:, 1 + 2, 1].numpy())
    back = checkpoint.scene_gs_out_from_ply(p, device="cpu")
    assert torch.allclose(back["opacity"], human["opacity"], atol=1e-6)
    assert torch.allclose(back["scales"], human["scales_canon"], rtol=1e-6)
    assert torch.equal(back["shs"], human["shs"])
