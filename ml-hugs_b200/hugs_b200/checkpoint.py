"""Trained-asset loaders on either side of the path (checkpoint half): the scene
model's `.pth` dictionary and 3DGS PLY files become the `scene_gs_out` dictionary that
renderer.render_human_scene consumes, and a posed human is exported as a 3DGS PLY.

Mirrors:
  SceneGS.state_dict / restore   hugs/models/scene.py:70-104   (keys of the .pth dictionary)
  SceneGS.forward                hugs/models/scene.py:147-160  (exp / normalize / sigmoid activations)
  vis.save_ply                   hugs/utils/vis.py:40-60       (human export: logit opacity, log scales)
Host-side file handling only; the tensors are placed on `device` for the CUDA path.
"""
import os

import numpy as np
import torch

from . import ply as _ply

SCENE_KEYS = ("active_sh_degree", "xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity",
              "max_radii2D", "xyz_gradient_accum", "denom", "optimizer", "spatial_lr_scale")


def make_scene_state(xyz, features_dc, features_rest, scaling, rotation, opacity, active_sh_degree=3,
                     max_radii2D=None, xyz_gradient_accum=None, denom=None, optimizer=None, spatial_lr_scale=1.0):
    """The dictionary SceneGS.state_dict() returns (scene.py:70-85), raw (pre-activation) parameters."""
    n = xyz.shape[0]
    z = lambda *s: torch.zeros(*s, device=xyz.device)  # noqa: E731
    return {
        "active_sh_degree": int(active_sh_degree), "xyz": xyz, "features_dc": features_dc,
        "features_rest": features_rest, "scaling": scaling, "rotation": rotation, "opacity": opacity,
        "max_radii2D": max_radii2D if max_radii2D is not None else z(n),
        "xyz_gradient_accum": xyz_gradient_accum if xyz_gradient_accum is not None else z(n, 1),
        "denom": denom if denom is not None else z(n, 1),
        "optimizer": optimizer if optimizer is not None else {}, "spatial_lr_scale": spatial_lr_scale,
    }


def scene_gs_out_from_state(state, device="cuda"):
    """scene.py:147-160 applied to a state dictionary."""
    missing = [k for k in ("xyz", "features_dc", "features_rest", "scaling", "rotation", "opacity") if k not in state]
    if missing:
        raise KeyError("scene checkpoint lacks %s" % ", ".join(missing))
    t = lambda k: torch.as_tensor(state[k]).detach().to(device=device, dtype=torch.float32)  # noqa: E731
    return {
        "xyz": t("xyz").contiguous(),
        "scales": torch.exp(t("scaling")).contiguous(),
        "rotq": torch.nn.functional.normalize(t("rotation")).contiguous(),
        "shs": torch.cat((t("features_dc"), t("features_rest")), dim=1).contiguous(),
        "opacity": torch.sigmoid(t("opacity")).contiguous(),
        "active_sh_degree": int(state.get("active_sh_degree", 3)),
    }


def save_scene_checkpoint(path, state):
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    torch.save(state, path)


def load_scene_checkpoint(path, device="cuda"):
    """torch.load of a scene_*.pth written by the trainer (gs_trainer.py:400-401) -> (scene_gs_out, state)."""
    state = torch.load(path, map_location="cpu", weights_only=False)
    return scene_gs_out_from_state(state, device), state


def scene_gs_out_from_ply(path, device="cuda", max_sh_degree=3, active_sh_degree=None):
    """scene.py:269-308 (load_ply) followed by :147-160 (forward)."""
    v = _ply.load_ply(path, max_sh_degree=max_sh_degree)
    state = {k: torch.from_numpy(v[k]) for k in ("xyz", "features_dc", "features_rest", "scaling", "rotation",
                                                  "opacity")}
    state["active_sh_degree"] = max_sh_degree if active_sh_degree is None else active_sh_degree
    return scene_gs_out_from_state(state, device)


@torch.no_grad()
def save_human_ply(human_gs_out, path):
    """hugs/utils/vis.py:40-60: canonical positions, logit(opacity), log(canonical scales), canonical
    quaternions and the SH coefficients of a human model's output as a 3DGS PLY."""
    shs = human_gs_out["shs"]
    op = human_gs_out["opacity"]
    _ply.save_ply(path, human_gs_out["xyz_canon"], shs[:, :1], shs[:, 1:], torch.log(op / (1 - op)),
                  torch.log(human_gs_out["scales_canon"]), human_gs_out["rotq_canon"])
