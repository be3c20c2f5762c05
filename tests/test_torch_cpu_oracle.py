"""The pure-PyTorch CPU evaluation (oracle/torch_cpu.py: BASELINE.json configs[0], bench.py's cpu_baseline for
the 256x256 workload) against the C oracle, which is pinned on the reference build's golden vectors
(tests/test_oracle_golden.py): image <= 1e-4 max-abs, gradients <= 1e-3 of the largest element."""
import numpy as np
import torch

import refutil
import scenes
from oracle import torch_cpu as TC


def _run_torch(sc):
    names = ("means3D", "scales", "rotations", "opacity", "shs")
    t = {k: sc[k].detach().clone().requires_grad_(True) for k in names}
    img = TC.render(t["means3D"], t["scales"], t["rotations"], t["opacity"], t["shs"], sc["degree"], sc["viewmatrix"],
                    sc["projmatrix"], sc["campos"], sc["bg"], sc["image_width"], sc["image_height"], sc["tanfovx"],
                    sc["tanfovy"])
    (img * sc["dL_dcolor"]).sum().backward()
    return img.detach(), {k: t[k].grad for k in names}


def test_torch_cpu_render_matches_c_oracle():
    sc = scenes.make_scene(1200, 800, 128, 96, 110.0, seed=41)
    img, grads = _run_torch(sc)
    orc = scenes.run_oracle(sc, backward=True)
    assert np.abs(img.numpy() - orc["fw"]["color"]).max() <= 1e-4
    g = orc["grads"]
    pairs = dict(means3D=g["dL_dmean3D"], scales=g["dL_dscale"], rotations=g["dL_drot"], opacity=g["dL_dopacity"],
                 shs=g["dL_dsh"])
    for k, ref in pairs.items():
        assert refutil.rel_err(grads[k].numpy().reshape(ref.shape), ref) <= 1e-3, k


def test_torch_cpu_lbs_matches_c_oracle():
    from hugs_b200 import synth
    from oracle import oracle as O
    h = synth.make_human_gaussians(3000, seed=42)
    A = h["A"][0]
    xyz, q, sc_ = TC.deform(A, h["xyz"], h["lbs_weights"], h["rotmat"], h["scales"], h["transl"][0])
    n = lambda t: t.numpy()  # noqa: E731
    oxyz, oq, _, _ = O.lbs_forward(n(A), n(h["xyz"]), n(h["lbs_weights"]), n(h["rotmat"]), n(h["scales"]),
                                   n(h["transl"][0]), 1.0)
    assert np.abs(oxyz - xyz.numpy()).max() < 1e-5
    # q and -q are the same rotation
    qa = q.numpy()
    sign = np.sign((qa * oq).sum(1, keepdims=True))
    assert np.abs(qa * sign - oq).max() < 1e-5
