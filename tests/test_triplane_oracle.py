"""The tri-plane oracle against the golden vectors the reference's own module produced
(tests/golden/make_golden_triplane.py): features bit-exact, gradients to float32 accuracy."""
import os

import numpy as np
import pytest

import refutil  # noqa: F401
from oracle import triplane_oracle as TO

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "triplane_golden.npz")


@pytest.mark.parametrize("tag", ["a", "b", "c", "d"])
def test_oracle_matches_reference_golden(tag):
    g = np.load(GOLD)
    planes = [g[tag + k] for k in ("_pxy", "_pxz", "_pyz")]
    feat = TO.forward(g[tag + "_x"], *planes)
    assert feat.dtype == np.float32 and np.array_equal(feat, g[tag + "_feat"])
    gx, gxy, gxz, gyz = TO.backward(g[tag + "_x"], *planes, g[tag + "_w"])
    assert np.abs(gx - g[tag + "_gx"]).max() <= 1e-5 * np.abs(g[tag + "_gx"]).max()
    for ours, key in ((gxy, "_gpxy"), (gxz, "_gpxz"), (gyz, "_gpyz")):
        assert np.abs(ours - g[tag + key]).max() <= 1e-5 * max(np.abs(g[tag + key]).max(), 1.0)


def test_oracle_batched_shape_and_zero_padding():
    g = np.load(GOLD)
    x = g["batched_x"]
    feat = TO.forward(x.reshape(-1, 3), g["batched_pxy"], g["batched_pxz"], g["batched_pyz"]).reshape(2, 7, -1)
    assert np.array_equal(feat, g["batched_feat"])
    # far outside the box every corner is out of range: zeros
    far = TO.forward(np.array([[5.0, -7.0, 9.0]], np.float32), g["batched_pxy"], g["batched_pxz"], g["batched_pyz"])
    assert not far.any()
