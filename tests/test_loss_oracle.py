"""The CPU oracle of the image-space losses against the golden vectors generated from the reference's
own hugs/losses/utils.py (tests/golden/make_golden_loss.py)."""
import os

import numpy as np
import pytest

import refutil  # noqa: F401  (puts the repo root on sys.path)
from oracle import loss_oracle as LO

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "loss_golden.npz")
TAGS = ["a", "b", "c", "d", "e"]


def test_window_matches_reference_construction():
    w = LO.window_1d()
    assert w.dtype == np.float32 and w.shape == (11,)
    assert abs(float(w.sum()) - 1.0) < 1e-6 and w[5] == w.max() and np.allclose(w, w[::-1])


@pytest.mark.parametrize("tag", TAGS)
def test_loss_oracle_matches_reference_python(tag):
    g = np.load(GOLD)
    o = LO.l1_ssim(g[tag + "_pred"], g[tag + "_gt"], g_l1=0.8, g_ssim=-0.2)
    assert abs(o["l1"] - float(g[tag + "_l1"])) <= 1e-6
    assert abs(o["ssim"] - float(g[tag + "_ssim"])) <= 1e-5
    ref = g[tag + "_grad"].astype(np.float64)
    assert np.abs(o["grad"] - ref).max() <= 1e-3 * np.abs(ref).max()
