"""The brute-force distCUDA2 oracle on cases worked out by hand."""
import numpy as np

import refutil  # noqa: F401
from oracle import knn_oracle as KO


def test_hand_computed_cases():
    # four points on a line at 0, 1, 3, 7: squared distances to the three others
    p = np.array([[0, 0, 0], [1, 0, 0], [3, 0, 0], [7, 0, 0]], dtype=np.float32)
    want = np.array([(1 + 9 + 49) / 3, (1 + 4 + 36) / 3, (9 + 4 + 16) / 3, (49 + 36 + 16) / 3], dtype=np.float32)
    assert np.allclose(KO.dist2_mean3(p), want)
    # a unit cube corner sees three neighbours at distance 1
    cube = np.array([[x, y, z] for x in (0, 1) for y in (0, 1) for z in (0, 1)], dtype=np.float32)
    assert np.allclose(KO.dist2_mean3(cube), 1.0)
    # coincident points count (distance 0)
    dup = np.array([[0, 0, 0], [0, 0, 0], [0, 0, 0], [0, 0, 0], [5, 5, 5]], dtype=np.float32)
    assert np.allclose(KO.dist2_mean3(dup)[:4], 0.0) and np.isclose(KO.dist2_mean3(dup)[4], 75.0)


def test_fewer_than_four_points_overflow_like_the_reference():
    assert np.isinf(KO.dist2_mean3(np.zeros((1, 3), np.float32))).all()
    assert np.isinf(KO.dist2_mean3(np.random.rand(2, 3).astype(np.float32))).all()
    # three points: two real neighbours + one FLT_MAX, no overflow
    v = KO.dist2_mean3(np.random.rand(3, 3).astype(np.float32))
    assert np.isfinite(v).all() and (v > 1e38).all()
