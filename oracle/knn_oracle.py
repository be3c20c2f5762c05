"""CPU oracle of simple-knn's distCUDA2 (TEST INFRASTRUCTURE only): brute-force exact 3-NN.

Restates the RESULT of submodules/simple-knn/simple_knn.cu:149-183 (boxMeanDist): for every point the
mean of the three smallest squared distances to the OTHER points (index-wise: coincident points count
with distance 0), missing neighbours counting as FLT_MAX (:156, the float sum then overflows to inf).
The reference's Morton ordering and box pruning only decide which candidates are skipped, never the
answer.  Parity unpinned against reference-authored vectors (the reference has no tests or fixtures
and its extension needs a GPU to run); pinned against hand-computed cases in tests/test_knn_oracle.py.
"""
import numpy as np

FLT_MAX = np.float32(3.4028234663852886e38)


def dist2_mean3(points, chunk=1024):
    p = np.asarray(points, dtype=np.float32)
    P = p.shape[0]
    out = np.empty(P, dtype=np.float32)
    p64 = p.astype(np.float64)
    for s in range(0, P, chunk):
        q = p64[s:s + chunk]
        d = ((q[:, None, :] - p64[None, :, :]) ** 2).sum(-1)          # exact enough: float64
        d[np.arange(q.shape[0]), np.arange(s, s + q.shape[0])] = np.inf  # not itself
        k = min(3, P - 1)
        best = np.sort(np.partition(d, k - 1, axis=1)[:, :k], axis=1) if k > 0 else np.empty((q.shape[0], 0))
        b = np.full((q.shape[0], 3), FLT_MAX, dtype=np.float32)
        b[:, :k] = best.astype(np.float32)
        with np.errstate(over="ignore"):
            out[s:s + chunk] = ((b[:, 0] + b[:, 1]) + b[:, 2]) / np.float32(3.0)
    return out
