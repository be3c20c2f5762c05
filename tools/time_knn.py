"""GPU-box timing of distCUDA2 (ours vs the unmodified reference simple-knn build when present)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "ml-hugs_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch

import refutil
from hugs_b200 import synth
from hugs_b200.knn import distCUDA2


def timeit(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    out = {}
    for P in (110_210, 1_000_000):
        cam = synth.make_camera(1920, 1080, 1400.0, seed=1002)
        pts = synth.make_scene_gaussians(P, 1004, cam["tanfovx"], cam["tanfovy"])["means3D"].cuda().contiguous()
        row = {"ours_ms": round(timeit(lambda: distCUDA2(pts)), 3)}
        if refutil.ref_knn_available():
            ref = refutil.load_ref_knn()
            row["reference_ms"] = round(timeit(lambda: ref(pts)), 3)
            row["speedup"] = round(row["reference_ms"] / row["ours_ms"], 2)
        out[str(P)] = row
    print(json.dumps({"op": "distCUDA2 (mean squared distance to 3 nearest neighbours)", "points": out}))


if __name__ == "__main__":
    main()
