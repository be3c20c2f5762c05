#!/usr/bin/env python3
"""Generate tests/golden/raster_golden.npz by running the UNMODIFIED reference rasterizer
(oracle/_ref, built from /root/reference by oracle/build_ref.py) on a B200.

Run on the GPU box:   gpurun -- 'python tests/golden/make_golden_raster.py'
It writes gpurun_out/golden/raster_golden.npz; copy that file to tests/golden/ and commit it.

For each small seeded scene the fixture holds every input tensor and, from the reference run:
radii, R, the per-Gaussian preprocess state (depths, means2D, conic_opacity, rgb, tiles_touched,
cov3D), the sorted 64-bit keys and Gaussian list, tile ranges, per-pixel n_contrib and final T,
the image, and all gradients.  These pin both the CPU oracle (tests/test_oracle_golden.py, no GPU)
and the CUDA path (tests/test_parity_gpu.py).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "ml-hugs_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

SCENES = {
    # tag: (P_scene, P_human, W, H, focal, seed, degree, mode)
    "g1": (1500, 300, 200, 120, 150.0, 11, 3, "sh"),
    "g2": (2400, 600, 192, 160, 140.0, 12, 2, "sh"),
    "g3": (900, 300, 160, 96, 120.0, 13, 0, "precomp"),
}


def main():
    import refutil
    import scenes

    assert torch.cuda.is_available() and refutil.ref_available()
    ref = refutil.load_ref_rasterizer()
    out = {}
    for tag, (ps, ph, W, H, focal, seed, degree, mode) in SCENES.items():
        sc = scenes.make_scene(ps, ph, W, H, focal, seed, degree=degree, device="cuda")
        P = sc["means3D"].shape[0]
        kw = {}
        if mode == "precomp":
            # cov3D from a scale/rotation run of the reference itself, colours seeded
            r0 = scenes.run_module(sc, ref, backward=False)
            d0 = refutil.decode_ref_buffers(r0["geom"], r0["binning"], r0["img"], P, r0["num_rendered"], W, H)
            g = torch.Generator().manual_seed(seed + 5)
            kw = dict(colors_precomp=torch.rand(P, 3, generator=g).cuda(), cov3D_precomp=d0["cov3D"].clone())
        r = scenes.run_module(sc, ref, backward=True, **kw)
        R = r["num_rendered"]
        d = refutil.decode_ref_buffers(r["geom"], r["binning"], r["img"], P, R, W, H)
        n = lambda t: t.detach().cpu().numpy()  # noqa: E731
        for k in ("means3D", "scales", "rotations", "opacity", "shs", "viewmatrix", "projmatrix", "campos", "bg",
                  "dL_dcolor"):
            out["%s_in_%s" % (tag, k)] = n(sc[k])
        for k, v in kw.items():
            out["%s_in_%s" % (tag, k)] = n(v)
        out[tag + "_meta"] = np.array([W, H, degree, R, P], dtype=np.int64)
        out[tag + "_tanfov"] = np.array([sc["tanfovx"], sc["tanfovy"]], dtype=np.float64)
        out[tag + "_radii"] = n(r["radii"])
        out[tag + "_color"] = n(r["color"])
        for k in ("depths", "means2D", "conic_opacity", "rgb", "tiles_touched", "cov3D", "clamped"):
            out["%s_pre_%s" % (tag, k)] = n(d[k])
        out[tag + "_keys"] = n(d["point_list_keys"]).view(np.uint64)
        out[tag + "_point_list"] = n(d["point_list"]).view(np.uint32)
        out[tag + "_ranges"] = n(d["ranges"]).view(np.uint32)
        out[tag + "_n_contrib"] = n(d["n_contrib"]).view(np.uint32)
        out[tag + "_final_T"] = n(d["accum_alpha"])
        for k, v in r["grads"].items():
            out["%s_grad_%s" % (tag, k)] = n(v)
        print(tag, "P", P, "R", R, "visible", int((r["radii"] > 0).sum()))
    dst = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(dst, exist_ok=True)
    path = os.path.join(dst, "raster_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
