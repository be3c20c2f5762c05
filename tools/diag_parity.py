"""Diagnostic (GPU box): our CUDA path vs the reference build vs the CPU oracle on seeded scenes.
Prints mismatch statistics instead of asserting; writes gpurun_out/diag_parity.json."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "ml-hugs_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch

import refutil
import scenes
import hugs_b200
import diff_gaussian_rasterization as ours

torch.backends.cuda.matmul.allow_tf32 = False


def bits(t):
    return t.contiguous().view(torch.int32)


def compare(name, sc, with_oracle=True):
    res = {"scene": name}
    W, H = sc["image_width"], sc["image_height"]
    P = sc["means3D"].shape[0]
    torch.cuda.synchronize()
    t0 = time.time()
    o = scenes.run_module(sc, ours)
    torch.cuda.synchronize()
    res["ours_s"] = time.time() - t0
    R = o["num_rendered"]
    od = refutil.decode_our_buffers(o["geom"], o["binning"], o["img"], P, R, W, H)
    res.update(P=P, R=R, visible=int((o["radii"] > 0).sum()))
    if refutil.ref_available():
        ref = refutil.load_ref_rasterizer()
        t0 = time.time()
        r = scenes.run_module(sc, ref)
        torch.cuda.synchronize()
        res["ref_s"] = time.time() - t0
        Rr = r["num_rendered"]
        rd = refutil.decode_ref_buffers(r["geom"], r["binning"], r["img"], P, Rr, W, H)
        vis = r["radii"] > 0
        res["ref_R"] = Rr
        res["radii_mismatch"] = int((o["radii"] != r["radii"]).sum())
        res["depth_bits_mismatch"] = int((bits(od["depths"])[vis] != bits(rd["depths"])[vis]).sum())
        res["means2D_bits_mismatch"] = int((bits(od["means2D"])[vis] != bits(rd["means2D"])[vis]).any(1).sum())
        cm = (bits(od["conic_opacity"])[vis] != bits(rd["conic_opacity"])[vis])
        res["conic_bits_mismatch"] = [int(cm[:, k].sum()) for k in range(4)]
        res["rgb_maxabs"] = float((od["rgb"][vis] - rd["rgb"][vis]).abs().max()) if vis.any() else 0.0
        res["tiles_mismatch"] = int((od["tiles_touched"][vis] != rd["tiles_touched"][vis]).sum())
        if R == Rr and R > 0:
            res["keys_mismatch"] = int((od["point_list_keys"] != rd["point_list_keys"]).sum())
            res["list_mismatch"] = int((od["point_list"] != rd["point_list"]).sum())
        res["ranges_mismatch"] = int((od["ranges"] != rd["ranges"]).any(1).sum())
        res["n_contrib_mismatch"] = int((od["n_contrib"] != rd["n_contrib"]).sum())
        res["final_T_maxabs"] = float((od["accum_alpha"] - rd["accum_alpha"]).abs().max())
        res["final_T_bits_mismatch"] = int((bits(od["accum_alpha"]) != bits(rd["accum_alpha"])).sum())
        res["color_maxabs"] = float((o["color"] - r["color"]).abs().max())
        res["grad_rel_vs_ref"] = {k: refutil.rel_err(o["grads"][k].cpu().numpy(), r["grads"][k].cpu().numpy())
                                  for k in o["grads"]}
        res["grad_robust_vs_ref"] = {k: refutil.robust_rel_err(o["grads"][k].cpu().numpy(), r["grads"][k].cpu().numpy())
                                     for k in o["grads"]}
    if with_oracle:
        t0 = time.time()
        orc = scenes.run_oracle(sc)
        res["oracle_s"] = time.time() - t0
        fw = orc["fw"]
        res["oracle_R"] = fw["binning"]["num_rendered"]
        res["oracle_radii_mismatch"] = int((o["radii"].cpu().numpy() != fw["pre"]["radii"]).sum())
        if fw["binning"]["num_rendered"] == R and R > 0:
            res["oracle_keys_mismatch"] = int((od["point_list_keys"].cpu().numpy().view(np.uint64) != fw["binning"]["keys"]).sum())
            res["oracle_list_mismatch"] = int((od["point_list"].cpu().numpy().view(np.uint32) != fw["binning"]["point_list"]).sum())
        res["oracle_ranges_mismatch"] = int((od["ranges"].cpu().numpy().view(np.uint32) != fw["binning"]["ranges"]).any(1).sum())
        res["oracle_n_contrib_mismatch"] = int((od["n_contrib"].cpu().numpy().view(np.uint32) != fw["n_contrib"]).sum())
        res["oracle_color_maxabs"] = float(np.abs(o["color"].cpu().numpy() - fw["color"]).max())
        g = orc["grads"]
        pairs = dict(means3D="dL_dmean3D", means2D="dL_dmean2D", opacity="dL_dopacity", shs="dL_dsh",
                     scales="dL_dscale", rotations="dL_drot")
        res["grad_rel_vs_oracle"] = {k: refutil.rel_err(o["grads"][k].cpu().numpy(), g[v]) for k, v in pairs.items()}
        if refutil.ref_available():
            res["ref_grad_rel_vs_oracle"] = {k: refutil.rel_err(r["grads"][k].cpu().numpy(), g[v]) for k, v in pairs.items()}
    print(json.dumps(res, indent=1), flush=True)
    return res


def main():
    dev = "cuda"
    out = []
    which = sys.argv[1:] or ["tiny", "small", "medium", "hs"]
    if "tiny" in which:
        out.append(compare("tiny_200x120", scenes.make_scene(1500, 300, 200, 120, 150.0, 11, device=dev)))
    if "small" in which:
        out.append(compare("small_256", scenes.make_scene(20000, 5000, 256, 256, 187.0, 12, device=dev)))
    if "medium" in which:
        out.append(compare("medium_640x360", scenes.make_scene(100000, 20000, 640, 360, 466.0, 13, device=dev)))
    if "hs" in which:
        out.append(compare("hs_1080p", scenes.make_scene(1000000, 110210, 1920, 1080, 1400.0, 1002, device=dev)))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "diag_parity.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
