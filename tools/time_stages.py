"""Quick per-stage timing on the GPU box (CUDA events around whole fwd / bwd, ours vs reference)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "ml-hugs_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch

import refutil
import scenes
import diff_gaussian_rasterization as ours


def time_module(sc, module, iters=10, warm=3):
    names = ["means3D", "scales", "rotations", "opacity", "shs"]
    t = {k: sc[k].detach().clone().requires_grad_(True) for k in names}
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    rast = module.GaussianRasterizer(scenes.settings_of(sc, module))
    fw, bw = [], []
    for it in range(iters + warm):
        for v in t.values():
            v.grad = None
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        color, radii = rast(means3D=t["means3D"], means2D=means2D, opacities=t["opacity"], shs=t["shs"],
                            scales=t["scales"], rotations=t["rotations"])
        e[1].record()
        (color * sc["dL_dcolor"]).sum().backward()
        e[2].record()
        torch.cuda.synchronize()
        if it >= warm:
            fw.append(e[0].elapsed_time(e[1]))
            bw.append(e[1].elapsed_time(e[2]))
    fw.sort()
    bw.sort()
    return fw[len(fw) // 2], bw[len(bw) // 2]


def main():
    dev = "cuda"
    cfgs = {
        "human_1080p": (0, 110210),
        "hs_1080p": (1000000, 110210),
    }
    res = {}
    for name, (ps, ph) in cfgs.items():
        sc = scenes.make_scene(ps, ph, 1920, 1080, 1400.0, 1002, device=dev)
        f, b = time_module(sc, ours)
        res[name] = dict(ours_fwd_ms=f, ours_bwd_ms=b)
        if refutil.ref_available():
            f, b = time_module(sc, refutil.load_ref_rasterizer())
            res[name].update(ref_fwd_ms=f, ref_bwd_ms=b)
        print(name, json.dumps(res[name]), flush=True)
    with open(os.path.join(ROOT, "gpurun_out", "time_stages.json"), "w") as f:
        json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
