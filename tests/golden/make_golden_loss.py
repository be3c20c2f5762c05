"""Generates tests/golden/loss_golden.npz by importing the REFERENCE's own loss functions
(hugs/losses/utils.py: l1_loss, ssim) on the CPU in the build container.  /root/reference does not
exist on the GPU box, so the vectors are committed.  pytorch3d (imported at the top of that module
for an unrelated regulariser) is not installable here and is stubbed.
usage: python tests/golden/make_golden_loss.py"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/hugs/losses/utils.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "loss_golden.npz")


def load_ref():
    p3d, ops = types.ModuleType("pytorch3d"), types.ModuleType("pytorch3d.ops")
    ops.laplacian = lambda *a, **k: None
    p3d.ops = ops
    sys.modules.setdefault("pytorch3d", p3d)
    sys.modules.setdefault("pytorch3d.ops", ops)
    spec = importlib.util.spec_from_file_location("hugs_ref_loss_utils", REF)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def main():
    ref = load_ref()
    out = {}
    cases = {"a": (3, 37, 53), "b": (3, 64, 48), "c": (3, 16, 16), "d": (3, 5, 7), "e": (1, 48, 80)}
    for tag, shape in cases.items():
        g = torch.Generator().manual_seed(100 + ord(tag))
        gt = torch.rand(shape, generator=g)
        # prediction = blurred/perturbed target so that SSIM is neither ~0 nor ~1
        pred = (gt + 0.25 * (torch.rand(shape, generator=g) - 0.5)).clamp(0, 1)
        pred[:, : shape[1] // 3] = gt[:, : shape[1] // 3]          # a region of exact equality (|x-y| = 0)
        pred = pred.clone().requires_grad_(True)
        l1 = ref.l1_loss(pred, gt)
        s = ref.ssim(pred, gt)
        loss = 0.8 * l1 + 0.2 * (1.0 - s)                           # the reference's weights (loss.py: l_l1_w, l_ssim_w)
        loss.backward()
        out[tag + "_pred"] = pred.detach().numpy()
        out[tag + "_gt"] = gt.numpy()
        out[tag + "_l1"] = np.float64(l1.item())
        out[tag + "_ssim"] = np.float64(s.item())
        out[tag + "_grad"] = pred.grad.numpy()
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items() if k.endswith("_grad")})


if __name__ == "__main__":
    main()
