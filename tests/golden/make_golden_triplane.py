"""Generates tests/golden/triplane_golden.npz by importing the REFERENCE's own TriPlane module
(hugs/models/modules/triplane.py) on the CPU in the build container and running its forward and
autograd backward.  /root/reference does not exist on the GPU box, so the vectors are committed.
loguru (imported at the top of that module, only for logging) is stubbed when absent.
usage: python tests/golden/make_golden_triplane.py"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/hugs/models/modules/triplane.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "triplane_golden.npz")

# tag: (features, resX, resY, resZ, points) -- unequal resolutions pin which coordinate indexes which axis
CASES = {"a": (4, 8, 8, 8, 64), "b": (32, 16, 12, 20, 300), "c": (3, 5, 9, 2, 100), "d": (40, 6, 6, 6, 50)}


def load_ref():
    try:
        import loguru  # noqa: F401
    except ImportError:
        lg = types.ModuleType("loguru")
        lg.logger = types.SimpleNamespace(info=print, warning=print, error=print, debug=print)
        sys.modules["loguru"] = lg
    spec = importlib.util.spec_from_file_location("hugs_ref_triplane", REF)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def points(n, g):
    x = torch.rand(n, 3, generator=g) * 2 - 1                 # the box center +- scale/2 = [-1, 1]
    x[0] = torch.tensor([-1.0, -1.0, -1.0])                   # exact corners and faces
    x[1] = torch.tensor([1.0, 1.0, 1.0])
    x[2] = torch.tensor([1.0, -1.0, 0.0])
    x[3] = torch.tensor([0.0, 0.0, 0.0])
    x[4] = torch.tensor([1.0015, -1.0015, 0.5])               # inside the EPS slack: partly zero-padded
    x[5] = torch.tensor([-1.0019, 1.0019, 1.0019])
    return x


def main():
    ref = load_ref()
    out = {}
    for tag, (feat, rx, ry, rz, n) in CASES.items():
        g = torch.Generator().manual_seed(500 + ord(tag))
        tp = ref.TriPlane(features=feat, resX=rx, resY=ry, resZ=rz)
        with torch.no_grad():
            for p in (tp.plane_xy, tp.plane_xz, tp.plane_yz):
                p.copy_(torch.randn(p.shape, generator=g))
        x = points(n, g).requires_grad_(True)
        w = torch.randn(n, 3 * feat, generator=g)
        y = tp(x)
        (y * w).sum().backward()
        out.update({tag + "_x": x.detach().numpy(), tag + "_w": w.numpy(), tag + "_feat": y.detach().numpy(),
                    tag + "_gx": x.grad.numpy(), tag + "_pxy": tp.plane_xy.detach().numpy(),
                    tag + "_pxz": tp.plane_xz.detach().numpy(), tag + "_pyz": tp.plane_yz.detach().numpy(),
                    tag + "_gpxy": tp.plane_xy.grad.numpy(), tag + "_gpxz": tp.plane_xz.grad.numpy(),
                    tag + "_gpyz": tp.plane_yz.grad.numpy()})
    # a batched input shape [2, 7, 3] (the module reshapes to [..., 3C])
    tp = ref.TriPlane(features=4, resX=8, resY=8, resZ=8)
    xb = torch.rand(2, 7, 3, generator=torch.Generator().manual_seed(9)) * 2 - 1
    out.update(batched_x=xb.numpy(), batched_feat=tp(xb).detach().numpy(), batched_pxy=tp.plane_xy.detach().numpy(),
               batched_pxz=tp.plane_xz.detach().numpy(), batched_pyz=tp.plane_yz.detach().numpy())
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
