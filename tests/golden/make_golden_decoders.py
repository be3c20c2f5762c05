"""Generates tests/golden/decoders_golden.npz by importing the REFERENCE's own decoder modules
(hugs/models/modules/decoders.py: AppearanceDecoder, DeformationDecoder, GeometryDecoder) on the CPU in the
build container: seeded parameters, seeded [rows, 96] inputs, their forward outputs and autograd gradients
(float32, as the reference runs them).  /root/reference does not exist on the GPU box, so the vectors are
committed.  loguru (imported for logging only) is stubbed when absent; the module's package-relative import
of .activation is served from the reference file next to it.
usage: python tests/golden/make_golden_decoders.py"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF_DIR = "/root/reference/hugs/models/modules"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "decoders_golden.npz")
ROWS = 333   # two full 128-row tiles and a ragged one


def load_ref():
    try:
        import loguru  # noqa: F401
    except ImportError:
        lg = types.ModuleType("loguru")
        lg.logger = types.SimpleNamespace(info=print, warning=print, error=print, debug=print)
        sys.modules["loguru"] = lg
    pkg = types.ModuleType("hugs_ref_modules")
    pkg.__path__ = [REF_DIR]
    sys.modules["hugs_ref_modules"] = pkg
    for name in ("activation", "decoders"):
        spec = importlib.util.spec_from_file_location("hugs_ref_modules." + name, os.path.join(REF_DIR, name + ".py"))
        m = importlib.util.module_from_spec(spec)
        sys.modules["hugs_ref_modules." + name] = m
        spec.loader.exec_module(m)
    return sys.modules["hugs_ref_modules.decoders"]


def main():
    ref = load_ref()
    torch.manual_seed(20240607)
    g = torch.Generator().manual_seed(7)
    out = {}
    x = torch.randn(ROWS, 96, generator=g) * 0.7
    out["x"] = x.numpy()
    mods = dict(appearance=ref.AppearanceDecoder(n_features=96),
                geometry=ref.GeometryDecoder(n_features=96),
                deformation=ref.DeformationDecoder(n_features=96, disable_posedirs=True))
    for tag, m in mods.items():
        # biases and weights away from their tiny default scale so every unit is exercised
        with torch.no_grad():
            for p in m.parameters():
                p.mul_(1.5)
        xi = x.clone().requires_grad_(True)
        res = m(xi)
        keys = [k for k, v in res.items() if v is not None]
        loss = 0.0
        for i, k in enumerate(keys):
            w = torch.randn(res[k].shape, generator=g)
            out["%s/w/%s" % (tag, k)] = w.numpy()
            out["%s/out/%s" % (tag, k)] = res[k].detach().numpy()
            loss = loss + (res[k] * w).sum()
        loss.backward()
        out["%s/dx" % tag] = xi.grad.numpy()
        for k, v in m.state_dict().items():
            out["%s/param/%s" % (tag, k)] = v.detach().numpy()
        for k, p in m.named_parameters():
            out["%s/grad/%s" % (tag, k)] = p.grad.numpy()
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, "%d arrays, %.1f KB" % (len(out), os.path.getsize(OUT) / 1024))


if __name__ == "__main__":
    main()
