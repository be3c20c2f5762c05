"""Errors of the fused decoder kernels against the reference golden vectors (GPU box).  usage: python tools/diag_decoders.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ml-hugs_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch

import test_decoders_gpu as T


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


g = T._golden()
for tag in ("appearance", "geometry", "deformation"):
    m = T._build(tag, g, "cuda")
    x = torch.from_numpy(g["x"]).cuda().requires_grad_(True)
    res = m(x)
    loss = 0.0
    for k in [k[len(tag) + 5:] for k in g if k.startswith(tag + "/out/")]:
        print(tag, "out", k, "%.2e" % rel(res[k].detach().cpu().numpy(), g["%s/out/%s" % (tag, k)]))
        loss = loss + (res[k] * torch.from_numpy(g["%s/w/%s" % (tag, k)]).cuda()).sum()
    loss.backward()
    print(tag, "dx", "%.2e" % rel(x.grad.cpu().numpy(), g[tag + "/dx"]))
    for name, p in m.named_parameters():
        print(tag, "grad", name, "%.2e" % rel(p.grad.cpu().numpy(), g["%s/grad/%s" % (tag, name)]), tuple(p.shape))

# ---- full size against torch autograd (fp32) ------------------------------------------------------------
from hugs_b200 import decoders as D
mods = [T._build(t, g, "cuda") for t in ("appearance", "geometry", "deformation")]
x0 = torch.randn(110_210, 96, device="cuda", generator=torch.Generator(device="cuda").manual_seed(3)) * 0.7


def run(fused):
    for m in mods:
        m.zero_grad()
    x = x0.clone().requires_grad_(True)
    if fused:
        a, b, c = D.decode_all(x, *mods[:2], mods[2])
    else:
        app, geo, dfm = mods
        h = app.net(x)
        a = {"shs": app.shs(h), "opacity": app.opacity(h)}
        h = geo.net(x)
        b = {"xyz": geo.xyz[1](h), "rotations": geo.rotations[1](h), "scales": torch.nn.functional.gelu(geo.scales[1](h))}
        h = dfm.net(x)
        c = {"lbs_weights": torch.nn.functional.gelu(dfm.skinning(torch.nn.functional.gelu(dfm.skinning_linear(h))))}
    outs = [a["shs"], a["opacity"], b["xyz"], b["rotations"], b["scales"], c["lbs_weights"]]
    gen = torch.Generator(device="cuda").manual_seed(11)
    loss = sum((o * torch.randn(o.shape, device="cuda", generator=gen)).sum() for o in outs)
    loss.backward()
    grads = {"x": x.grad.clone()}
    for t, m in zip(("a", "g", "d"), mods):
        for n, p in m.named_parameters():
            grads[t + "." + n] = p.grad.clone()
    return grads


gf, gt = run(True), run(False)
gt64 = None
for k in gt:
    print("full", k, "%.2e" % rel(gf[k].cpu().numpy(), gt[k].cpu().numpy()))
