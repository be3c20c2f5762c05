"""Timing of the fused decoder kernel against torch's evaluation of the same modules (GPU box).
usage: python tools/time_decoders.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ml-hugs_b200"))
import torch

from hugs_b200 import _native, decoders as D


def timed(fn, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    torch.manual_seed(0)
    N = 110210
    x = torch.randn(N, 96, device="cuda")
    app, geo, dfm = D.AppearanceDecoder(96).cuda(), D.GeometryDecoder(96).cuda(), D.DeformationDecoder(96, disable_posedirs=True).cuda()

    def torch_path():
        with torch.no_grad():
            h = app.net(x); app.shs(h); app.opacity(h)
            # the reference evaluates the shared trunk once per head (Sequential(self.net, Linear))
            geo.xyz(x); geo.rotations(x); torch.nn.functional.gelu(geo.scales(x))
            h = dfm.net(x)
            torch.nn.functional.gelu(dfm.skinning(torch.nn.functional.gelu(dfm.skinning_linear(h))))

    def fused():
        with torch.no_grad():
            D.decode_all(x, app, geo, dfm)

    xg = x.clone().requires_grad_(True)

    def torch_train():
        h = app.net(xg); o = [app.shs(h), app.opacity(h)]
        o += [geo.xyz(xg), geo.rotations(xg), torch.nn.functional.gelu(geo.scales(xg))]
        h = dfm.net(xg)
        o.append(torch.nn.functional.gelu(dfm.skinning(torch.nn.functional.gelu(dfm.skinning_linear(h)))))
        sum(t.sum() for t in o).backward()

    def fused_train():
        a, b, c = D.decode_all(xg, app, geo, dfm)
        (a["shs"].sum() + a["opacity"].sum() + b["xyz"].sum() + b["rotations"].sum() + b["scales"].sum()
         + c["lbs_weights"].sum()).backward()

    out = {"rows": N, "torch_fp32_ms": timed(torch_path), "fused_ms_with_packing": timed(fused),
           "torch_fp32_fwd_bwd_ms": timed(torch_train, 20), "fused_fwd_bwd_ms": timed(fused_train, 20)}
    # the fused launch alone (weights already packed): stage a call by hand
    specs = [d.spec() for d in (app, geo, dfm)]
    with torch.no_grad():
        fused()
    torch.cuda.synchronize()
    flops = 0
    for sp in specs:
        for W, _ in list(sp.hidden) + [sp.head]:
            flops += 2 * N * W.shape[0] * W.shape[1]
    out["useful_gflop"] = flops / 1e9
    out["fused_tflops_incl_packing"] = flops / (out["fused_ms_with_packing"] * 1e-3) / 1e12
    print(json.dumps(out))


if __name__ == "__main__":
    main()
