"""The decoder MLPs on tcgen05 (csrc/mlp_tc.cu, hugs_b200/decoders.py) against golden vectors generated from
the reference's own modules (tests/golden/make_golden_decoders.py imports
hugs/models/modules/decoders.py:24-110 on the CPU, float32).  Tolerances: forward outputs 2e-5 of the
output's largest magnitude (3xTF32 products, fp32 accumulation in a different order than the CPU GEMM);
gradients 1e-3 of the gradient's largest magnitude (BASELINE.json's bar for gradients): the backward kernel
rounds dZ and the activations to tf32 (nearest) for its two products -- measured 2e-4 .. 5e-4 on the 333-row
golden case and 2e-4 .. 5e-4 at the model's 110 210 rows (a weight gradient is a sum of zero-mean products:
it grows like the square root of the row count, and so does the rounding error -- the ratio stays ~2^-12 * sqrt 2)."""
import os

import numpy as np
import pytest
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "decoders_golden.npz")


def _golden():
    return dict(np.load(GOLD))


def _build(tag, g, device):
    from hugs_b200 import decoders as D
    m = {"appearance": lambda: D.AppearanceDecoder(n_features=96),
         "geometry": lambda: D.GeometryDecoder(n_features=96),
         "deformation": lambda: D.DeformationDecoder(n_features=96, disable_posedirs=True)}[tag]()
    sd = {k[len(tag) + 7:]: torch.from_numpy(v) for k, v in g.items() if k.startswith(tag + "/param/")}
    m.load_state_dict(sd)   # same keys as the reference's state_dict
    return m.to(device)


def test_state_dict_keys_are_the_reference_ones():
    g = _golden()
    for tag in ("appearance", "geometry", "deformation"):
        m = _build(tag, g, "cpu")
        want = sorted(k[len(tag) + 7:] for k in g if k.startswith(tag + "/param/"))
        assert sorted(m.state_dict().keys()) == want


def _close(a, b, tol):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max()) <= tol * max(float(np.abs(b).max()), 1e-30)


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["appearance", "geometry", "deformation"])
def test_decoder_forward_and_backward_match_reference_golden(tag):
    g = _golden()
    m = _build(tag, g, "cuda")
    x = torch.from_numpy(g["x"]).cuda().requires_grad_(True)
    res = m(x)
    keys = [k[len(tag) + 5:] for k in g if k.startswith(tag + "/out/")]
    assert sorted(k for k, v in res.items() if v is not None) == sorted(keys)
    loss = 0.0
    for k in keys:
        assert res[k].shape == g["%s/out/%s" % (tag, k)].shape
        assert _close(res[k].detach().cpu().numpy(), g["%s/out/%s" % (tag, k)], 2e-5), k
        loss = loss + (res[k] * torch.from_numpy(g["%s/w/%s" % (tag, k)]).cuda()).sum()
    loss.backward()
    assert _close(x.grad.cpu().numpy(), g[tag + "/dx"], 1e-3)
    for name, p in m.named_parameters():
        assert p.grad is not None, name
        assert _close(p.grad.cpu().numpy(), g["%s/grad/%s" % (tag, name)], 1e-3), name


@pytest.mark.gpu
def test_decode_all_is_one_launch_and_equals_the_separate_calls():
    from hugs_b200 import _native, decoders as D
    g = _golden()
    app, geo, dfm = (_build(t, g, "cuda") for t in ("appearance", "geometry", "deformation"))
    x = torch.from_numpy(g["x"]).cuda()
    sep = (app(x), geo(x), dfm(x))
    torch.cuda.synchronize()
    _native.launch_count(1)
    a, b, c = D.decode_all(x, app, geo, dfm)
    torch.cuda.synchronize()
    launches = _native.launch_count(0)
    assert launches == 10 + 1     # one weight-packing launch per (concatenated) Linear, ONE fused launch for the rest
    for got, want in zip((a, b, c), sep):
        for k, v in want.items():
            assert (got[k] is None and v is None) or torch.equal(got[k], v), k


@pytest.mark.gpu
@pytest.mark.parametrize("rows", [0, 1, 127, 128, 129, 110_210])
def test_decoder_row_counts(rows):
    """ragged tiles, and the model's own size against torch's fp32 evaluation of the same module"""
    g = _golden()
    geo = _build("geometry", g, "cuda")
    x = torch.randn(rows, 96, device="cuda", generator=torch.Generator(device="cuda").manual_seed(rows + 1))
    got = geo(x)
    h = geo.net(x)
    want = {"xyz": geo.xyz[1](h), "rotations": geo.rotations[1](h), "scales": torch.nn.functional.gelu(geo.scales[1](h))}
    for k, v in want.items():
        assert got[k].shape == v.shape
        if rows:
            assert _close(got[k].detach().cpu().numpy(), v.detach().cpu().numpy(), 2e-5), k


@pytest.mark.gpu
def test_decoder_full_size_gradients():
    """110 210 rows, all three decoders in one launch each way, against torch autograd through the same
    modules' fp32 arithmetic (torch.backends.cuda.matmul.allow_tf32 is off)."""
    from hugs_b200 import decoders as D
    assert not torch.backends.cuda.matmul.allow_tf32
    g = _golden()
    mods = [_build(t, g, "cuda") for t in ("appearance", "geometry", "deformation")]
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
            b = {"xyz": geo.xyz[1](h), "rotations": geo.rotations[1](h),
                 "scales": torch.nn.functional.gelu(geo.scales[1](h))}
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
        return [o.detach() for o in outs], grads

    outs_f, gf = run(True)
    outs_t, gt = run(False)
    for a, b in zip(outs_f, outs_t):
        assert _close(a.cpu().numpy(), b.cpu().numpy(), 2e-5)
    for k in gt:
        assert _close(gf[k].cpu().numpy(), gt[k].cpu().numpy(), 1e-3), k


@pytest.mark.gpu
def test_canon_forward_keys_and_shapes():
    """tri-plane gather -> decoders -> the dictionary of hugs_trimlp.py:226-238"""
    from hugs_b200 import decoders as D
    from hugs_b200.triplane import TriPlane
    g = _golden()
    app, geo, dfm = (_build(t, g, "cuda") for t in ("appearance", "geometry", "deformation"))
    tri = TriPlane(32, resX=64, resY=64, resZ=64).cuda()
    xyz = (torch.rand(5000, 3, device="cuda") - 0.5) * 0.9
    out = D.canon_forward(tri, xyz, app, geo, dfm, scaling_multiplier=0.5)
    assert set(out) == {"xyz_offsets", "scales", "rot6d_canon", "shs", "opacity", "lbs_weights", "posedirs"}
    assert out["shs"].shape == (5000, 16, 3) and out["opacity"].shape == (5000, 1)
    assert out["xyz_offsets"].shape == (5000, 3) and out["rot6d_canon"].shape == (5000, 6) and out["scales"].shape == (5000, 3)
    assert out["lbs_weights"].shape == (5000, 24) and out["posedirs"] is None
    assert torch.allclose(out["lbs_weights"].sum(-1), torch.ones(5000, device="cuda"), atol=1e-5)
    feats = tri(xyz)
    assert torch.equal(out["scales"], geo(feats)["scales"] * 0.5)
