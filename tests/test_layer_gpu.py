"""The separate human image from the joint pass's lists (csrc/layer.cu, hugs_b200/layer.py)
against what it replaces: the reference's TWO render() calls of hugs/renderer/gs_renderer.py:56-82 on the
unmodified reference build -- joint image, human-only image against its own background, and the gradients
autograd sums over both.  Bars: contributor counts / ranges of the human-only pass bit-exact, images
<= 1e-4 max-abs, gradients <= 1e-3 (the three acceptance views of refutil.assert_grad_close)."""
import ctypes as C

import numpy as np
import pytest
import torch

import refutil
import scenes

pytestmark = pytest.mark.gpu

NAMES = ("means3D", "scales", "rotations", "opacity", "shs")


def _leaves(sc):
    return {k: sc[k].detach().clone().float().requires_grad_(True) for k in NAMES}


def _two_passes(sc, module, n_h, hbg, G_h):
    """The reference's structure: render(all) + render(human alone), both on the same leaf tensors."""
    t = _leaves(sc)
    settings = scenes.settings_of(sc, module)
    m2d = torch.zeros_like(t["means3D"], requires_grad=True)
    color, radii = module.GaussianRasterizer(settings)(
        means3D=t["means3D"], means2D=m2d, opacities=t["opacity"], shs=t["shs"], scales=t["scales"],
        rotations=t["rotations"])
    m2d_h = torch.zeros(n_h, 3, device=color.device, requires_grad=True)
    color_h, radii_h = module.GaussianRasterizer(settings._replace(bg=hbg))(
        means3D=t["means3D"][:n_h], means2D=m2d_h, opacities=t["opacity"][:n_h], shs=t["shs"][:n_h],
        scales=t["scales"][:n_h], rotations=t["rotations"][:n_h])
    node_h = color_h.grad_fn   # saved tensors are freed by backward: take the buffers now
    state_h = (node_h.saved_tensors[7], node_h.saved_tensors[8], node_h.saved_tensors[9], int(node_h.num_rendered))
    ((color * sc["dL_dcolor"]).sum() + (color_h * G_h).sum()).backward()
    grads = {k: t[k].grad for k in NAMES}
    grads["means2D"] = m2d.grad
    return dict(color=color.detach(), color_h=color_h.detach(), radii=radii, radii_h=radii_h, grads=grads,
                state_h=state_h)


def _ours(sc, n_h, hbg, G_h, two_segments):
    import diff_gaussian_rasterization as dgr
    from hugs_b200 import layer
    t = _leaves(sc)
    settings = scenes.settings_of(sc, dgr)
    m2d = torch.zeros_like(t["means3D"], requires_grad=True)
    if two_segments:
        h = {k: t[k][:n_h].detach().clone().requires_grad_(True) for k in NAMES}
        s = {k: t[k][n_h:].detach().clone().requires_grad_(True) for k in NAMES}
        seg = lambda d: (d["means3D"], d["shs"], d["opacity"], d["scales"], d["rotations"])  # noqa: E731
        color, color_h, radii = layer.rasterize_two_segments_with_layer(seg(h), seg(s), m2d, settings, hbg)
    else:
        color, color_h, radii = layer.rasterize_with_layer(t["means3D"], m2d, t["opacity"], t["scales"],
                                                           t["rotations"], settings, n_h, hbg, shs=t["shs"])
    saved = color.grad_fn.saved_tensors
    state = _layer_state(saved[2], saved[4], n_h, sc["image_width"], sc["image_height"])
    ((color * sc["dL_dcolor"]).sum() + (color_h * G_h).sum()).backward()
    if two_segments:
        grads = {k: torch.cat([h[k].grad, s[k].grad]) for k in NAMES}
    else:
        grads = {k: t[k].grad for k in NAMES}
    grads["means2D"] = m2d.grad
    return dict(color=color.detach(), color_h=color_h.detach(), radii=radii, grads=grads, state=state)


def _layer_state(binning, layer, n_h, W, H):
    """n_contrib / final_T / ranges of the layer, decoded from the layer buffer the autograd node keeps."""
    from hugs_b200 import _native as N
    offs = (C.c_size_t * 6)()
    assert N.layer_layout_of(n_h, int(binning.numel()), W, H, offs) == 6
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    take = lambda off, n, dt: layer[off:off + 4 * n].view(dt)  # noqa: E731
    return dict(accum_alpha=take(offs[2], W * H, torch.float32).clone(),
                n_contrib=take(offs[3], W * H, torch.int32).clone(),
                ranges=take(offs[4], 2 * tiles, torch.int32).reshape(tiles, 2).clone())


def _compare(sc, n_h, two_segments, full_checks=True):
    ref_mod = refutil.load_ref_rasterizer()
    W, H = sc["image_width"], sc["image_height"]
    dev = sc["means3D"].device
    hbg = torch.tensor([0.9, 0.2, 0.6], device=dev)
    G_h = torch.randn(3, H, W, generator=torch.Generator(device="cpu").manual_seed(5)).to(dev)
    ref = _two_passes(sc, ref_mod, n_h, hbg, G_h)
    our = _ours(sc, n_h, hbg, G_h, two_segments)
    assert torch.equal(our["radii"], ref["radii"])
    assert torch.equal(our["radii"][:n_h], ref["radii_h"])
    assert float((our["color"] - ref["color"]).abs().max()) <= 1e-4
    assert float((our["color_h"] - ref["color_h"]).abs().max()) <= 1e-4
    # state of the human-only pass: the reference's second call keeps it in ITS image buffer
    st = our["state"]
    g_h, b_h, i_h, R_h = ref["state_h"]
    rdec = refutil.decode_ref_buffers(g_h, b_h, i_h, n_h, R_h, W, H)
    assert torch.equal(st["n_contrib"], rdec["n_contrib"])
    assert float((st["accum_alpha"] - rdec["accum_alpha"]).abs().max()) <= 1e-6
    ours_len = (st["ranges"][:, 1] - st["ranges"][:, 0]).cpu()
    ref_len = (rdec["ranges"][:, 1] - rdec["ranges"][:, 0]).cpu()
    assert torch.equal(ours_len, ref_len)
    assert int(ours_len.sum()) == R_h
    for k in list(NAMES) + ["means2D"]:
        refutil.assert_grad_close(our["grads"][k].cpu().numpy(), ref["grads"][k].cpu().numpy(), name=k)


@pytest.mark.parametrize("two_segments", [False, True])
@pytest.mark.parametrize("n_s,n_h,W,H,focal,seed", [(2600, 1800, 320, 200, 300.0, 21), (0, 2500, 200, 120, 180.0, 22),
                                                    (60000, 20000, 640, 360, 500.0, 23)])
def test_layer_matches_the_reference_second_pass(n_s, n_h, W, H, focal, seed, two_segments):
    sc = scenes.make_scene(n_s, n_h, W, H, focal, seed=seed, device="cuda")
    _compare(sc, n_h, two_segments)


@pytest.mark.parametrize("view", [0, 3])
def test_layer_full_size_human_scene_1080p(view):
    sc = scenes.make_scene(1_000_000, 110_210, 1920, 1080, 1400.0, seed=1, device="cuda", view=view, n_views=8)
    _compare(sc, 110_210, two_segments=True)


def test_layer_of_nothing_and_of_everything():
    """n_layer = 0: the layer image is its background; n_layer = P: the layer image is the main image
    rendered against the other background."""
    import diff_gaussian_rasterization as dgr
    from hugs_b200 import layer
    sc = scenes.make_scene(2000, 1000, 256, 160, 220.0, seed=31, device="cuda")
    hbg = torch.tensor([0.1, 0.7, 0.3], device="cuda")
    settings = scenes.settings_of(sc, dgr)
    P = sc["means3D"].shape[0]
    args = (sc["means3D"], torch.zeros(P, 3, device="cuda"), sc["opacity"], sc["scales"], sc["rotations"], settings)
    color, lay, radii = layer.rasterize_with_layer(*args, 0, hbg, shs=sc["shs"])
    assert torch.equal(lay, hbg[:, None, None].expand(3, 160, 256))
    color2, lay2, _ = layer.rasterize_with_layer(*args, P, hbg, shs=sc["shs"])
    whole = dgr.GaussianRasterizer(settings._replace(bg=hbg))(
        means3D=sc["means3D"], means2D=torch.zeros(P, 3, device="cuda"), opacities=sc["opacity"], shs=sc["shs"],
        scales=sc["scales"], rotations=sc["rotations"])[0]
    assert torch.equal(color2, color) and torch.equal(lay2, whole)
    with pytest.raises(ValueError):
        layer.rasterize_with_layer(*args, P + 1, hbg, shs=sc["shs"])


def test_layer_with_precomputed_colors_and_unused_layer_gradient():
    import diff_gaussian_rasterization as dgr
    from hugs_b200 import layer
    n_h, n_s = 900, 1500
    sc = scenes.make_scene(n_s, n_h, 256, 160, 220.0, seed=32, device="cuda")
    P = n_h + n_s
    rgb = torch.rand(P, 3, device="cuda", generator=torch.Generator(device="cuda").manual_seed(3))
    hbg = torch.zeros(3, device="cuda")
    settings = scenes.settings_of(sc, dgr)

    def run(use_layer_grad):
        t = {k: sc[k].detach().clone().requires_grad_(True) for k in ("means3D", "scales", "rotations", "opacity")}
        c = rgb.clone().requires_grad_(True)
        m2d = torch.zeros(P, 3, device="cuda", requires_grad=True)
        color, lay, _ = layer.rasterize_with_layer(t["means3D"], m2d, t["opacity"], t["scales"], t["rotations"],
                                                   settings, n_h, hbg, colors_precomp=c)
        loss = (color * sc["dL_dcolor"]).sum()
        if use_layer_grad:
            loss = loss + lay.sum()
        loss.backward()
        return color.detach(), lay.detach(), c.grad, t["means3D"].grad, m2d.grad

    color, lay, gc, gm, g2 = run(True)
    color0, lay0, gc0, gm0, g20 = run(False)
    honly = dgr.GaussianRasterizer(settings._replace(bg=hbg))(
        means3D=sc["means3D"][:n_h], means2D=torch.zeros(n_h, 3, device="cuda"), opacities=sc["opacity"][:n_h],
        colors_precomp=rgb[:n_h], scales=sc["scales"][:n_h], rotations=sc["rotations"][:n_h])[0]
    assert torch.equal(lay, honly) and torch.equal(lay0, honly) and torch.equal(color, color0)
    # the layer's loss reaches the human rows only, and never the reported screen-space gradient
    # (atomics: two runs agree to summation order, not to the bit)
    n = lambda x: x.cpu().numpy()  # noqa: E731
    assert refutil.rel_err(n(gc[n_h:]), n(gc0[n_h:])) < 1e-5 and float((gc[:n_h] - gc0[:n_h]).abs().max()) > 1e-3
    assert refutil.rel_err(n(gm[n_h:]), n(gm0[n_h:])) < 1e-5
    assert refutil.rel_err(g2.cpu().numpy(), g20.cpu().numpy()) < 1e-5
