"""Gradient sinks of the two-segment rasterizer (hugs_b200_grad_sinks; view-batched training): what the
per-Gaussian backward ADDS into the parameters' accumulation buffers over several views equals what
torch's AccumulateGrad sums from the plain call (floating-point summation order aside: 1e-5 of the
largest element), with and without the separate human image, with views on two CUDA streams."""
import pytest
import torch

import refutil
import scenes

pytestmark = pytest.mark.gpu
NAMES = ("means3D", "shs", "opacity", "scales", "rotations")


def _segments(sc, n_h):
    h = {k: sc[k][:n_h].detach().clone().contiguous().requires_grad_(True) for k in NAMES}
    s = {k: sc[k][n_h:].detach().clone().contiguous().requires_grad_(True) for k in NAMES}
    return h, s


def _render(sc, h, s, with_layer):
    import diff_gaussian_rasterization as dgr
    from hugs_b200 import layer, segments
    seg = lambda d: (d["means3D"], d["shs"], d["opacity"], d["scales"], d["rotations"])  # noqa: E731
    P = h["means3D"].shape[0] + s["means3D"].shape[0]
    m2d = torch.empty(P, 3, device="cuda", requires_grad=True)
    settings = scenes.settings_of(sc, dgr)
    if with_layer:
        hbg = torch.tensor([0.2, 0.9, 0.4], device="cuda")
        color, color_h, _ = layer.rasterize_two_segments_with_layer(seg(h), seg(s), m2d, settings, hbg)
        return (color * sc["dL_dcolor"]).sum() + (color_h * sc["dL_dcolor"].flip(2)).sum()
    color, _ = segments.rasterize_two_segments(seg(h), seg(s), m2d, settings)
    return (color * sc["dL_dcolor"]).sum()


@pytest.mark.parametrize("with_layer", [False, True])
@pytest.mark.parametrize("sunk", ["all", "scene_only", "sh_and_opacity"])
def test_sinks_accumulate_like_autograd(with_layer, sunk):
    import hugs_b200
    n_h, n_s, W, H = 4000, 9000, 320, 200
    views = [scenes.make_scene(n_s, n_h, W, H, 300.0, seed=61, device="cuda", view=v, n_views=3) for v in range(3)]
    # plain: autograd accumulates the three views
    h0, s0 = _segments(views[0], n_h)
    for sc in views:
        _render(sc, h0, s0, with_layer).backward()
    # sinks
    h1, s1 = _segments(views[0], n_h)
    chosen = {"all": [(h1, k) for k in NAMES] + [(s1, k) for k in NAMES],
              "scene_only": [(s1, k) for k in NAMES],
              "sh_and_opacity": [(h1, "shs"), (h1, "opacity"), (s1, "shs"), (s1, "opacity")]}[sunk]
    for d, k in chosen:
        hugs_b200.attach_grad_sink(d[k])
    try:
        streams = [torch.cuda.Stream(), torch.cuda.Stream()]
        main = torch.cuda.current_stream()
        for st in streams:
            st.wait_stream(main)
        for i, sc in enumerate(views):
            with torch.cuda.stream(streams[i % 2]):
                _render(sc, h1, s1, with_layer).backward()
        for st in streams:
            main.wait_stream(st)
        torch.cuda.synchronize()
        for d0, d1, tag in ((h0, h1, "human"), (s0, s1, "scene")):
            for k in NAMES:
                assert d1[k].grad is not None, (tag, k)
                err = refutil.rel_err(d1[k].grad.cpu().numpy(), d0[k].grad.cpu().numpy())
                assert err < 1e-5, (tag, k, err)
    finally:
        for d, k in chosen:
            hugs_b200.detach_grad_sink(d[k])


def test_flat_grad_buffer_with_sinks_and_an_empty_scene():
    import hugs_b200
    from hugs_b200 import dp
    n_h = 3000
    sc = scenes.make_scene(0, n_h, 256, 160, 220.0, seed=62, device="cuda")
    h0, s0 = _segments(sc, n_h)
    _render(sc, h0, s0, False).backward()
    h1, s1 = _segments(sc, n_h)
    buf = dp.FlatGradBuffer({"h_" + k: h1[k] for k in NAMES})
    buf.attach_sinks(["h_" + k for k in NAMES])
    try:
        buf.zero_()
        _render(sc, h1, s1, False).backward()
        _render(sc, h1, s1, False).backward()
        torch.cuda.synchronize()
        for k in NAMES:
            assert h1[k].grad.data_ptr() == buf.grads()["h_" + k].data_ptr()
            err = refutil.rel_err(h1[k].grad.cpu().numpy(), 2.0 * h0[k].grad.cpu().numpy())
            assert err < 1e-5, (k, err)
    finally:
        for k in NAMES:
            hugs_b200.detach_grad_sink(h1[k])


@pytest.mark.parametrize("n_views,two_streams,with_layer", [(3, False, False), (5, True, False), (3, True, True)])
def test_deferred_scene_backward_over_all_views(n_views, two_streams, with_layer):
    """hugs_b200.segments.defer_scene_backward / hugs_b200_scene_backward_views: the scene segment's per-Gaussian
    backward run once for all views of a step leaves the same sums in the sinks as one run per view, and the
    same per-view screen-space gradients (means2D.grad), scene rows included."""
    import diff_gaussian_rasterization as dgr
    import hugs_b200
    from hugs_b200 import segments
    n_h, n_s, W, H = 3000, 9000, 320, 200
    views = [scenes.make_scene(n_s, n_h, W, H, 300.0, seed=63, device="cuda", view=v, n_views=n_views)
             for v in range(n_views)]
    seg = lambda d: (d["means3D"], d["shs"], d["opacity"], d["scales"], d["rotations"])  # noqa: E731

    def run(deferred):
        h, s = _segments(views[0], n_h)
        for d in (h, s):
            for k in NAMES:
                hugs_b200.attach_grad_sink(d[k])
        m2ds = []
        try:
            streams = [torch.cuda.Stream(), torch.cuda.Stream()] if two_streams else [torch.cuda.current_stream()]
            main = torch.cuda.current_stream()
            for st in streams:
                st.wait_stream(main)
            ctx = segments.defer_scene_backward() if deferred else contextlib.nullcontext()
            with ctx:
                for i, sc in enumerate(views):
                    with torch.cuda.stream(streams[i % len(streams)]):
                        m2d = torch.empty(n_h + n_s, 3, device="cuda", requires_grad=True)
                        if with_layer:
                            from hugs_b200 import layer
                            hbg = torch.tensor([0.2, 0.9, 0.4], device="cuda")
                            color, color_h, _ = layer.rasterize_two_segments_with_layer(
                                seg(h), seg(s), m2d, scenes.settings_of(sc, dgr), hbg)
                            ((color * sc["dL_dcolor"]).sum() + (color_h * sc["dL_dcolor"].flip(2)).sum()).backward()
                        else:
                            color, _ = segments.rasterize_two_segments(seg(h), seg(s), m2d, scenes.settings_of(sc, dgr))
                            (color * sc["dL_dcolor"]).sum().backward()
                        m2ds.append(m2d)
                for st in streams:
                    main.wait_stream(st)
            torch.cuda.synchronize()
            return ({("h", k): h[k].grad.clone() for k in NAMES} | {("s", k): s[k].grad.clone() for k in NAMES},
                    [m.grad.clone() for m in m2ds])
        finally:
            for d in (h, s):
                for k in NAMES:
                    hugs_b200.detach_grad_sink(d[k])

    import contextlib
    g0, m0 = run(False)
    g1, m1 = run(True)
    for key in g0:
        assert float(g0[key].abs().max()) > 0, key
        err = refutil.rel_err(g1[key].cpu().numpy(), g0[key].cpu().numpy())
        assert err < 1e-5, (key, err)
    for a, b in zip(m0, m1):   # (the blend backward's reductions differ in order between the two runs)
        assert refutil.rel_err(b.cpu().numpy(), a.cpu().numpy()) < 1e-5
        assert torch.equal(a == 0, b == 0)


@pytest.mark.parametrize("with_layer", [False, True])
def test_scene_prepass_matches_the_per_view_preprocess(with_layer):
    """hugs_b200.segments.scene_prepass / hugs_b200_preprocess_scene_views: the scene segment preprocessed for
    all cameras of a step in one launch gives every view the same radii (bit for bit), the same image (to 1e-6:
    the SH evaluation is compiled into another kernel) and the same gradients as its own preprocess."""
    import diff_gaussian_rasterization as dgr
    from hugs_b200 import segments
    n_h, n_s, W, H, n_views = 3000, 9000, 320, 200, 5
    views = [scenes.make_scene(n_s, n_h, W, H, 300.0, seed=64, device="cuda", view=v, n_views=n_views)
             for v in range(n_views)]
    seg = lambda d: (d["means3D"], d["shs"], d["opacity"], d["scales"], d["rotations"])  # noqa: E731
    settings = [scenes.settings_of(sc, dgr) for sc in views]

    def run(prepass):
        h, s = _segments(views[0], n_h)
        out = []
        ctx = segments.scene_prepass(settings, seg(s), n_h) if prepass else contextlib.nullcontext()
        with ctx as pp:
            for sc, st in zip(views, settings):
                m2d = torch.empty(n_h + n_s, 3, device="cuda", requires_grad=True)
                if with_layer:
                    from hugs_b200 import layer
                    hbg = torch.tensor([0.2, 0.9, 0.4], device="cuda")
                    color, color_h, radii = layer.rasterize_two_segments_with_layer(seg(h), seg(s), m2d, st, hbg)
                    ((color * sc["dL_dcolor"]).sum() + (color_h * sc["dL_dcolor"].flip(2)).sum()).backward()
                    color = torch.cat([color, color_h])
                else:
                    color, radii = segments.rasterize_two_segments(seg(h), seg(s), m2d, st)
                    (color * sc["dL_dcolor"]).sum().backward()
                out.append((color.detach().clone(), radii.clone(), m2d.grad.clone()))
            if prepass:
                assert not pp.entries   # every view picked its workspace up
        torch.cuda.synchronize()
        return out, {("h", k): h[k].grad.clone() for k in NAMES} | {("s", k): s[k].grad.clone() for k in NAMES}

    import contextlib
    o0, g0 = run(False)
    o1, g1 = run(True)
    for (c0, r0, m0), (c1, r1, m1) in zip(o0, o1):
        assert torch.equal(r0, r1)
        assert float((c0 - c1).abs().max()) < 1e-6
        assert refutil.rel_err(m1.cpu().numpy(), m0.cpu().numpy()) < 1e-5
    for key in g0:
        assert refutil.rel_err(g1[key].cpu().numpy(), g0[key].cpu().numpy()) < 1e-5, key
