"""View-batched gradient accumulation on the GPU (the data-parallel path of SURVEY.md 8(e)):
with an SH gradient sink attached, the backward pass of every view writes / adds its SH gradient
straight into the sink buffer; the result must equal torch's own accumulation of the per-view
gradients, and every other gradient must be untouched by the sink."""
import pytest
import torch

import refutil
import scenes

pytestmark = pytest.mark.gpu


def _views(n_views, P_scene=30000, P_human=3000, W=320, H=240, focal=230.0, seed=21, degree=3):
    base = scenes.make_scene(P_scene, P_human, W, H, focal, seed, degree=degree, device="cuda")
    out = []
    for v in range(n_views):
        sc = scenes.make_scene(P_scene, P_human, W, H, focal, seed, degree=degree, device="cuda", view=v,
                               n_views=max(n_views, 2))
        # same Gaussians, different camera and loss gradient
        for k in ("means3D", "scales", "rotations", "opacity", "shs"):
            sc[k] = base[k]
        sc["dL_dcolor"] = torch.rand(3, H, W, device="cuda", generator=torch.Generator("cuda").manual_seed(v)) * 2 - 1
        out.append(sc)
    return out


def _render_all(views, leaves, means2D, module):
    for sc in views:
        rast = module.GaussianRasterizer(scenes.settings_of(sc, module))
        color, _ = rast(means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacity"], shs=leaves["shs"],
                        scales=leaves["scales"], rotations=leaves["rotations"])
        (color * sc["dL_dcolor"]).sum().backward()


def _leaves(sc):
    names = ["means3D", "scales", "rotations", "opacity", "shs"]
    return {k: sc[k].detach().clone().float().contiguous().requires_grad_(True) for k in names}


@pytest.mark.parametrize("degree", [3, 1])
def test_sink_equals_autograd_accumulation(degree):
    import diff_gaussian_rasterization as m
    import hugs_b200

    views = _views(3, degree=degree)
    ref = _leaves(views[0])
    _render_all(views, ref, torch.zeros_like(ref["means3D"], requires_grad=True), m)

    got = _leaves(views[0])
    # garbage in the buffer: the first backward pass after reset() must overwrite every row
    sink = hugs_b200.attach_sh_grad_sink(got["shs"], torch.full_like(got["shs"], float("nan")))
    try:
        _render_all(views, got, torch.zeros_like(got["means3D"], requires_grad=True), m)
        assert got["shs"].grad is sink.buffer
        assert torch.isfinite(sink.buffer).all()
        assert refutil.rel_err(sink.buffer.cpu().numpy(), ref["shs"].grad.cpu().numpy()) <= 1e-5
        for k in ("means3D", "scales", "rotations", "opacity"):
            # (the blend backward sums with atomics: equal up to the order of the additions)
            assert refutil.rel_err(got[k].grad.cpu().numpy(), ref[k].grad.cpu().numpy()) <= 1e-5, k
        # a second step after reset() starts from scratch
        first = sink.buffer.clone()
        sink.reset()
        for k in ("means3D", "scales", "rotations", "opacity"):
            got[k].grad = None
        _render_all(views, got, torch.zeros_like(got["means3D"], requires_grad=True), m)
        assert refutil.rel_err(sink.buffer.cpu().numpy(), first.cpu().numpy()) <= 1e-5
    finally:
        hugs_b200.detach_sh_grad_sink(got["shs"])


def test_sink_rows_of_culled_gaussians_stay_zero():
    import diff_gaussian_rasterization as m
    import hugs_b200

    views = _views(2)
    got = _leaves(views[0])
    sink = hugs_b200.attach_sh_grad_sink(got["shs"])
    try:
        radii = []
        for sc in views:
            rast = m.GaussianRasterizer(scenes.settings_of(sc, m))
            color, r = rast(means3D=got["means3D"], means2D=torch.zeros_like(got["means3D"], requires_grad=True),
                            opacities=got["opacity"], shs=got["shs"], scales=got["scales"], rotations=got["rotations"])
            (color * sc["dL_dcolor"]).sum().backward()
            radii.append(r)
        never = (radii[0] <= 0) & (radii[1] <= 0)
        assert int(never.sum()) > 0
        assert float(sink.buffer[never].abs().max()) == 0.0
    finally:
        hugs_b200.detach_sh_grad_sink(got["shs"])


def test_without_sink_nothing_changes():
    import diff_gaussian_rasterization as m

    views = _views(2)
    a = _leaves(views[0])
    _render_all(views[:1], a, torch.zeros_like(a["means3D"], requires_grad=True), m)
    b = _leaves(views[0])
    _render_all(views[:1], b, torch.zeros_like(b["means3D"], requires_grad=True), m)
    assert refutil.rel_err(a["shs"].grad.cpu().numpy(), b["shs"].grad.cpu().numpy()) <= 1e-5
