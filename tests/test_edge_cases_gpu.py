"""GPU edge cases of the drop-in boundary, checked against the CPU oracle (no reference build
needed): empty input, everything culled, SH degree / coefficient-count variants, strided inputs,
a non-default stream, debug mode, markVisible."""
import numpy as np
import pytest
import torch

import refutil
import scenes

pytestmark = pytest.mark.gpu


def ours():
    import diff_gaussian_rasterization as m
    return m


def n(t):
    return t.detach().cpu().numpy()


def _vs_oracle(sc, backward=True):
    o = scenes.run_module(sc, ours(), backward=backward)
    orc = scenes.run_oracle(sc, backward=backward)
    fw = orc["fw"]
    W, H = sc["image_width"], sc["image_height"]
    P = sc["means3D"].shape[0]
    R = o["num_rendered"]
    assert fw["binning"]["num_rendered"] == R
    d = refutil.decode_our_buffers(o["geom"], o["binning"], o["img"], P, R, W, H)
    assert (n(o["radii"]) == fw["pre"]["radii"]).all()
    if R > 0:
        assert (n(d["point_list_keys"]).view(np.uint64) == fw["binning"]["keys"]).all()
        assert (n(d["point_list"]).view(np.uint32) == fw["binning"]["point_list"]).all()
    assert (n(d["ranges"]).view(np.uint32) == fw["binning"]["ranges"]).all()
    bad = int((n(d["n_contrib"]).view(np.uint32) != fw["n_contrib"]).sum())
    assert bad <= max(1, int(2e-5 * W * H))   # CPU expf vs CUDA expf, see oracle header
    assert np.abs(n(o["color"]) - fw["color"]).max() <= 1e-4
    if backward:
        pairs = dict(means3D="dL_dmean3D", means2D="dL_dmean2D", opacity="dL_dopacity", shs="dL_dsh",
                     scales="dL_dscale", rotations="dL_drot")
        for k, v in pairs.items():
            refutil.assert_grad_close(n(o["grads"][k]), orc["grads"][v].reshape(n(o["grads"][k]).shape), k)
    return o


def test_empty_input():
    m = ours()
    sc = scenes.make_scene(100, 0, 64, 48, 60.0, 1, device="cuda")
    z = lambda *s: torch.zeros(*s, device="cuda", requires_grad=True)  # noqa: E731
    means = z(0, 3)
    rast = m.GaussianRasterizer(scenes.settings_of(sc, m))
    color, radii = rast(means3D=means, means2D=z(0, 3), opacities=z(0, 1), shs=z(0, 16, 3), scales=z(0, 3),
                        rotations=z(0, 4))
    assert color.shape == (3, 48, 64) and radii.shape == (0,) and radii.dtype == torch.int32
    assert float(color.abs().max()) == 0.0   # rasterize_points.cu:81: nothing is launched for P == 0
    color.sum().backward()
    assert means.grad.shape == (0, 3)


def test_everything_culled_gives_background():
    sc = scenes.make_scene(500, 0, 80, 64, 70.0, 2, device="cuda")
    sc["means3D"][:, 2] = -sc["means3D"][:, 2].abs() - 1.0
    o = _vs_oracle(sc)
    assert o["num_rendered"] == 0 and int((o["radii"] > 0).sum()) == 0
    bg = sc["bg"]
    assert torch.allclose(o["color"], bg[:, None, None].expand_as(o["color"]))
    for g in o["grads"].values():
        assert float(g.abs().max()) == 0.0


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_sh_degrees(deg):
    sc = scenes.make_scene(3000, 500, 96, 80, 90.0, 50 + deg, degree=deg, device="cuda")
    o = _vs_oracle(sc)
    nz = (deg + 1) ** 2
    if nz < 16:
        assert float(o["grads"]["shs"][:, nz:].abs().max()) == 0.0   # bands above the active degree get zeros


@pytest.mark.parametrize("M,deg", [(4, 1), (9, 2), (1, 0)])
def test_sh_coefficient_counts_other_than_16(M, deg):
    sc = scenes.make_scene(2000, 0, 96, 64, 80.0, 60 + M, degree=deg, device="cuda")
    sc["shs"] = sc["shs"][:, :M].contiguous()
    _vs_oracle(sc)


def test_strided_inputs_and_side_stream():
    m = ours()
    sc = scenes.make_scene(4000, 0, 128, 96, 100.0, 71, device="cuda")
    base = scenes.run_module(sc, m, backward=False)
    sc2 = dict(sc)
    wide = torch.zeros(sc["means3D"].shape[0], 6, device="cuda")
    wide[:, ::2] = sc["means3D"]
    sc2["means3D"] = wide[:, ::2]               # non-contiguous view
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        alt = scenes.run_module(sc2, m, backward=False)
    s.synchronize()
    assert torch.equal(alt["color"], base["color"]) and torch.equal(alt["radii"], base["radii"])


def test_debug_mode_and_determinism_of_forward():
    m = ours()
    sc = scenes.make_scene(3000, 0, 128, 96, 100.0, 72, device="cuda")
    a = scenes.run_module(sc, m, backward=False)
    rast = m.GaussianRasterizer(scenes.settings_of(sc, m, debug=True))
    color, radii = rast(means3D=sc["means3D"], means2D=torch.zeros_like(sc["means3D"]), opacities=sc["opacity"],
                        shs=sc["shs"], scales=sc["scales"], rotations=sc["rotations"])
    assert torch.equal(color, a["color"]) and torch.equal(radii, a["radii"])


def test_mark_visible_matches_oracle():
    from oracle import oracle as O

    m = ours()
    sc = scenes.make_scene(5000, 0, 128, 96, 100.0, 73, device="cuda")
    rast = m.GaussianRasterizer(scenes.settings_of(sc, m))
    vis = rast.markVisible(sc["means3D"])
    assert vis.dtype == torch.bool and vis.shape == (5000,)
    assert (n(vis) == O.mark_visible(n(sc["means3D"]), n(sc["viewmatrix"]))).all()
    assert 0 < int(vis.sum()) < 5000


def test_finite_difference_on_sh_dc_coefficient():
    """Independent of any reference: the image is exactly linear in an unclamped SH DC coefficient, so
    d(loss)/d(shs[i,0,c]) must equal a central finite difference."""
    m = ours()
    sc = scenes.make_scene(60, 0, 48, 48, 50.0, 74, degree=1, device="cuda")
    sc["scales"] *= 6.0
    sc["shs"][:, 0, :] = sc["shs"][:, 0, :].abs() + 2.0     # keep colours away from the clamp at 0
    o = scenes.run_module(sc, m)
    g = o["grads"]["shs"]
    idx = torch.nonzero(o["radii"] > 0).flatten()[:6]
    assert idx.numel() > 0
    for i in idx.tolist():
        for c in range(3):
            vals = []
            for eps in (0.25, -0.25):
                sc2 = dict(sc)
                sh = sc["shs"].clone()
                sh[i, 0, c] += eps
                sc2["shs"] = sh
                col = scenes.run_module(sc2, m, backward=False)["color"]
                vals.append(float((col.double() * sc["dL_dcolor"].double()).sum()))
            fd = (vals[0] - vals[1]) / 0.5
            assert abs(fd - float(g[i, 0, c])) <= 2e-3 * max(abs(fd), 1e-2), (i, c, fd, float(g[i, 0, c]))


def test_workspaces_are_freed_without_the_cyclic_gc():
    """The three workspaces (hundreds of MB at full size) must die by reference counting: with the
    cyclic collector off (bench.py pauses it while timing), device memory must not grow per call."""
    import gc

    import diff_gaussian_rasterization as dgr
    import scenes

    sc = scenes.make_scene(20000, 2000, 320, 240, 250.0, 5, device="cuda")

    def once():
        o = scenes.run_module(sc, dgr, backward=True)
        del o

    once()
    gc.collect()
    gc.disable()
    try:
        torch.cuda.synchronize()
        base = torch.cuda.memory_allocated()
        for _ in range(10):
            once()
        torch.cuda.synchronize()
        grown = torch.cuda.memory_allocated() - base
    finally:
        gc.enable()
    assert grown < (1 << 20), "device memory grew by %d bytes over 10 calls with the cyclic GC off" % grown


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_two_devices_in_one_process():
    """Per-device state of the library (count events, stage-timer event pools, shared-memory opt-ins): a
    process that rasterizes on cuda:0 and then on cuda:1 gets the same image on both."""
    import diff_gaussian_rasterization as dgr
    imgs = []
    for dev in ("cuda:0", "cuda:1"):
        sc = scenes.make_scene(3000, 1000, 320, 200, 300.0, seed=77, device=dev)
        with torch.cuda.device(dev):
            out = scenes.run_module(sc, dgr, backward=True)
            torch.cuda.synchronize(dev)
        imgs.append(out["color"].cpu())
    assert torch.equal(imgs[0], imgs[1])
