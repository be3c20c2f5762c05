"""Renderer glue + post-ops against a torch restatement of the reference's own lines
(hugs/renderer/gs_renderer.py:19-161, hugs/trainer/gs_trainer.py:406-411, hugs/models/scene.py:460-462).
The densification statistics are floating point: tolerance 1e-6 relative on the accumulated norms;
the max radii and the counters are exact."""
import pytest
import torch

import refutil  # noqa: F401
import scenes

pytestmark = pytest.mark.gpu


def _data_of(sc):
    return dict(image_height=sc["image_height"], image_width=sc["image_width"], fovx=sc["fovx"], fovy=sc["fovy"],
                world_view_transform=sc["viewmatrix"], full_proj_transform=sc["projmatrix"],
                camera_center=sc["campos"])


def _gs_out(sc, lo, hi, grad=True):
    names = dict(xyz="means3D", shs="shs", opacity="opacity", scales="scales", rotq="rotations")
    out = {k: sc[v][lo:hi].detach().clone().requires_grad_(grad) for k, v in names.items()}
    out["active_sh_degree"] = sc["degree"]
    return out


def test_render_is_rasterizer_plus_clamp_and_filter():
    import diff_gaussian_rasterization as dgr
    from hugs_b200 import renderer
    sc = scenes.make_scene(3000, 2000, 320, 200, 300.0, seed=11, device="cuda")
    ref = scenes.run_module(sc, dgr, backward=False)
    pkg = renderer.render(sc["means3D"], sc["shs"], sc["opacity"], sc["scales"], sc["rotations"], _data_of(sc),
                          bg_color=sc["bg"], active_sh_degree=sc["degree"])
    assert set(pkg) == {"render", "viewspace_points", "visibility_filter", "radii"}
    assert torch.equal(pkg["render"], ref["color"].clamp(0.0, 1.0))
    assert torch.equal(pkg["radii"], ref["radii"])
    assert torch.equal(pkg["visibility_filter"], ref["radii"] > 0)
    assert pkg["viewspace_points"].shape == sc["means3D"].shape


def test_render_precomputed_colors_and_default_background():
    from hugs_b200 import renderer
    sc = scenes.make_scene(1500, 0, 160, 120, 150.0, seed=12, device="cuda")
    rgb = torch.rand(1500, 3, device="cuda")
    pkg = renderer.render(sc["means3D"], rgb, sc["opacity"], sc["scales"], sc["rotations"], _data_of(sc))
    assert pkg["render"].shape == (3, 120, 160)
    assert float(pkg["render"].min()) >= 0.0 and float(pkg["render"].max()) <= 1.0


@pytest.mark.parametrize("separate", [False, True])
def test_render_human_scene_modes(separate):
    from hugs_b200 import renderer
    n_h, n_s = 1800, 2600
    sc = scenes.make_scene(n_s, n_h, 320, 200, 300.0, seed=13, device="cuda")
    data = _data_of(sc)
    human, scene = _gs_out(sc, 0, n_h), _gs_out(sc, n_h, n_h + n_s)
    hbg = torch.tensor([1.0, 1.0, 1.0], device="cuda")
    pkg = renderer.render_human_scene(data, human, scene, sc["bg"], human_bg_color=hbg,
                                      render_mode="human_scene", render_human_separate=separate)
    whole = renderer.render(sc["means3D"], sc["shs"], sc["opacity"], sc["scales"], sc["rotations"], data,
                            bg_color=sc["bg"], active_sh_degree=sc["degree"])
    assert torch.equal(pkg["render"], whole["render"])
    assert torch.equal(pkg["scene_radii"], whole["radii"][n_h:])
    assert torch.equal(pkg["scene_visibility_filter"], whole["radii"][n_h:] > 0)
    if separate:
        honly = renderer.render(human["xyz"], human["shs"], human["opacity"], human["scales"], human["rotq"], data,
                                bg_color=hbg, active_sh_degree=sc["degree"])
        assert torch.equal(pkg["human_img"], honly["render"])
        assert torch.equal(pkg["human_radii"], honly["radii"])
    else:
        assert "human_img" not in pkg
        assert torch.equal(pkg["human_radii"], whole["radii"][:n_h])
    # gradients reach both models through the concatenation, and viewspace_points keeps its .grad
    loss = (pkg["render"] * sc["dL_dcolor"]).sum()
    if separate:
        loss = loss + pkg["human_img"].sum()
    loss.backward()
    assert pkg["viewspace_points"].grad is not None and pkg["viewspace_points"].grad.shape == (n_h + n_s, 3)
    for k in ("xyz", "shs", "opacity", "scales", "rotq"):
        assert human[k].grad is not None and scene[k].grad is not None
        assert float(human[k].grad.abs().sum()) > 0 and float(scene[k].grad.abs().sum()) > 0

    only_h = renderer.render_human_scene(data, human, scene, sc["bg"], render_mode="human")
    assert torch.equal(only_h["human_radii"], only_h["radii"]) and only_h["radii"].shape[0] == n_h
    only_s = renderer.render_human_scene(data, human, scene, sc["bg"], render_mode="scene")
    assert torch.equal(only_s["scene_visibility_filter"], only_s["radii"] > 0) and only_s["radii"].shape[0] == n_s
    with pytest.raises(ValueError, match="Unknown render mode"):
        renderer.render_human_scene(data, human, scene, sc["bg"], render_mode="both")


def _reference_stats(radii, grad, acc, den, mx):
    """The reference's lines, verbatim semantics (boolean-mask indexing)."""
    vis = radii > 0
    mx[vis] = torch.max(mx[vis], radii[vis])
    acc[vis] += torch.norm(grad[:vis.shape[0]][vis, :2], dim=-1, keepdim=True)
    den[vis] += 1


@pytest.mark.parametrize("n,n_grad", [(0, 0), (1, 1), (1000, 1000), (70001, 90000), (1_000_003, 1_000_003)])
def test_densification_stats_match_reference_lines(n, n_grad):
    from hugs_b200 import renderer
    g = torch.Generator(device="cuda").manual_seed(n + 1)
    radii = torch.randint(-1, 40, (n,), generator=g, device="cuda", dtype=torch.int32).clamp_min(0)
    grad = torch.randn(n_grad, 3, generator=g, device="cuda") * 1e-3
    acc = torch.rand(n, 1, generator=g, device="cuda")
    den = torch.randint(0, 9, (n, 1), generator=g, device="cuda").float()
    mx = torch.randint(0, 40, (n,), generator=g, device="cuda").float()
    r_acc, r_den, r_mx = acc.clone(), den.clone(), mx.clone()
    _reference_stats(radii, grad, r_acc, r_den, r_mx)
    renderer.add_densification_stats(radii, grad, xyz_gradient_accum=acc, denom=den, max_radii2D=mx)
    assert torch.equal(mx, r_mx)
    assert torch.equal(den, r_den)
    assert torch.allclose(acc, r_acc, rtol=1e-6, atol=0.0)
    invisible = radii <= 0
    if n:
        assert torch.equal(acc[invisible], r_acc[invisible])


def test_densification_stats_after_a_render_on_slices():
    """Scene statistics from the scene slice of radii and the FIRST n_scene gradient rows (the
    reference's grad[:update_filter.shape[0]], scene.py:461), human ones from the human slice."""
    from hugs_b200 import renderer
    n_h, n_s = 1200, 2200
    sc = scenes.make_scene(n_s, n_h, 320, 200, 300.0, seed=14, device="cuda")
    human, scene = _gs_out(sc, 0, n_h), _gs_out(sc, n_h, n_h + n_s)
    pkg = renderer.render_human_scene(_data_of(sc), human, scene, sc["bg"])
    (pkg["render"] * sc["dL_dcolor"]).sum().backward()
    grad = pkg["viewspace_points"].grad
    for radii, n in ((pkg["scene_radii"], n_s), (pkg["human_radii"], n_h)):
        acc, den, mx = (torch.zeros(n, 1, device="cuda"), torch.zeros(n, 1, device="cuda"),
                        torch.zeros(n, device="cuda"))
        r = [acc.clone(), den.clone(), mx.clone()]
        _reference_stats(radii, grad, *r)
        renderer.add_densification_stats(radii, grad, acc, den, mx)
        assert torch.equal(mx, r[2]) and torch.equal(den, r[1])
        assert torch.allclose(acc, r[0], rtol=1e-6, atol=0.0)
        assert float(den.sum()) > 0


def test_densification_stats_optional_outputs_and_errors():
    from hugs_b200 import renderer
    radii = torch.tensor([3, 0, 7], dtype=torch.int32, device="cuda")
    mx = torch.tensor([5.0, 9.0, 1.0], device="cuda")
    renderer.add_densification_stats(radii, None, max_radii2D=mx)
    assert mx.tolist() == [5.0, 9.0, 7.0]
    den = torch.zeros(3, device="cuda")
    renderer.add_densification_stats(radii, None, denom=den)
    assert den.tolist() == [1.0, 0.0, 1.0]
    with pytest.raises(RuntimeError, match="viewspace_point_grad"):
        renderer.add_densification_stats(radii, torch.zeros(2, 3, device="cuda"),
                                         xyz_gradient_accum=torch.zeros(3, device="cuda"))
    with pytest.raises(RuntimeError, match="int32"):
        renderer.add_densification_stats(radii.float(), None, denom=den)
    with pytest.raises(RuntimeError, match="float32"):
        renderer.add_densification_stats(radii, None, denom=den.double())


def test_checkpoint_and_ply_assets_render_like_the_tensors_they_hold(tmp_path):
    """scene_*.pth / *_splat.ply -> scene_gs_out -> render_human_scene(mode='scene') gives the image of
    the same Gaussians passed directly (N4 loaders feeding the path)."""
    from hugs_b200 import checkpoint, ply, renderer
    n = 4000
    sc = scenes.make_scene(n, 0, 320, 200, 300.0, seed=15, device="cuda")
    raw = dict(xyz=sc["means3D"], features_dc=sc["shs"][:, :1].contiguous(), features_rest=sc["shs"][:, 1:].contiguous(),
               scaling=torch.log(sc["scales"]), rotation=sc["rotations"] * 1.7,
               opacity=torch.log(sc["opacity"] / (1 - sc["opacity"])))
    p_ckpt, p_ply = str(tmp_path / "scene_final.pth"), str(tmp_path / "scene_final_splat.ply")
    checkpoint.save_scene_checkpoint(p_ckpt, checkpoint.make_scene_state(**{k: v.cpu() for k, v in raw.items()}))
    ply.save_ply(p_ply, raw["xyz"], raw["features_dc"], raw["features_rest"], raw["opacity"], raw["scaling"],
                 raw["rotation"])
    direct = checkpoint.scene_gs_out_from_state(raw)
    want = renderer.render_human_scene(_data_of(sc), None, direct, sc["bg"], render_mode="scene")["render"]
    from_ckpt, _ = checkpoint.load_scene_checkpoint(p_ckpt)
    from_ply = checkpoint.scene_gs_out_from_ply(p_ply)
    for gs in (from_ckpt, from_ply):
        got = renderer.render_human_scene(_data_of(sc), None, gs, sc["bg"], render_mode="scene")
        assert torch.equal(got["render"], want)
        assert int(got["scene_visibility_filter"].sum()) > 0


def test_stats_sync_single_process_two_views():
    """dp.DensificationStatsSync on the GPU (world size 1: no exchange): two views of a step give the
    statistics of the reference's lines applied view after view."""
    from hugs_b200 import dp
    n = 5000
    g = torch.Generator(device="cuda").manual_seed(21)
    acc, den, mx = torch.zeros(n, 1, device="cuda"), torch.zeros(n, 1, device="cuda"), torch.zeros(n, device="cuda")
    r = [acc.clone(), den.clone(), mx.clone()]
    sync = dp.DensificationStatsSync(n, device="cuda")
    sync.begin_step()
    for _ in range(2):
        radii = torch.randint(-5, 50, (n,), generator=g, device="cuda", dtype=torch.int32).clamp_min(0)
        grad = torch.randn(n, 3, generator=g, device="cuda")
        sync.add_view(radii, grad, mx)
        _reference_stats(radii, grad, *r)
    sync.finish_step(acc, den, mx)
    assert torch.equal(mx, r[2]) and torch.equal(den, r[1])
    assert torch.allclose(acc, r[0], rtol=1e-6, atol=0.0)


@pytest.mark.parametrize("n_h,n_s", [(1800, 2600), (0, 3000), (2500, 0), (33, 1), (110210, 200000)])
def test_two_segments_equal_the_concatenation(n_h, n_s):
    """hugs_b200_rasterize_*_seg2: image, radii and num_rendered bit-identical to the rasterizer on
    torch.cat(human, scene); every gradient within 1e-4 of its largest element (the blend backward's
    atomics are unordered in both; two runs of the SAME call differ by as much)."""
    import diff_gaussian_rasterization as dgr
    from hugs_b200.segments import rasterize_two_segments
    big = n_h + n_s > 100000
    W, H, focal = (1920, 1080, 1400.0) if big else (320, 200, 300.0)
    sc = scenes.make_scene(n_s, n_h, W, H, focal, seed=16, device="cuda")
    settings = scenes.settings_of(sc, dgr)
    names = ("means3D", "shs", "opacity", "scales", "rotations")
    whole = [sc[k].detach().clone().requires_grad_(True) for k in names]
    m2d_w = torch.zeros(n_h + n_s, 3, device="cuda", requires_grad=True)
    color_w, radii_w = dgr.GaussianRasterizer(settings)(
        means3D=whole[0], means2D=m2d_w, shs=whole[1], opacities=whole[2], scales=whole[3], rotations=whole[4])
    (color_w * sc["dL_dcolor"]).sum().backward()

    seg0 = [sc[k][:n_h].detach().clone().requires_grad_(True) for k in names]
    seg1 = [sc[k][n_h:].detach().clone().requires_grad_(True) for k in names]
    m2d = torch.zeros(n_h + n_s, 3, device="cuda", requires_grad=True)
    color, radii = rasterize_two_segments(seg0, seg1, m2d, settings)
    assert torch.equal(color, color_w) and torch.equal(radii, radii_w)
    assert color.grad_fn.num_rendered == color_w.grad_fn.num_rendered
    (color * sc["dL_dcolor"]).sum().backward()

    def close(a, b, what):
        scale, err = float(b.abs().max()), float((a - b).abs().max())
        assert err <= 1e-4 * max(scale, 1e-30), "%s: max|diff| %.3e against max|grad| %.3e" % (what, err, scale)
    close(m2d.grad, m2d_w.grad, "means2D")
    for k, w, a, b in zip(names, whole, seg0, seg1):
        if n_h:
            close(a.grad, w.grad[:n_h], k + " (segment 0)")
        if n_s:
            close(b.grad, w.grad[n_h:], k + " (segment 1)")


def test_two_segments_argument_errors():
    import diff_gaussian_rasterization as dgr
    from hugs_b200.segments import rasterize_two_segments
    sc = scenes.make_scene(40, 0, 64, 64, 60.0, seed=17, device="cuda")
    settings = scenes.settings_of(sc, dgr)
    names = ("means3D", "shs", "opacity", "scales", "rotations")
    seg = [sc[k] for k in names]
    m2d = torch.zeros(80, 3, device="cuda")
    bad = list(seg)
    bad[3] = seg[3][:10]
    with pytest.raises(RuntimeError, match="row count"):
        rasterize_two_segments(seg, bad, m2d, settings)
    bad = list(seg)
    bad[1] = seg[1][:, :9].contiguous()
    with pytest.raises(RuntimeError, match="same M"):
        rasterize_two_segments(seg, bad, m2d, settings)
    with pytest.raises(RuntimeError, match="no CPU path"):
        rasterize_two_segments([x.cpu() for x in seg], seg, m2d, settings)


@pytest.mark.parametrize("separate", [False, True])
def test_fused_clamp_equals_torch_clamp_forward_and_backward(separate):
    """hugs/renderer/gs_renderer.py:153: the clamp inside the blend kernels (HUGS_B200_FLAG_CLAMP01) gives the
    image of torch.clamp bit for bit and its gradient mask (SH colours exceed [0, 1] on many pixels of this
    scene, a zero background puts others exactly on 0)."""
    import diff_gaussian_rasterization as dgr
    from hugs_b200 import layer, renderer
    n_h, n_s = 1500, 2500
    sc = scenes.make_scene(n_s, n_h, 320, 200, 300.0, seed=15, device="cuda")
    sc["shs"] = sc["shs"] * 2.5          # push colours out of range
    sc["bg"] = torch.zeros(3, device="cuda")
    data = _data_of(sc)
    human, scene = _gs_out(sc, 0, n_h), _gs_out(sc, n_h, n_h + n_s)
    hbg = torch.tensor([1.0, 0.0, 0.5], device="cuda")
    pkg = renderer.render_human_scene(data, human, scene, sc["bg"], human_bg_color=hbg, render_human_separate=separate)
    # the same through the unfused route: raw images + torch.clamp
    human2, scene2 = _gs_out(sc, 0, n_h), _gs_out(sc, n_h, n_h + n_s)
    settings = scenes.settings_of(sc, dgr)
    seg = lambda o: (o['xyz'], o['shs'], o['opacity'], o['scales'], o['rotq'])  # noqa: E731
    m2d = torch.zeros(n_h + n_s, 3, device="cuda", requires_grad=True)
    raw, raw_h, _ = layer.rasterize_two_segments_with_layer(seg(human2), seg(scene2), m2d, settings, hbg)
    want, want_h = raw.clamp(0.0, 1.0), raw_h.clamp(0.0, 1.0)
    assert float((raw < 0).float().mean()) + float((raw > 1).float().mean()) > 0.01   # the clamp does something
    assert torch.equal(pkg["render"], want)
    loss, loss2 = (pkg["render"] * sc["dL_dcolor"]).sum(), (want * sc["dL_dcolor"]).sum()
    if separate:
        assert torch.equal(pkg["human_img"], want_h)
        loss, loss2 = loss + (pkg["human_img"] * sc["dL_dcolor"].flip(1)).sum(), loss2 + (want_h * sc["dL_dcolor"].flip(1)).sum()
    loss.backward()
    loss2.backward()
    for k in ("xyz", "shs", "opacity", "scales", "rotq"):
        for a, b in ((human[k], human2[k]), (scene[k], scene2[k])):
            assert refutil.rel_err(a.grad.cpu().numpy(), b.grad.cpu().numpy()) < 1e-5, k
    assert refutil.rel_err(pkg["viewspace_points"].grad.cpu().numpy(), m2d.grad.cpu().numpy()) < 1e-5
