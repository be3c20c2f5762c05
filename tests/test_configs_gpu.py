"""The two inference configurations of BASELINE.json at full size:
configs[3]  3 M scene Gaussians at 1920x1080, forward only;
configs[4]  110 210 human Gaussians x pose frames: batched LBS + forward render, frames sharded
            over ranks with no collective.
Bit-exact against the reference build where it is present, size-independent properties always."""
import pytest
import torch

import refutil
import scenes

pytestmark = pytest.mark.gpu


def ours():
    import diff_gaussian_rasterization as m
    return m


def bits(t):
    return t.contiguous().view(torch.int32)


def _structure_checks(sc, o, W, H):
    P = sc["means3D"].shape[0]
    R = o["num_rendered"]
    d = refutil.decode_our_buffers(o["geom"], o["binning"], o["img"], P, R, W, H)
    keys = d["point_list_keys"]
    assert bool((keys[1:] >= keys[:-1]).all())                       # sortedness
    rg = d["ranges"].long()
    assert int((rg[:, 1] - rg[:, 0]).sum()) == R                      # ranges partition the list
    cnt = torch.bincount(d["point_list"].long(), minlength=P)
    assert torch.equal(cnt.int(), d["tiles_touched"] * (o["radii"] > 0).int())  # every pair emitted once
    # contributor counts never exceed the tile's list length
    tiles_x = (W + 15) // 16
    ty = torch.arange(H, device="cuda") // 16
    tx = torch.arange(W, device="cuda") // 16
    tile = (ty[:, None] * tiles_x + tx[None, :]).flatten()
    length = (rg[:, 1] - rg[:, 0])[tile]
    assert bool((d["n_contrib"].long().flatten() <= length).all())
    assert bool(torch.isfinite(o["color"]).all())
    return d


def test_scene_3M_forward_only_1080p():
    W, H = 1920, 1080
    sc = scenes.make_scene(3_000_000, 0, W, H, 1400.0, 1003, degree=3, device="cuda")
    o = scenes.run_module(sc, ours(), backward=False)
    d = _structure_checks(sc, o, W, H)
    if refutil.ref_available():
        r = scenes.run_module(sc, refutil.load_ref_rasterizer(), backward=False)
        R = r["num_rendered"]
        assert o["num_rendered"] == R
        rd = refutil.decode_ref_buffers(r["geom"], r["binning"], r["img"], sc["means3D"].shape[0], R, W, H)
        assert torch.equal(o["radii"], r["radii"])
        assert torch.equal(d["point_list_keys"], rd["point_list_keys"])
        assert torch.equal(d["point_list"], rd["point_list"])
        assert torch.equal(d["ranges"], rd["ranges"])
        assert torch.equal(d["n_contrib"], rd["n_contrib"])
        assert float((o["color"] - r["color"]).abs().max()) <= 1e-4


def test_pose_animation_batched_lbs_and_render():
    """A shard of the 300-frame animation: batched LBS of all frames in one launch, then one forward
    render per frame; every frame equals the per-frame path (LBS called alone) bit for bit, and the
    reference build on the same deformed Gaussians where present."""
    import hugs_b200
    from hugs_b200 import dp, synth

    F_total, world = 300, 8
    frames = dp.shard_views(F_total, 3, world)[:4]          # rank 3's first frames: 3, 11, 19, 27
    assert frames == [3, 11, 19, 27]
    W, H = 1920, 1080
    cam = synth.make_camera(W, H, 1400.0, seed=1004, device="cuda")
    h = synth.make_human_gaussians(synth.N_HUMAN, seed=1005, n_frames=max(frames) + 1, device="cuda")
    A = h["A"][frames].contiguous()
    transl = h["transl"][frames].contiguous()
    bx, bq, bs = hugs_b200.lbs_deform_frames(A, h["xyz"], h["lbs_weights"], h["rotmat"], h["scales"], transl, 1.0)
    m = ours()
    settings = m.GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=cam["bg"],
        scale_modifier=1.0, viewmatrix=cam["viewmatrix"], projmatrix=cam["projmatrix"], sh_degree=3,
        campos=cam["campos"], prefiltered=False, debug=False)
    ref = refutil.load_ref_rasterizer() if refutil.ref_available() else None
    with torch.no_grad():
        for i, f in enumerate(frames):
            x, q, s = hugs_b200.deform_gaussians(h["A"][f], h["xyz"], h["lbs_weights"], h["rotmat"], h["scales"],
                                                 smpl_scale=1.0, transl=h["transl"][f])
            assert torch.equal(bx[i], x) and torch.equal(bq[i], q) and torch.equal(bs[i], s)
            kw = dict(means3D=bx[i], means2D=torch.zeros_like(bx[i]), opacities=h["opacity"], shs=h["shs"],
                      scales=bs[i], rotations=bq[i])
            color, radii = m.GaussianRasterizer(settings)(**kw)
            assert color.shape == (3, H, W) and bool(torch.isfinite(color).all())
            assert int((radii > 0).sum()) > 0.9 * synth.N_HUMAN          # the human is in view
            if ref is not None:
                rs = ref.GaussianRasterizationSettings(**settings._asdict())
                rc, rr = ref.GaussianRasterizer(rs)(**kw)
                assert torch.equal(radii, rr)
                assert float((color - rc).abs().max()) <= 1e-4
        # frames differ (the pose changes the image)
        c0, _ = m.GaussianRasterizer(settings)(means3D=bx[0], means2D=torch.zeros_like(bx[0]), opacities=h["opacity"],
                                               shs=h["shs"], scales=bs[0], rotations=bq[0])
        c1, _ = m.GaussianRasterizer(settings)(means3D=bx[1], means2D=torch.zeros_like(bx[1]), opacities=h["opacity"],
                                               shs=h["shs"], scales=bs[1], rotations=bq[1])
        assert float((c0 - c1).abs().max()) > 1e-3
