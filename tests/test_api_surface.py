"""The Python face keeps the reference's diff_gaussian_rasterization API
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py:157-220)."""
import inspect

import pytest
import torch


def test_settings_namedtuple_fields_in_reference_order():
    import diff_gaussian_rasterization as dgr

    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")


def test_rasterizer_signature_and_exports():
    import diff_gaussian_rasterization as dgr

    sig = inspect.signature(dgr.GaussianRasterizer.forward)
    assert list(sig.parameters) == ["self", "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales",
                                    "rotations", "cov3D_precomp"]
    assert all(sig.parameters[k].default is None for k in ("shs", "colors_precomp", "scales", "rotations",
                                                           "cov3D_precomp"))
    assert hasattr(dgr.GaussianRasterizer, "markVisible")
    assert list(inspect.signature(dgr.rasterize_gaussians).parameters) == [
        "means3D", "means2D", "sh", "colors_precomp", "opacities", "scales", "rotations", "cov3Ds_precomp",
        "raster_settings"]
    for name in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert hasattr(dgr._C, name)
    assert issubclass(dgr._RasterizeGaussians, torch.autograd.Function)


def _settings(dgr):
    return dgr.GaussianRasterizationSettings(
        image_height=32, image_width=32, tanfovx=1.0, tanfovy=1.0, bg=torch.zeros(3), scale_modifier=1.0,
        viewmatrix=torch.eye(4), projmatrix=torch.eye(4), sh_degree=0, campos=torch.zeros(3), prefiltered=False,
        debug=False)


def test_exactly_one_of_validation_messages():
    import diff_gaussian_rasterization as dgr

    r = dgr.GaussianRasterizer(_settings(dgr))
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, torch.ones(4, 1), scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, torch.ones(4, 1), shs=torch.ones(4, 1, 3), colors_precomp=torch.ones(4, 3), scales=torch.ones(4, 3),
          rotations=torch.ones(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.ones(4, 1), shs=torch.ones(4, 1, 3), scales=torch.ones(4, 3))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.ones(4, 1), shs=torch.ones(4, 1, 3), scales=torch.ones(4, 3), rotations=torch.ones(4, 4),
          cov3D_precomp=torch.ones(4, 6))


def test_no_cpu_fallback():
    """CPU tensors are refused loudly: there is no CPU path in the product."""
    import diff_gaussian_rasterization as dgr
    import hugs_b200

    r = dgr.GaussianRasterizer(_settings(dgr))
    m = torch.zeros(4, 3)
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(m, m, torch.ones(4, 1), shs=torch.ones(4, 1, 3), scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(RuntimeError, match="dimensions"):
        r(torch.zeros(4, 2), m, torch.ones(4, 1), shs=torch.ones(4, 1, 3), scales=torch.ones(4, 3),
          rotations=torch.ones(4, 4))
    with pytest.raises(RuntimeError, match="no CPU path"):
        hugs_b200.deform_gaussians(torch.eye(4).repeat(24, 1, 1), torch.zeros(5, 3), torch.ones(5, 24) / 24,
                                   torch.eye(3).repeat(5, 1, 1))


def test_lbs_extra_keeps_reference_signature():
    import hugs_b200

    assert list(inspect.signature(hugs_b200.lbs_extra).parameters) == [
        "A", "v_shaped", "posedirs", "lbs_weights", "pose", "disable_posedirs", "pose2rot"]


def test_renderer_glue_keeps_the_reference_signatures():
    """hugs/renderer/gs_renderer.py:19-28 (render_human_scene) and :103 (render)."""
    from hugs_b200 import renderer

    assert list(inspect.signature(renderer.render_human_scene).parameters) == [
        "data", "human_gs_out", "scene_gs_out", "bg_color", "human_bg_color", "scaling_modifier", "render_mode",
        "render_human_separate"]
    p = inspect.signature(renderer.render_human_scene).parameters
    assert p["render_mode"].default == "human_scene" and p["render_human_separate"].default is False
    assert list(inspect.signature(renderer.render).parameters) == [
        "means3D", "feats", "opacity", "scales", "rotations", "data", "scaling_modifier", "bg_color",
        "active_sh_degree"]
    with pytest.raises(RuntimeError, match="no CPU path"):
        renderer.render(torch.zeros(2, 3), torch.zeros(2, 3), torch.zeros(2, 1), torch.zeros(2, 3),
                        torch.zeros(2, 4), {})
