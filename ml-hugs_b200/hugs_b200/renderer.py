"""Renderer glue and post-ops around the rasterizer (the step right after the path).

Mirrors hugs/renderer/gs_renderer.py with the same names, arguments and result keys:
  render              :103-161   settings from the camera dict, rasterize, clamp(0,1), radii > 0
  render_human_scene  :19-100    human / scene / human_scene modes, optional separate human render
In human_scene mode with SH colours the two models' tensors go to the rasterizer as two segments
(segments.py) instead of through five torch.cat calls.  The trainer's densification bookkeeping is ONE kernel without boolean-mask indexing:
  add_densification_stats   hugs/models/scene.py:460-462 + hugs/trainer/gs_trainer.py:406-411,:429-435
"""
import ctypes as C
import math

import torch

from . import _native as N
from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer, output_clamp


def render(means3D, feats, opacity, scales, rotations, data, scaling_modifier=1.0, bg_color=None,
           active_sh_degree=0):
    """hugs/renderer/gs_renderer.py:103-161."""
    device = means3D.device
    if not means3D.is_cuda:
        raise RuntimeError("hugs_b200 has no CPU path: render inputs must be CUDA tensors")
    if bg_color is None:
        bg_color = torch.zeros(3, dtype=torch.float32, device=device)

    screenspace_points = torch.zeros_like(means3D, dtype=means3D.dtype, requires_grad=True, device=device) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass

    shs, rgb = (None, feats) if feats.dim() == 2 else (feats, None)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(data['image_height']),
        image_width=int(data['image_width']),
        tanfovx=math.tan(data['fovx'] * 0.5),
        tanfovy=math.tan(data['fovy'] * 0.5),
        bg=bg_color,
        scale_modifier=scaling_modifier,
        viewmatrix=data['world_view_transform'],
        projmatrix=data['full_proj_transform'],
        sh_degree=active_sh_degree,
        campos=data['camera_center'],
        prefiltered=False,
        debug=False,
    )
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    # torch.clamp(rendered_image, 0.0, 1.0) of gs_renderer.py:153 happens inside the blend kernels
    with output_clamp():
        rendered_image, radii = rasterizer(
            means3D=means3D,
            means2D=screenspace_points,
            shs=shs,
            opacities=opacity,
            scales=scales,
            rotations=rotations,
            colors_precomp=rgb,
        )
    return {
        "render": rendered_image,
        "viewspace_points": screenspace_points,
        "visibility_filter": radii > 0,
        "radii": radii,
    }


def _render_two_segments(human_gs_out, scene_gs_out, data, scaling_modifier, bg_color, active_sh_degree,
                         human_bg_color=None, human_separate=False):
    """`render` (gs_renderer.py:103-161) of human || scene without building the concatenation; with
    `human_separate` also the human-only image of gs_renderer.py:68-82, from the same sorted lists
    (layer.py) instead of a second rasterizer pass."""
    from .segments import rasterize_two_segments
    from .layer import rasterize_two_segments_with_layer
    device = human_gs_out['xyz'].device
    if bg_color is None:
        bg_color = torch.zeros(3, dtype=torch.float32, device=device)
    n = human_gs_out['xyz'].shape[0] + scene_gs_out['xyz'].shape[0]
    screenspace_points = torch.zeros(n, 3, dtype=human_gs_out['xyz'].dtype, device=device, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    settings = GaussianRasterizationSettings(
        image_height=int(data['image_height']), image_width=int(data['image_width']),
        tanfovx=math.tan(data['fovx'] * 0.5), tanfovy=math.tan(data['fovy'] * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=data['world_view_transform'],
        projmatrix=data['full_proj_transform'], sh_degree=active_sh_degree, campos=data['camera_center'],
        prefiltered=False, debug=False)
    seg = lambda o: (o['xyz'], o['shs'], o['opacity'], o['scales'], o['rotq'])  # noqa: E731
    human_image = None
    with output_clamp():   # both images come back clamped to [0, 1] (gs_renderer.py:153)
        if human_separate:
            rendered_image, human_image, radii = rasterize_two_segments_with_layer(
                seg(human_gs_out), seg(scene_gs_out), screenspace_points, settings,
                human_bg_color if human_bg_color is not None else bg_color)
        else:
            rendered_image, radii = rasterize_two_segments(seg(human_gs_out), seg(scene_gs_out), screenspace_points,
                                                           settings)
    pkg = {
        "render": rendered_image,
        "viewspace_points": screenspace_points,
        "visibility_filter": radii > 0,
        "radii": radii,
    }
    if human_image is not None:
        _add_human_layer(pkg, human_image, radii, human_gs_out['xyz'].shape[0])
    return pkg


def _add_human_layer(pkg, human_image, radii, n_human):
    # the keys gs_renderer.py:80-82 copies from its second render(); radii are per Gaussian and do not
    # depend on which other Gaussians are rendered with it
    pkg['human_img'] = human_image   # clamped by the kernel
    pkg['human_radii'] = radii[:n_human]
    pkg['human_visibility_filter'] = pkg['human_radii'] > 0


def _render_with_human_layer(args, n_human, data, scaling_modifier, bg_color, human_bg_color, active_sh_degree):
    """`render` of the concatenated tensors plus the human-only image (precomputed-colour / mixed inputs)."""
    from .layer import rasterize_with_layer
    means3D = args['means3D']
    device = means3D.device
    if bg_color is None:
        bg_color = torch.zeros(3, dtype=torch.float32, device=device)
    screenspace_points = torch.zeros_like(means3D, dtype=means3D.dtype, requires_grad=True, device=device) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    feats = args['feats']
    shs, rgb = (None, feats) if feats.dim() == 2 else (feats, None)
    settings = GaussianRasterizationSettings(
        image_height=int(data['image_height']), image_width=int(data['image_width']),
        tanfovx=math.tan(data['fovx'] * 0.5), tanfovy=math.tan(data['fovy'] * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=data['world_view_transform'],
        projmatrix=data['full_proj_transform'], sh_degree=active_sh_degree, campos=data['camera_center'],
        prefiltered=False, debug=False)
    with output_clamp():
        rendered_image, human_image, radii = rasterize_with_layer(
            means3D, screenspace_points, args['opacity'], args['scales'], args['rotations'], settings, n_human,
            human_bg_color if human_bg_color is not None else bg_color, shs=shs, colors_precomp=rgb)
    pkg = {
        "render": rendered_image,
        "viewspace_points": screenspace_points,
        "visibility_filter": radii > 0,
        "radii": radii,
    }
    _add_human_layer(pkg, human_image, radii, n_human)
    return pkg


_KEYS = (("shs", "feats"), ("xyz", "means3D"), ("opacity", "opacity"), ("scales", "scales"), ("rotq", "rotations"))


def _pick(gs_out):
    return {arg: gs_out[key] for key, arg in _KEYS}


def render_human_scene(data, human_gs_out, scene_gs_out, bg_color, human_bg_color=None, scaling_modifier=1.0,
                       render_mode='human_scene', render_human_separate=False):
    """hugs/renderer/gs_renderer.py:19-100."""
    if render_mode == 'human_scene':
        active_sh_degree = human_gs_out['active_sh_degree']
        if human_gs_out['shs'].dim() == 3 and scene_gs_out['shs'].dim() == 3:
            args = None     # SH colours: the two segments go to the rasterizer as they are, no torch.cat
        else:
            args = {arg: torch.cat([human_gs_out[key], scene_gs_out[key]], dim=0) for key, arg in _KEYS}
    elif render_mode == 'human':
        args, active_sh_degree = _pick(human_gs_out), human_gs_out['active_sh_degree']
    elif render_mode == 'scene':
        args, active_sh_degree = _pick(scene_gs_out), scene_gs_out['active_sh_degree']
    else:
        raise ValueError(f'Unknown render mode: {render_mode}')

    # gs_renderer.py:68-82 renders the human a second time for `human_img`; here it is a second blend over
    # the joint pass's lists (ONE rasterizer forward and backward per step)
    separate = bool(render_human_separate) and render_mode == 'human_scene'
    if args is None:
        render_pkg = _render_two_segments(human_gs_out, scene_gs_out, data, scaling_modifier, bg_color,
                                          active_sh_degree, human_bg_color, separate)
    elif separate:
        render_pkg = _render_with_human_layer(args, human_gs_out['xyz'].shape[0], data, scaling_modifier, bg_color,
                                              human_bg_color, active_sh_degree)
    else:
        render_pkg = render(data=data, scaling_modifier=scaling_modifier, bg_color=bg_color,
                            active_sh_degree=active_sh_degree, **args)

    if render_mode == 'human':
        render_pkg['human_visibility_filter'] = render_pkg['visibility_filter']
        render_pkg['human_radii'] = render_pkg['radii']
    elif render_mode == 'human_scene':
        human_n_gs = human_gs_out['xyz'].shape[0]
        scene_n_gs = scene_gs_out['xyz'].shape[0]
        render_pkg['scene_visibility_filter'] = render_pkg['visibility_filter'][human_n_gs:]
        render_pkg['scene_radii'] = render_pkg['radii'][human_n_gs:]
        if 'human_visibility_filter' not in render_pkg:
            render_pkg['human_visibility_filter'] = render_pkg['visibility_filter'][:-scene_n_gs]
            render_pkg['human_radii'] = render_pkg['radii'][:-scene_n_gs]
    elif render_mode == 'scene':
        render_pkg['scene_visibility_filter'] = render_pkg['visibility_filter']
        render_pkg['scene_radii'] = render_pkg['radii']
    return render_pkg


def _f32_stat(t, n, name):
    if t is None:
        return None
    if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == n):
        raise RuntimeError("%s must be a contiguous CUDA float32 tensor with %d elements" % (name, n))
    return t


@torch.no_grad()
def add_densification_stats(radii, viewspace_point_grad, xyz_gradient_accum=None, denom=None, max_radii2D=None):
    """In place, for every i with radii[i] > 0 (i < len(radii)):
        max_radii2D[i] = max(max_radii2D[i], radii[i])                       gs_trainer.py:406-411
        xyz_gradient_accum[i] += ||viewspace_point_grad[:len(radii)][i, :2]||   scene.py:460-461
        denom[i] += 1                                                        scene.py:462
    One kernel, no nonzero()/host sync.  `radii` may be a slice of the renderer's radii (scene_radii,
    human_radii); the gradient rows used are the FIRST len(radii), as in the reference."""
    n = int(radii.shape[0])
    if not (radii.is_cuda and radii.dtype == torch.int32):
        raise RuntimeError("radii must be a CUDA int32 tensor")
    radii = radii.contiguous()
    g = viewspace_point_grad
    if xyz_gradient_accum is not None:
        if g is None or g.dim() != 2 or g.shape[0] < n or g.shape[1] < 2:
            raise RuntimeError("viewspace_point_grad must be [>=%d, >=2]" % n)
        g = g.detach().float()
        if g.stride(1) != 1:
            g = g.contiguous()
    stride = int(g.stride(0)) if g is not None else 0
    acc = _f32_stat(xyz_gradient_accum, n, "xyz_gradient_accum")
    den = _f32_stat(denom, n, "denom")
    mx = _f32_stat(max_radii2D, n, "max_radii2D")
    device = radii.device
    with torch.cuda.device(device):
        rc = N.densification_stats(n, N.ptr(radii), N.ptr(g), stride, N.ptr(acc), N.ptr(den), N.ptr(mx),
                                   C.c_void_p(torch.cuda.current_stream(device).cuda_stream))
    N.check(rc, "densification_stats")
