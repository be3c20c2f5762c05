"""The separate human image from the joint pass's sorted lists.

hugs/renderer/gs_renderer.py:68-82 calls ``render`` a second time on the human Gaussians alone whenever
the trainer wants ``human_img`` (hugs/trainer/gs_trainer.py:255-271): a second complete rasterizer
forward and, later, backward.  Here the joint pass is followed by ``hugs_b200_layer_forward``: the joint
tile lists are filtered to the first ``n_layer`` Gaussians and blended against the layer's own background
(csrc/layer.cu); the backward pass consumes both image gradients in one call
(``hugs_b200_rasterize_backward[_seg2]_layer``).  One preprocess, one binning, one per-Gaussian backward.

``rasterize_with_layer``            one set of tensors (SH or precomputed colours, scales + rotations)
``rasterize_two_segments_with_layer``  the human || scene pair of segments.py
Both return (color, layer_color, radii).  Every input gradient is the sum over the two images, as autograd
gives for the reference's two calls; ``means2D.grad`` is the main image's alone (the reference's second
pass has a screen-space tensor of its own that nothing reads).
"""
import ctypes as C

import torch

from . import _native as N
from .rasterizer import _Workspace, _f32c, _stream_ptr, clamp_active, flags_of


def _layer_forward(P, n_layer, bg_layer, W, H, colors, geom, binning, img, flags, device):
    nbytes = int(N.layer_bytes(n_layer, int(binning.numel()), W, H))
    layer = torch.empty(nbytes, dtype=torch.uint8, device=device)
    out = torch.empty(3, H, W, dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        rc = N.layer_forward(P, n_layer, N.ptr(bg_layer), W, H, N.ptr(colors), N.ptr(geom), N.ptr(binning),
                             int(binning.numel()), N.ptr(img), layer.data_ptr(), nbytes, out.data_ptr(),
                             int(flags), _stream_ptr(device))
    N.check(rc, "layer_forward")
    return out, layer


def _run_forward(call, device, geom_preset=None):
    geom, binning, img = _Workspace(device, preset=geom_preset), _Workspace(device), _Workspace(device)
    num_rendered = C.c_int(0)
    with torch.cuda.device(device):
        rc = call(geom, binning, img, num_rendered)
    for ws in (geom, binning, img):
        ws.release()
        if ws.error is not None:
            raise RuntimeError("workspace allocation failed: %r" % (ws.error,))
    N.check(rc, "rasterize_with_layer")
    return int(num_rendered.value), geom.tensor, binning.tensor, img.tensor


def _grad_buffers(P, M, device):
    opts = dict(dtype=torch.float32, device=device)
    return {k: torch.empty(shape, **opts) for k, shape in (
        ("means2D", (P, 3)), ("conic", (P, 2, 2)), ("opacity", (P, 1)), ("colors", (P, 3)), ("means3D", (P, 3)),
        ("cov3D", (P, 6)), ("sh", (P, M, 3)), ("scales", (P, 3)), ("rotations", (P, 4)))}


def _grad_ptrs(g):
    return [g[k].data_ptr() for k in ("means2D", "conic", "opacity", "colors", "means3D", "cov3D", "sh", "scales",
                                      "rotations")]


def _check_layer(n_layer, P):
    n_layer = int(n_layer)
    if n_layer < 0 or n_layer > P:
        raise ValueError("n_layer must lie in [0, %d]" % P)
    return n_layer


class _RasterizeWithLayer(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, raster_settings, n_layer,
                bg_layer):
        if not means3D.is_cuda:
            raise RuntimeError("hugs_b200 has no CPU path: means3D must be a CUDA tensor")
        ctx.set_materialize_grads(False)   # an image nobody differentiates costs no blend backward
        ctx.clamp = clamp_active()
        s, device = raster_settings, means3D.device
        P = int(means3D.shape[0])
        if P == 0:
            raise RuntimeError("rasterize_with_layer needs at least one Gaussian")
        n_layer = _check_layer(n_layer, P)
        H, W = int(s.image_height), int(s.image_width)
        means3D, sh, colors_precomp, opacities, scales, rotations = (
            _f32c(x, device) for x in (means3D, sh, colors_precomp, opacities, scales, rotations))
        bg, view, proj, campos, bg_layer = (_f32c(x, device) for x in (s.bg, s.viewmatrix, s.projmatrix, s.campos,
                                                                       bg_layer))
        M = int(sh.size(1)) if (sh is not None and sh.numel() != 0) else 0
        radii = torch.empty(P, dtype=torch.int32, device=device)
        color = torch.empty(3, H, W, dtype=torch.float32, device=device)
        R, geom, binning, img = _run_forward(lambda g, b, i, nr: N.rasterize_forward(
            g.callback, None, b.callback, None, i.callback, None, P, int(s.sh_degree), M, N.ptr(bg), W, H,
            N.ptr(means3D), N.ptr(sh), N.ptr(colors_precomp), N.ptr(opacities), N.ptr(scales),
            float(s.scale_modifier), N.ptr(rotations), None, N.ptr(view), N.ptr(proj), N.ptr(campos),
            float(s.tanfovx), float(s.tanfovy), int(bool(s.prefiltered)), color.data_ptr(), radii.data_ptr(),
            flags_of(s.debug, ctx.clamp), _stream_ptr(device), C.byref(nr)), device)
        layer_color, layer = _layer_forward(P, n_layer, bg_layer, W, H, colors_precomp, geom, binning, img,
                                            flags_of(s.debug, ctx.clamp), device)
        ctx.raster_settings, ctx.num_rendered, ctx.sizes = s, R, (P, M, H, W, n_layer)
        ctx.save_for_backward(radii, geom, binning, img, layer, bg, bg_layer, view, proj, campos, means3D, sh,
                              colors_precomp, scales, rotations)
        ctx.mark_non_differentiable(radii)
        return color, layer_color, radii

    @staticmethod
    def backward(ctx, grad_color, grad_layer, _):
        (radii, geom, binning, img, layer, bg, bg_layer, view, proj, campos, means3D, sh, colors_precomp, scales,
         rotations) = ctx.saved_tensors
        P, M, H, W, n_layer = ctx.sizes
        s, device = ctx.raster_settings, radii.device
        g = _grad_buffers(P, M, device)
        dL = _f32c(grad_color, device) if grad_color is not None else torch.zeros(3, H, W, device=device)
        dLl = _f32c(grad_layer, device) if grad_layer is not None else None
        with torch.cuda.device(device):
            rc = N.rasterize_backward_layer(
                P, int(s.sh_degree), M, int(ctx.num_rendered), N.ptr(bg), W, H, N.ptr(means3D), N.ptr(sh),
                N.ptr(colors_precomp), N.ptr(scales), float(s.scale_modifier), N.ptr(rotations), None, N.ptr(view),
                N.ptr(proj), N.ptr(campos), float(s.tanfovx), float(s.tanfovy), radii.data_ptr(), N.ptr(geom),
                N.ptr(binning), int(binning.numel()), N.ptr(img), dL.data_ptr(), *_grad_ptrs(g),
                flags_of(s.debug, ctx.clamp), _stream_ptr(device), 0, n_layer, N.ptr(bg_layer), layer.data_ptr(), int(layer.numel()), N.ptr(dLl))
        N.check(rc, "rasterize_with_layer backward")
        has = lambda t: t is not None and t.numel() != 0  # noqa: E731
        return (g["means3D"], g["means2D"], g["sh"] if has(sh) else None, g["colors"] if has(colors_precomp) else None,
                g["opacity"], g["scales"], g["rotations"], None, None, None)


def rasterize_with_layer(means3D, means2D, opacities, scales, rotations, raster_settings, n_layer, bg_layer,
                         shs=None, colors_precomp=None):
    """GaussianRasterizer(raster_settings)(...) plus the image of Gaussians [0, n_layer) alone against
    ``bg_layer``.  Returns (color [3,H,W], layer_color [3,H,W], radii int32 [P])."""
    if (shs is None) == (colors_precomp is None):
        raise Exception('Please provide excatly one of either SHs or precomputed colors!')
    empty = torch.Tensor([])
    return _RasterizeWithLayer.apply(means3D, means2D, empty if shs is None else shs,
                                     empty if colors_precomp is None else colors_precomp, opacities, scales,
                                     rotations, raster_settings, n_layer, bg_layer)


class _RasterizeTwoSegmentsWithLayer(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means2D, raster_settings, n_layer, bg_layer, *t):
        a, b = t[:5], t[5:]
        if not all(x.is_cuda for x in t):
            raise RuntimeError("hugs_b200 has no CPU path: segment tensors must be CUDA tensors")
        ctx.set_materialize_grads(False)
        ctx.clamp = clamp_active()
        from .segments import batch_for, sinks_of
        ctx.sinks, ctx.sunk = sinks_of(a, b)
        ctx.batch = batch_for(ctx.sunk, int(b[0].shape[0]))
        device = a[0].device
        a = [_f32c(x, device) for x in a]
        b = [_f32c(x, device) for x in b]
        P0, P1 = int(a[0].shape[0]), int(b[0].shape[0])
        P = P0 + P1
        if P == 0:
            raise RuntimeError("rasterize_two_segments_with_layer needs at least one Gaussian")
        n_layer = _check_layer(n_layer, P)
        if a[1].dim() != 3 or b[1].dim() != 3 or a[1].shape[1:] != b[1].shape[1:]:
            raise RuntimeError("shs must be [P,M,3] with the same M in both segments")
        s = raster_settings
        H, W, M = int(s.image_height), int(s.image_width), int(a[1].shape[1])
        bg, view, proj, campos, bg_layer = (_f32c(x, device) for x in (s.bg, s.viewmatrix, s.projmatrix, s.campos,
                                                                       bg_layer))
        from .segments import take_prepared
        ready = take_prepared(view, proj, campos, P0, P1, H, W, b)   # segments.scene_prepass
        radii = ready[1] if ready is not None else torch.empty(P, dtype=torch.int32, device=device)
        color = torch.empty(3, H, W, dtype=torch.float32, device=device)
        fwd_flags = flags_of(s.debug, ctx.clamp) | (N.FLAG_SEG1_PREPROCESSED if ready is not None else 0)
        R, geom, binning, img = _run_forward(lambda g, bn, i, nr: N.rasterize_forward_seg2(
            g.callback, None, bn.callback, None, i.callback, None, int(s.sh_degree), M, N.ptr(bg), W, H,
            P0, *[N.ptr(x) for x in a], P1, *[N.ptr(x) for x in b], float(s.scale_modifier), N.ptr(view),
            N.ptr(proj), N.ptr(campos), float(s.tanfovx), float(s.tanfovy), int(bool(s.prefiltered)),
            color.data_ptr(), radii.data_ptr(), fwd_flags, _stream_ptr(device), C.byref(nr)), device,
            geom_preset=ready[0] if ready is not None else None)
        layer_color, layer = _layer_forward(P, n_layer, bg_layer, W, H, None, geom, binning, img,
                                            flags_of(s.debug, ctx.clamp), device)
        ctx.raster_settings, ctx.num_rendered, ctx.sizes = s, R, (P0, P1, M, H, W, n_layer)
        ctx.save_for_backward(radii, geom, binning, img, layer, bg, bg_layer, view, proj, campos, *a, *b)
        ctx.mark_non_differentiable(radii)
        return color, layer_color, radii

    @staticmethod
    def backward(ctx, grad_color, grad_layer, _):
        radii, geom, binning, img, layer, bg, bg_layer, view, proj, campos, *t = ctx.saved_tensors
        a, b = t[:5], t[5:]
        P0, P1, M, H, W, n_layer = ctx.sizes
        s, P, device = ctx.raster_settings, P0 + P1, radii.device
        g = _grad_buffers(P, M, device)
        dL = _f32c(grad_color, device) if grad_color is not None else torch.zeros(3, H, W, device=device)
        dLl = _f32c(grad_layer, device) if grad_layer is not None else None
        with torch.cuda.device(device):
            rc = N.rasterize_backward_seg2_ex(
                int(s.sh_degree), M, int(ctx.num_rendered), N.ptr(bg), W, H,
                P0, N.ptr(a[0]), N.ptr(a[1]), N.ptr(a[3]), N.ptr(a[4]),
                P1, N.ptr(b[0]), N.ptr(b[1]), N.ptr(b[3]), N.ptr(b[4]),
                float(s.scale_modifier), N.ptr(view), N.ptr(proj), N.ptr(campos), float(s.tanfovx), float(s.tanfovy),
                radii.data_ptr(), N.ptr(geom), N.ptr(binning), int(binning.numel()), N.ptr(img), dL.data_ptr(),
                *_grad_ptrs(g), flags_of(s.debug, ctx.clamp) | (N.FLAG_DEFER_SEG1 if ctx.batch is not None else 0),
                _stream_ptr(device),
                n_layer, N.ptr(bg_layer), layer.data_ptr(), int(layer.numel()), N.ptr(dLl),
                C.byref(ctx.sinks) if ctx.sinks is not None else None)
        N.check(rc, "rasterize_two_segments_with_layer backward")
        from .segments import defer_record, segment_grads
        if ctx.batch is not None:
            defer_record(ctx.batch, s, view, proj, campos, radii, geom, b, P0, P1, M, H, W, ctx.sinks)
        return (g["means2D"], None, None, None, *segment_grads(g, P0, ctx.sunk))


def rasterize_two_segments_with_layer(segment0, segment1, means2D, raster_settings, bg_layer, n_layer=None):
    """segments as in ``rasterize_two_segments``; the layer is segment0 (the human) unless ``n_layer`` says
    otherwise.  Returns (color, layer_color, radii)."""
    from .segments import _NAMES

    def tensors(seg):
        return [seg[k] for k in _NAMES] if isinstance(seg, dict) else list(seg)
    a, b = tensors(segment0), tensors(segment1)
    if n_layer is None:
        n_layer = int(a[0].shape[0])
    return _RasterizeTwoSegmentsWithLayer.apply(means2D, raster_settings, n_layer, bg_layer, *a, *b)
