"""Rasterizing two input segments without concatenating them.

hugs/renderer/gs_renderer.py:33-37 joins the human and the scene Gaussians with five torch.cat calls per
render -- 236 B per Gaussian read and written again, plus the split of every gradient in backward.
``rasterize_two_segments`` hands both sets of tensors to the library as they are
(hugs_b200_rasterize_forward_seg2 / _backward_seg2): Gaussians [0, P0) are segment 0 (the human),
[P0, P0+P1) segment 1 (the scene).  Image, radii and gradients are those of GaussianRasterizer on the
concatenation, bit for bit; each input's gradient comes back as a view of the joint gradient array.
SH colours with scales / rotations only (what that renderer passes).
"""
import ctypes as C

import torch

from . import _native as N
from .rasterizer import _Workspace, _f32c, _stream_ptr, clamp_active, flags_of, grad_sink_of

_NAMES = ("means3D", "shs", "opacities", "scales", "rotations")
_SINK_FIELDS = ("means3D", "sh", "opacity", "scales", "rotations")   # hugs_b200_grad_sinks, in _NAMES order


def sinks_of(raw_a, raw_b):
    """(GradSinks struct or None, [10 bools]) for the raw input tensors of the two segments: an input that
    has a gradient sink attached (hugs_b200.attach_grad_sink) gets its gradient delivered there."""
    st, flags, any_sink = N.GradSinks(), [], False
    for seg, raw in enumerate((raw_a, raw_b)):
        for field, t in zip(_SINK_FIELDS, raw):
            sink = grad_sink_of(t)
            flags.append(sink is not None)
            if sink is not None:
                getattr(st, field)[seg] = sink.buffer.data_ptr()
                any_sink = True
    return (st if any_sink else None), flags


class _RasterizeTwoSegments(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means2D, raster_settings, *t):
        a, b = t[:5], t[5:]
        if not all(x.is_cuda for x in t):
            raise RuntimeError("hugs_b200 has no CPU path: segment tensors must be CUDA tensors")
        ctx.sinks, ctx.sunk = sinks_of(a, b)
        ctx.clamp = clamp_active()
        device = a[0].device
        a = [_f32c(x, device) for x in a]
        b = [_f32c(x, device) for x in b]
        P0, P1 = int(a[0].shape[0]), int(b[0].shape[0])
        for seg, n in ((a, P0), (b, P1)):
            if any(int(x.shape[0]) != n for x in seg) or seg[0].dim() != 2 or seg[0].shape[1] != 3:
                raise RuntimeError("every tensor of a segment needs the segment's row count; means3D is [P,3]")
        if a[1].dim() != 3 or b[1].dim() != 3 or a[1].shape[1:] != b[1].shape[1:]:
            raise RuntimeError("shs must be [P,M,3] with the same M in both segments")
        s = raster_settings
        H, W, M = int(s.image_height), int(s.image_width), int(a[1].shape[1])
        bg, view, proj, campos = (_f32c(x, device) for x in (s.bg, s.viewmatrix, s.projmatrix, s.campos))
        P = P0 + P1
        radii = torch.empty(P, dtype=torch.int32, device=device)
        geom, binning, img = _Workspace(device), _Workspace(device), _Workspace(device)
        if P == 0:
            for ws in (geom, binning, img):
                ws.release()
            color, R = torch.zeros(3, H, W, dtype=torch.float32, device=device), 0
        else:
            color = torch.empty(3, H, W, dtype=torch.float32, device=device)
            num_rendered = C.c_int(0)
            with torch.cuda.device(device):
                rc = N.rasterize_forward_seg2(
                    geom.callback, None, binning.callback, None, img.callback, None,
                    int(s.sh_degree), M, N.ptr(bg), W, H,
                    P0, *[N.ptr(x) for x in a], P1, *[N.ptr(x) for x in b],
                    float(s.scale_modifier), N.ptr(view), N.ptr(proj), N.ptr(campos),
                    float(s.tanfovx), float(s.tanfovy), int(bool(s.prefiltered)),
                    color.data_ptr(), radii.data_ptr(), flags_of(s.debug, ctx.clamp), _stream_ptr(device),
                    C.byref(num_rendered))
            for ws in (geom, binning, img):
                ws.release()
                if ws.error is not None:
                    raise RuntimeError("workspace allocation failed: %r" % (ws.error,))
            N.check(rc, "rasterize_two_segments")
            R = int(num_rendered.value)
        ctx.raster_settings, ctx.num_rendered, ctx.sizes = s, R, (P0, P1, M, H, W)
        ctx.save_for_backward(radii, geom.tensor, binning.tensor, img.tensor, bg, view, proj, campos, *a, *b)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_color, _):
        radii, geom, binning, img, bg, view, proj, campos, *t = ctx.saved_tensors
        a, b = t[:5], t[5:]
        P0, P1, M, H, W = ctx.sizes
        s, P = ctx.raster_settings, P0 + P1
        device = radii.device
        opts = dict(dtype=torch.float32, device=device)
        alloc = torch.empty if P > 0 else torch.zeros
        g = {k: alloc(shape, **opts) for k, shape in (
            ("means2D", (P, 3)), ("conic", (P, 2, 2)), ("opacity", (P, 1)), ("colors", (P, 3)), ("means3D", (P, 3)),
            ("cov3D", (P, 6)), ("sh", (P, M, 3)), ("scales", (P, 3)), ("rotations", (P, 4)))}
        if P > 0:
            dL = _f32c(grad_color, device)
            with torch.cuda.device(device):
                rc = N.rasterize_backward_seg2_ex(
                    int(s.sh_degree), M, int(ctx.num_rendered), N.ptr(bg), W, H,
                    P0, N.ptr(a[0]), N.ptr(a[1]), N.ptr(a[3]), N.ptr(a[4]),
                    P1, N.ptr(b[0]), N.ptr(b[1]), N.ptr(b[3]), N.ptr(b[4]),
                    float(s.scale_modifier), N.ptr(view), N.ptr(proj), N.ptr(campos),
                    float(s.tanfovx), float(s.tanfovy),
                    radii.data_ptr(), N.ptr(geom), N.ptr(binning), int(binning.numel()), N.ptr(img),
                    dL.data_ptr(),
                    g["means2D"].data_ptr(), g["conic"].data_ptr(), g["opacity"].data_ptr(), g["colors"].data_ptr(),
                    g["means3D"].data_ptr(), g["cov3D"].data_ptr(), g["sh"].data_ptr(), g["scales"].data_ptr(),
                    g["rotations"].data_ptr(), flags_of(s.debug, ctx.clamp), _stream_ptr(device),
                    0, None, None, 0, None, C.byref(ctx.sinks) if ctx.sinks is not None else None)
            N.check(rc, "rasterize_two_segments backward")
        return (g["means2D"], None, *segment_grads(g, P0, ctx.sunk))


def segment_grads(g, P0, sunk):
    """The ten per-segment input gradients as views of the joint arrays; None where a sink took delivery."""
    order = ("means3D", "sh", "opacity", "scales", "rotations")
    out = [g[k][:P0] for k in order] + [g[k][P0:] for k in order]
    return [None if s else v for v, s in zip(out, sunk)]


def rasterize_two_segments(segment0, segment1, means2D, raster_settings):
    """segment0 / segment1: dicts (or sequences in this order) of means3D [P,3], shs [P,M,3],
    opacities [P,1], scales [P,3], rotations [P,4].  means2D: the [P0+P1,3] dummy whose .grad receives
    the screen-space gradients.  Returns (color [3,H,W], radii int32 [P0+P1])."""
    def tensors(seg):
        return [seg[k] for k in _NAMES] if isinstance(seg, dict) else list(seg)
    return _RasterizeTwoSegments.apply(means2D, raster_settings, *tensors(segment0), *tensors(segment1))
