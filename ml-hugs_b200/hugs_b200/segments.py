"""Rasterizing two input segments without concatenating them.

hugs/renderer/gs_renderer.py:33-37 joins the human and the scene Gaussians with five torch.cat calls per
render -- 236 B per Gaussian read and written again, plus the split of every gradient in backward.
``rasterize_two_segments`` hands both sets of tensors to the library as they are
(hugs_b200_rasterize_forward_seg2 / _backward_seg2): Gaussians [0, P0) are segment 0 (the human),
[P0, P0+P1) segment 1 (the scene).  Image, radii and gradients are those of GaussianRasterizer on the
concatenation, bit for bit; each input's gradient comes back as a view of the joint gradient array.
SH colours with scales / rotations only (what that renderer passes).
"""
import contextlib
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


class SceneBackwardBatch:
    """The scene segment's per-Gaussian backward of a whole view-batched step, run ONCE
    (hugs_b200_scene_backward_views).  While a batch is active (``defer_scene_backward``), the backward pass of
    every two-segment render whose segment-1 inputs all have gradient sinks runs the blend backward and the
    human segment's per-Gaussian backward only, and leaves its record here; ``finish()`` -- on a stream that is
    ordered after every view's backward pass -- then reads each scene Gaussian's parameters and SH row once,
    walks the views and adds the summed gradients to the sinks once (~120 B per Gaussian and view instead of
    744).  A view's ``means2D.grad`` is complete when its own backward pass returns."""

    def __init__(self):
        self.views, self.common = [], None

    def add(self, view, common):
        if self.common is not None and self.common["key"] != common["key"]:
            self.finish()   # another pair of models: what has been collected goes first
        self.common = common
        self.views.append(view)

    def finish(self):
        if not self.views:
            return
        c, views = self.common, self.views
        self.views, self.common = [], None
        arr = (N.SceneView * len(views))()
        for rec, v in zip(arr, views):
            rec.viewmatrix, rec.projmatrix, rec.cam_pos = v["view"].data_ptr(), v["proj"].data_ptr(), v["campos"].data_ptr()
            rec.tan_fovx, rec.tan_fovy = v["tanfovx"], v["tanfovy"]
            rec.radii, rec.geom_buffer = v["radii"].data_ptr(), v["geom"].data_ptr()
        b = c["b"]
        device = b[0].device
        with torch.cuda.device(device):
            rc = N.scene_backward_views(len(views), arr, c["P0"], c["P1"], c["D"], c["M"], c["W"], c["H"],
                                        N.ptr(b[0]), N.ptr(b[1]), N.ptr(b[3]), c["scale_modifier"], N.ptr(b[4]),
                                        C.byref(c["sinks"]), _stream_ptr(device))
        N.check(rc, "scene_backward_views")


_batch = None   # the active SceneBackwardBatch (set on the thread that runs the forward passes)


def batch_for(sunk, P1):
    """The active batch if this render's scene segment can be deferred (all five of its inputs sunk), else None.
    Called in the forward pass, on the caller's thread: the backward pass runs on autograd's."""
    return _batch if (_batch is not None and all(sunk[5:]) and P1 > 0) else None


def defer_record(batch, s, view, proj, campos, radii, geom, b, P0, P1, M, H, W, sinks):
    batch.add(
        dict(view=view, proj=proj, campos=campos, tanfovx=float(s.tanfovx), tanfovy=float(s.tanfovy), radii=radii,
             geom=geom),
        dict(key=(tuple(x.data_ptr() for x in b), P0, P1, M, H, W, int(s.sh_degree)), b=b, P0=P0, P1=P1,
             D=int(s.sh_degree), M=M, W=W, H=H, scale_modifier=float(s.scale_modifier), sinks=sinks))


@contextlib.contextmanager
def defer_scene_backward():
    """``with defer_scene_backward() as batch:`` around the views of a step; ``batch.finish()`` once every
    view's backward pass has been enqueued and the current stream is ordered after all of them (the context
    exit calls it too)."""
    global _batch
    prev, _batch = _batch, SceneBackwardBatch()
    try:
        yield _batch
    finally:
        batch, _batch = _batch, prev
        batch.finish()


class ScenePrepass:
    """The scene segment's preprocess of a whole view-batched step, run ONCE (hugs_b200_preprocess_scene_views):
    every scene Gaussian's parameters and SH row are read once and the cameras walked, instead of one full
    pass per view.  ``scene_prepass(settings_list, segment1, P0)`` fills, for every view, the scene rows of a
    geometry workspace and of ``radii``; the two-segment renders made inside the context with one of those
    cameras (same viewmatrix / projmatrix / campos tensors, same scene tensors) pick their workspace up here,
    preprocess the human segment only and carry on.  Bit-identical lists and ranges; the image to 1e-6 (the SH
    evaluation is the same source compiled into another kernel)."""

    def __init__(self, settings_list, segment1, P0):
        b = [segment1[k] for k in _NAMES] if isinstance(segment1, dict) else list(segment1)
        device = b[0].device
        b = [_f32c(x, device) for x in b]
        s0 = settings_list[0]
        P1, M = int(b[0].shape[0]), int(b[1].shape[1])
        H, W = int(s0.image_height), int(s0.image_width)
        lay = N.GeomLayout()
        N.geom_layout_of(P0 + P1, C.byref(lay))
        self.entries, self.scene_ptrs = {}, tuple(x.data_ptr() for x in b)
        arr = (N.SceneView * len(settings_list))()
        keep = []
        for rec, s in zip(arr, settings_list):
            if (int(s.image_height), int(s.image_width), int(s.sh_degree), float(s.scale_modifier), bool(s.prefiltered)) \
                    != (H, W, int(s0.sh_degree), float(s0.scale_modifier), bool(s0.prefiltered)):
                raise ValueError("scene_prepass: the views must share image size, SH degree, scale modifier")
            view, proj, campos = (_f32c(x, device) for x in (s.viewmatrix, s.projmatrix, s.campos))
            geom = torch.empty(int(lay.total), dtype=torch.uint8, device=device)
            radii = torch.empty(P0 + P1, dtype=torch.int32, device=device)
            rec.viewmatrix, rec.projmatrix, rec.cam_pos = view.data_ptr(), proj.data_ptr(), campos.data_ptr()
            rec.tan_fovx, rec.tan_fovy = float(s.tanfovx), float(s.tanfovy)
            rec.radii, rec.geom_buffer = radii.data_ptr(), geom.data_ptr()
            keep.append((view, proj, campos))
            self.entries[(view.data_ptr(), proj.data_ptr(), campos.data_ptr(), P0, P1, H, W)] = (geom, radii)
        with torch.cuda.device(device):
            rc = N.preprocess_scene_views(len(settings_list), arr, P0, P1, int(s0.sh_degree), M, W, H,
                                          N.ptr(b[0]), N.ptr(b[1]), N.ptr(b[2]), N.ptr(b[3]),
                                          float(s0.scale_modifier), N.ptr(b[4]), int(bool(s0.prefiltered)),
                                          _stream_ptr(device))
        N.check(rc, "preprocess_scene_views")

    def take(self, view, proj, campos, P0, P1, H, W, b):
        """(geometry workspace, radii) prepared for this camera and these scene tensors, or None (used once)."""
        if tuple(x.data_ptr() for x in b) != self.scene_ptrs:
            return None
        return self.entries.pop((view.data_ptr(), proj.data_ptr(), campos.data_ptr(), P0, P1, H, W), None)


_prepass = None   # the active ScenePrepass


def take_prepared(view, proj, campos, P0, P1, H, W, b):
    """(geometry workspace, radii) the active scene_prepass holds for this camera and these scene tensors, or None."""
    if _prepass is None or P1 <= 0:
        return None
    return _prepass.take(view, proj, campos, P0, P1, H, W, b)


@contextlib.contextmanager
def scene_prepass(settings_list, segment1, P0):
    """``with scene_prepass([settings of every view of the step], scene_segment, n_human):`` around the step's
    two-segment renders; must be entered on a stream the views' streams are ordered after."""
    global _prepass
    prev, _prepass = _prepass, ScenePrepass(settings_list, segment1, P0)
    try:
        yield _prepass
    finally:
        _prepass = prev


class _RasterizeTwoSegments(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means2D, raster_settings, *t):
        a, b = t[:5], t[5:]
        if not all(x.is_cuda for x in t):
            raise RuntimeError("hugs_b200 has no CPU path: segment tensors must be CUDA tensors")
        ctx.sinks, ctx.sunk = sinks_of(a, b)
        ctx.clamp = clamp_active()
        ctx.batch = batch_for(ctx.sunk, int(b[0].shape[0]))
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
        ready = take_prepared(view, proj, campos, P0, P1, H, W, b)
        radii = ready[1] if ready is not None else torch.empty(P, dtype=torch.int32, device=device)
        geom = _Workspace(device, preset=ready[0] if ready is not None else None)
        binning, img = _Workspace(device), _Workspace(device)
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
                    color.data_ptr(), radii.data_ptr(),
                    flags_of(s.debug, ctx.clamp) | (N.FLAG_SEG1_PREPROCESSED if ready is not None else 0),
                    _stream_ptr(device), C.byref(num_rendered))
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
                    g["rotations"].data_ptr(),
                    flags_of(s.debug, ctx.clamp) | (N.FLAG_DEFER_SEG1 if ctx.batch is not None else 0),
                    _stream_ptr(device),
                    0, None, None, 0, None, C.byref(ctx.sinks) if ctx.sinks is not None else None)
            N.check(rc, "rasterize_two_segments backward")
            if ctx.batch is not None:
                defer_record(ctx.batch, s, view, proj, campos, radii, geom, b, P0, P1, M, H, W, ctx.sinks)
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
