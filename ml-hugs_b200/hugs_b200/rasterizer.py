"""Python face of the B200 rasterizer: the reference's ``diff_gaussian_rasterization`` API, kept
verbatim (names, argument order, defaults, return shapes/dtypes, error messages), on top of the
C-ABI library.

Reference being mirrored:
  submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py
    rasterize_gaussians :21-42, _RasterizeGaussians :44-155,
    GaussianRasterizationSettings :157-169, GaussianRasterizer :171-220
  submodules/diff-gaussian-rasterization/rasterize_points.cu
    RasterizeGaussiansCUDA :35-115, RasterizeGaussiansBackwardCUDA :117-196, markVisible :198-217

All device memory (inputs, outputs, the three opaque workspaces) is torch-owned; the native side
receives raw pointers plus torch's current CUDA stream.
"""
import contextlib
import ctypes as C
import threading
from types import SimpleNamespace
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _native as N


def cpu_deep_copy_tuple(input_tuple):
    copied_tensors = [item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple]
    return tuple(copied_tensors)


def _f32c(t, device=None):
    """contiguous float32 (the reference calls .contiguous().data<float>() on everything,
    rasterize_points.cu:94-111); absent optionals (numel()==0) pass through untouched."""
    if t is None or t.numel() == 0:
        return t
    if device is not None and t.device != device:
        t = t.to(device)
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


_opts = threading.local()


@contextlib.contextmanager
def output_clamp(enabled=True):
    """Rasterizer calls made inside this context return the image already clamped to [0, 1]
    (hugs/renderer/gs_renderer.py:153 does ``torch.clamp(rendered_image, 0.0, 1.0)`` right after the call) and
    their backward pass treats the incoming gradient as that of the clamped image: the clamp and its
    backward mask run inside the blend kernels (HUGS_B200_FLAG_CLAMP01) instead of as two elementwise passes."""
    prev = getattr(_opts, "clamp", False)
    _opts.clamp = bool(enabled)
    try:
        yield
    finally:
        _opts.clamp = prev


def clamp_active():
    return getattr(_opts, "clamp", False)


def flags_of(debug, clamp):
    return (N.FLAG_DEBUG if debug else 0) | (N.FLAG_CLAMP01 if clamp else 0)


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class _Workspace:
    """Resize callback target: allocates a uint8 CUDA tensor of the requested size and hands its
    pointer to the native side (the reference's resizeFunctional, rasterize_points.cu:27-33)."""

    def __init__(self, device, preset=None):
        self.device = device
        self.tensor = torch.empty(0, dtype=torch.uint8, device=device)
        self.error = None

        def _resize(_user, nbytes):
            try:
                if preset is not None:   # a buffer that is already (partly) filled: segments.scene_prepass
                    if int(preset.numel()) != int(nbytes):
                        raise RuntimeError("preset workspace of %d bytes, %d wanted" % (preset.numel(), nbytes))
                    self.tensor = preset
                    return self.tensor.data_ptr()
                self.tensor = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
                return self.tensor.data_ptr()
            except Exception as ex:  # surfaced by the caller as RuntimeError
                self.error = ex
                return None

        self.callback = N.RESIZE_FN(_resize)

    def release(self):
        """Drop the ctypes thunk.  It closes over `self`, a reference cycle: without this the workspace
        tensor (hundreds of MB) would live until the cyclic garbage collector happens to run."""
        self.callback = None


def rasterize_gaussians_forward(
    background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
    viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
    prefiltered, debug,
):
    """Same 19 arguments and 6 return values as the reference's ``_C.rasterize_gaussians``
    (rasterize_points.h:18-38): (num_rendered, color[3,H,W], radii[P], geomBuffer, binningBuffer,
    imgBuffer)."""
    if means3D.ndimension() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("hugs_b200 has no CPU path: means3D must be a CUDA tensor")
    device = means3D.device
    P = int(means3D.size(0))
    H, W = int(image_height), int(image_width)

    means3D = _f32c(means3D)
    colors = _f32c(colors, device)
    opacity = _f32c(opacity, device)
    scales = _f32c(scales, device)
    rotations = _f32c(rotations, device)
    cov3D_precomp = _f32c(cov3D_precomp, device)
    sh = _f32c(sh, device)
    background = _f32c(background, device)
    viewmatrix = _f32c(viewmatrix, device)
    projmatrix = _f32c(projmatrix, device)
    campos = _f32c(campos, device)

    radii = torch.empty(P, dtype=torch.int32, device=device)
    geom, binning, img = _Workspace(device), _Workspace(device), _Workspace(device)
    if P == 0:
        # rasterize_points.cu:81: nothing is launched; colour stays at its zero fill.
        out_color = torch.zeros(3, H, W, dtype=torch.float32, device=device)
        for ws in (geom, binning, img):
            ws.release()
        return 0, out_color, radii, geom.tensor, binning.tensor, img.tensor
    out_color = torch.empty(3, H, W, dtype=torch.float32, device=device)

    M = int(sh.size(1)) if (sh is not None and sh.numel() != 0) else 0
    num_rendered = C.c_int(0)
    with torch.cuda.device(device):
        rc = N.rasterize_forward(
            geom.callback, None, binning.callback, None, img.callback, None,
            P, int(degree), M,
            N.ptr(background), W, H,
            N.ptr(means3D), N.ptr(sh), N.ptr(colors), N.ptr(opacity),
            N.ptr(scales), float(scale_modifier), N.ptr(rotations), N.ptr(cov3D_precomp),
            N.ptr(viewmatrix), N.ptr(projmatrix), N.ptr(campos),
            float(tan_fovx), float(tan_fovy), int(bool(prefiltered)),
            out_color.data_ptr(), radii.data_ptr(), flags_of(debug, clamp_active()), _stream_ptr(device),
            C.byref(num_rendered),
        )
    for ws in (geom, binning, img):
        ws.release()
        if ws.error is not None:
            raise RuntimeError("workspace allocation failed: %r" % (ws.error,))
    N.check(rc, "rasterize_gaussians")
    return int(num_rendered.value), out_color, radii, geom.tensor, binning.tensor, img.tensor


def rasterize_gaussians_backward(
    background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
    projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R, binningBuffer,
    imageBuffer, debug, sh_grad_out=None, sh_grad_accumulate=False, clamp=False,
):
    """Same 21 arguments and 8 return values as the reference's ``_C.rasterize_gaussians_backward``
    (rasterize_points.h:40-62): (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D,
    dL_dsh, dL_dscales, dL_drotations).

    ``sh_grad_out`` / ``sh_grad_accumulate`` (not in the reference) serve view-batched training: the SH
    gradient is written (or, with ``sh_grad_accumulate``, added for the visible Gaussians) straight into
    the caller's [P,M,3] buffer, which is also what is returned as dL_dsh."""
    device = means3D.device
    P = int(means3D.size(0))
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    M = int(sh.size(1)) if (sh is not None and sh.numel() != 0) else 0

    means3D = _f32c(means3D)
    colors = _f32c(colors, device)
    scales = _f32c(scales, device)
    rotations = _f32c(rotations, device)
    cov3D_precomp = _f32c(cov3D_precomp, device)
    sh = _f32c(sh, device)
    background = _f32c(background, device)
    viewmatrix = _f32c(viewmatrix, device)
    projmatrix = _f32c(projmatrix, device)
    campos = _f32c(campos, device)
    dL_dout_color = _f32c(dL_dout_color, device)
    radii = radii.contiguous()

    opts = dict(dtype=torch.float32, device=device)
    # No zero fill: the native backward writes every element (rasterize_points.cu:151-159 memsets
    # ~300 B per Gaussian here).
    alloc = torch.empty if P > 0 else torch.zeros
    dL_dmeans3D = alloc((P, 3), **opts)
    dL_dmeans2D = alloc((P, 3), **opts)
    dL_dcolors = alloc((P, 3), **opts)
    dL_dconic = alloc((P, 2, 2), **opts)
    dL_dopacity = alloc((P, 1), **opts)
    dL_dcov3D = alloc((P, 6), **opts)
    if sh_grad_out is not None:
        if (sh_grad_out.dtype != torch.float32 or sh_grad_out.device != device or not sh_grad_out.is_contiguous()
                or tuple(sh_grad_out.shape) != (P, M, 3)):
            raise ValueError("sh_grad_out must be a contiguous float32 [P,M,3] tensor on the inputs' device")
        dL_dsh = sh_grad_out
    else:
        if sh_grad_accumulate:
            raise ValueError("sh_grad_accumulate needs sh_grad_out")
        dL_dsh = alloc((P, M, 3), **opts)
    dL_dscales = alloc((P, 3), **opts)
    dL_drotations = alloc((P, 4), **opts)

    if P != 0:
        with torch.cuda.device(device):
            rc = N.rasterize_backward_acc(
                P, int(degree), M, int(R),
                N.ptr(background), W, H,
                N.ptr(means3D), N.ptr(sh), N.ptr(colors),
                N.ptr(scales), float(scale_modifier), N.ptr(rotations), N.ptr(cov3D_precomp),
                N.ptr(viewmatrix), N.ptr(projmatrix), N.ptr(campos),
                float(tan_fovx), float(tan_fovy),
                radii.data_ptr(), N.ptr(geomBuffer), N.ptr(binningBuffer), int(binningBuffer.numel()),
                N.ptr(imageBuffer),
                dL_dout_color.data_ptr(),
                dL_dmeans2D.data_ptr(), dL_dconic.data_ptr(), dL_dopacity.data_ptr(), dL_dcolors.data_ptr(),
                dL_dmeans3D.data_ptr(), dL_dcov3D.data_ptr(), N.ptr(dL_dsh), dL_dscales.data_ptr(),
                dL_drotations.data_ptr(),
                flags_of(debug, clamp), _stream_ptr(device), int(bool(sh_grad_accumulate)),
            )
        N.check(rc, "rasterize_gaussians_backward")
    return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations


def mark_visible(means3D, viewmatrix, projmatrix):
    """``_C.mark_visible`` (rasterize_points.cu:198-217): bool[P], True where view-space z > 0.2."""
    device = means3D.device
    P = int(means3D.size(0))
    present = torch.zeros(P, dtype=torch.bool, device=device)
    if P != 0:
        means3D = _f32c(means3D)
        viewmatrix = _f32c(viewmatrix, device)
        projmatrix = _f32c(projmatrix, device)
        with torch.cuda.device(device):
            rc = N.mark_visible(P, means3D.data_ptr(), viewmatrix.data_ptr(), projmatrix.data_ptr(),
                                present.data_ptr(), _stream_ptr(device))
        N.check(rc, "mark_visible")
    return present


# The reference's extension module object, for code that reaches for ``_C`` directly.
_C = SimpleNamespace(
    rasterize_gaussians=rasterize_gaussians_forward,
    rasterize_gaussians_backward=rasterize_gaussians_backward,
    mark_visible=mark_visible,
)


class ShGradSink:
    """Where the SH gradient of a parameter tensor goes during view-batched training.

    The reference renders one view per optimiser step.  When a data-parallel rank renders several
    views per step (``hugs_b200.dp``), torch's AccumulateGrad adds every view's [P,16,3] gradient to
    ``param.grad`` with a separate elementwise kernel (768 B of traffic per Gaussian and view on top
    of the 192 B the backward pass wrote).  With a sink attached, the backward pass of every render
    whose ``shs`` argument IS ``param`` writes (first view after ``reset()``) or adds (later views;
    visible Gaussians only) its SH gradient directly into ``buffer``; autograd sees no gradient for
    ``shs``.  ``param.grad`` is pointed at ``buffer``."""

    def __init__(self, param, buffer=None):
        if not (param.is_leaf and param.dtype == torch.float32 and param.dim() == 3 and param.is_contiguous()):
            raise ValueError("an SH gradient sink needs a contiguous float32 [P,M,3] leaf tensor")
        self.param = param
        self.buffer = torch.zeros_like(param) if buffer is None else buffer
        if tuple(self.buffer.shape) != tuple(param.shape) or self.buffer.dtype != torch.float32 \
                or self.buffer.device != param.device or not self.buffer.is_contiguous():
            raise ValueError("sink buffer must match the parameter (contiguous float32, same shape and device)")
        self.fresh = True  # the next backward pass overwrites the buffer instead of adding to it
        param.grad = self.buffer

    def reset(self):
        """Start of a step: the next backward pass overwrites the buffer (no zero fill needed)."""
        self.fresh = True
        if self.param.grad is not self.buffer:
            self.param.grad = self.buffer


    def zero_(self):
        """Start of a step whose views run on several CUDA streams: clear the buffer and let EVERY
        backward pass add into it (additions commute; an overwriting first pass would have to be
        ordered before the others)."""
        self.buffer.zero_()
        self.fresh = False
        if self.param.grad is not self.buffer:
            self.param.grad = self.buffer


_sh_sinks = {}


def attach_sh_grad_sink(param, buffer=None):
    sink = ShGradSink(param, buffer)
    _sh_sinks[id(param)] = sink
    return sink


def detach_sh_grad_sink(param):
    _sh_sinks.pop(id(param), None)


def _sink_of(sh):
    sink = _sh_sinks.get(id(sh))
    return sink if (sink is not None and sink.param is sh) else None


class GradSink:
    """Where the gradient of ONE leaf parameter of a two-segment render goes during view-batched training
    (means3D / shs / opacities / scales / rotations of either segment): the per-Gaussian backward ADDS the
    rows of the visible Gaussians straight into ``buffer`` (hugs_b200_grad_sinks) and autograd sees no
    gradient for that input -- no zero rows written, no AccumulateGrad kernel per view and parameter.
    The owner zeroes ``buffer`` at the start of a step (``hugs_b200.dp.FlatGradBuffer.zero_``)."""

    def __init__(self, param, buffer=None):
        if not (param.is_leaf and param.dtype == torch.float32 and param.is_contiguous() and param.is_cuda):
            raise ValueError("a gradient sink needs a contiguous float32 CUDA leaf tensor")
        self.param = param
        self.buffer = torch.zeros_like(param) if buffer is None else buffer
        b = self.buffer
        if (tuple(b.shape) != tuple(param.shape) or b.dtype != torch.float32 or b.device != param.device
                or not b.is_contiguous() or b.data_ptr() % 16):
            raise ValueError("sink buffer must match the parameter (contiguous float32, same shape and device, "
                             "16-byte aligned)")
        param.grad = self.buffer


_grad_sinks = {}


def attach_grad_sink(param, buffer=None):
    sink = GradSink(param, buffer)
    _grad_sinks[id(param)] = sink
    return sink


def detach_grad_sink(param):
    _grad_sinks.pop(id(param), None)


def grad_sink_of(t):
    sink = _grad_sinks.get(id(t))
    return sink if (sink is not None and sink.param is t) else None


def rasterize_gaussians(
    means3D,
    means2D,
    sh,
    colors_precomp,
    opacities,
    scales,
    rotations,
    cov3Ds_precomp,
    raster_settings,
):
    return _RasterizeGaussians.apply(
        means3D,
        means2D,
        sh,
        colors_precomp,
        opacities,
        scales,
        rotations,
        cov3Ds_precomp,
        raster_settings,
    )


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(
        ctx,
        means3D,
        means2D,
        sh,
        colors_precomp,
        opacities,
        scales,
        rotations,
        cov3Ds_precomp,
        raster_settings,
    ):
        # Restructure arguments the way the native side expects them (reference __init__.py:59-80)
        args = (
            raster_settings.bg,
            means3D,
            colors_precomp,
            opacities,
            scales,
            rotations,
            raster_settings.scale_modifier,
            cov3Ds_precomp,
            raster_settings.viewmatrix,
            raster_settings.projmatrix,
            raster_settings.tanfovx,
            raster_settings.tanfovy,
            raster_settings.image_height,
            raster_settings.image_width,
            sh,
            raster_settings.sh_degree,
            raster_settings.campos,
            raster_settings.prefiltered,
            raster_settings.debug,
        )

        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)  # Copy them before they can be corrupted
            try:
                num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _C.rasterize_gaussians(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = _C.rasterize_gaussians(*args)

        # Keep relevant tensors for backward
        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.sh_sink = _sink_of(sh)
        ctx.clamp = clamp_active()   # the image handed out is clamped: backward masks its gradient
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _):
        num_rendered = ctx.num_rendered
        raster_settings = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
         geomBuffer, binningBuffer, imgBuffer) = ctx.saved_tensors

        args = (raster_settings.bg,
                means3D,
                radii,
                colors_precomp,
                scales,
                rotations,
                raster_settings.scale_modifier,
                cov3Ds_precomp,
                raster_settings.viewmatrix,
                raster_settings.projmatrix,
                raster_settings.tanfovx,
                raster_settings.tanfovy,
                grad_out_color,
                sh,
                raster_settings.sh_degree,
                raster_settings.campos,
                geomBuffer,
                num_rendered,
                binningBuffer,
                imgBuffer,
                raster_settings.debug)

        # view-batched training: the SH gradient goes straight into the attached sink (ShGradSink)
        sink = ctx.sh_sink
        kw = {} if sink is None else dict(sh_grad_out=sink.buffer, sh_grad_accumulate=not sink.fresh)
        if ctx.clamp:
            kw["clamp"] = True
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)  # Copy them before they can be corrupted
            try:
                (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp,
                 grad_sh, grad_scales, grad_rotations) = rasterize_gaussians_backward(*args, **kw)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp,
             grad_sh, grad_scales, grad_rotations) = rasterize_gaussians_backward(*args, **kw)
        if sink is not None:
            sink.fresh = False
            grad_sh = None  # delivered through the sink

        def present(inp, g):
            # the reference hands back [P,...] zero tensors even for inputs that were passed as the
            # empty placeholder torch.Tensor([]); autograd ignores those.  Return None instead.
            return g if (inp is not None and inp.numel() != 0) else None

        grads = (
            grad_means3D,
            grad_means2D,
            present(sh, grad_sh) if grad_sh is not None else None,
            present(colors_precomp, grad_colors_precomp),
            grad_opacities,
            present(scales, grad_scales),
            present(rotations, grad_rotations),
            present(cov3Ds_precomp, grad_cov3Ds_precomp),
            None,
        )
        return grads


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        # Mark visible points (based on frustum culling for camera) with a boolean
        with torch.no_grad():
            raster_settings = self.raster_settings
            visible = _C.mark_visible(
                positions,
                raster_settings.viewmatrix,
                raster_settings.projmatrix)

        return visible

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):

        raster_settings = self.raster_settings

        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')

        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])

        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])

        # Invoke C++/CUDA rasterization routine
        return rasterize_gaussians(
            means3D,
            means2D,
            shs,
            colors_precomp,
            opacities,
            scales,
            rotations,
            cov3D_precomp,
            raster_settings,
        )
