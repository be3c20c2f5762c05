"""Image-space losses on the fused CUDA kernels (the step right after the
rasterizer in the reference's training loop).

Mirrors, for the unmasked / size-averaged path the trainer uses on full images:
  l1_loss   hugs/losses/utils.py:52-56
  ssim      hugs/losses/utils.py:74-83, _ssim :86-106 (window 11, sigma 1.5, zero padding)
with the same names and arguments.  ``l1_ssim`` returns both means from ONE pass over the images
(what hugs/losses/loss.py:87-107 computes with two separate calls); its backward is one kernel that
writes dL/dimage in the [3,H,W] layout the rasterizer's backward reads.
"""
import ctypes as C

import torch

from . import _native as N


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class _L1Ssim(torch.autograd.Function):
    """(img1[...,H,W], img2[...,H,W]) -> (mean |img1 - img2|, mean SSIM(img1, img2)); gradient w.r.t. img1."""

    @staticmethod
    def forward(ctx, img1, img2):
        if not (img1.is_cuda and img2.is_cuda):
            raise RuntimeError("hugs_b200 has no CPU path: loss inputs must be CUDA tensors")
        if img1.shape != img2.shape or img1.dim() < 2:
            raise RuntimeError("l1_ssim: images must have the same [..., H, W] shape")
        device = img1.device
        a = img1.detach().float().contiguous()
        b = img2.detach().to(device).float().contiguous()
        H, W = int(a.shape[-2]), int(a.shape[-1])
        Cn = int(a.numel() // max(H * W, 1)) if H * W > 0 else 0
        need_grad = ctx.needs_input_grad[0]
        dmap = torch.empty((3,) + tuple(a.shape), dtype=torch.float32, device=device) if need_grad else None
        sums = torch.empty(2, dtype=torch.float64, device=device)
        with torch.cuda.device(device):
            rc = N.l1_ssim_forward(Cn, H, W, N.ptr(a), N.ptr(b), N.ptr(dmap), sums.data_ptr(), _stream_ptr(device))
        N.check(rc, "l1_ssim_forward")
        means = (sums / float(max(a.numel(), 1))).float()
        ctx.save_for_backward(a, b, dmap)
        ctx.dims = (Cn, H, W)
        ctx.in_dtype = img1.dtype
        return means[0], means[1]

    @staticmethod
    def backward(ctx, g_l1, g_ssim):
        a, b, dmap = ctx.saved_tensors
        Cn, H, W = ctx.dims
        device = a.device
        z = torch.zeros((), dtype=torch.float32, device=device)
        g = torch.stack([(g_l1 if g_l1 is not None else z).float().reshape(()),
                         (g_ssim if g_ssim is not None else z).float().reshape(())]).contiguous()
        grad = torch.empty_like(a)
        if a.numel() > 0:
            with torch.cuda.device(device):
                rc = N.l1_ssim_backward(Cn, H, W, a.data_ptr(), b.data_ptr(), dmap.data_ptr(), g.data_ptr(),
                                        grad.data_ptr(), _stream_ptr(device))
            N.check(rc, "l1_ssim_backward")
        return grad.to(ctx.in_dtype), None


def l1_ssim(img1, img2):
    """(l1 mean, ssim mean) of two [C,H,W] (or [B,C,H,W]) images in one pass."""
    return _L1Ssim.apply(img1, img2)


def l1_loss(network_output, gt, mask=None):
    """hugs/losses/utils.py:52-56."""
    l1, _ = _L1Ssim.apply(network_output, gt)
    if mask is not None:
        return l1 * float(network_output.numel()) / mask.sum()
    return l1


def ssim(img1, img2, window_size=11, size_average=True, mask=None):
    """hugs/losses/utils.py:74-83.  ``mask`` is accepted and ignored, as in the reference (_ssim never
    reads it); only the trainer's configuration (window 11, size_average) runs on the fused kernel."""
    if window_size != 11 or not size_average:
        raise NotImplementedError("hugs_b200.losses.ssim implements window_size=11, size_average=True")
    _, s = _L1Ssim.apply(img1, img2)
    return s
