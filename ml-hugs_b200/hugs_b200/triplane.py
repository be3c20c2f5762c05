"""Tri-plane feature sampler on the CUDA kernels (the step right before the path).

Mirrors hugs/models/modules/triplane.py:14-39 -- same class name, constructor arguments, parameter
names and shapes (so a reference state_dict loads), same forward contract (x [...,3] in the box
center +- scale/2 -> features [..., 3*features]) and the same range assertion.  The three planes are
kept in torch.channels_last memory format: logically [1,C,R,R] like the reference's, physically
[R][R][C], the layout the kernels read with one coalesced request per bilinear corner.
"""
import ctypes as C

import torch
import torch.nn as nn

from . import _native as N

EPS = 1e-3


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _texels(plane):
    """[1,C,H,W] parameter -> dense [H,W,C] view (a copy only if it is not channels_last already)."""
    v = plane[0].permute(1, 2, 0)
    return v if v.is_contiguous() else v.contiguous()


class _TriPlaneSample(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, plane_xy, plane_xz, plane_yz, center, scale, check_range):
        if not (xyz.is_cuda and plane_xy.is_cuda and plane_xz.is_cuda and plane_yz.is_cuda):
            raise RuntimeError("hugs_b200 has no CPU path: TriPlane inputs must be CUDA tensors")
        device = xyz.device
        pts = xyz.detach().float().contiguous()
        planes = [_texels(p.detach().float()) for p in (plane_xy, plane_xz, plane_yz)]
        Cn = int(plane_xy.shape[1])
        resX, resY, resZ = int(plane_xy.shape[2]), int(plane_xy.shape[3]), int(plane_xz.shape[3])
        if tuple(plane_xz.shape) != (1, Cn, resX, resZ) or tuple(plane_yz.shape) != (1, Cn, resY, resZ):
            raise RuntimeError("TriPlane: plane shapes must be [1,C,X,Y], [1,C,X,Z], [1,C,Y,Z]")
        n = int(pts.shape[0])
        out = torch.empty(n, 3 * Cn, dtype=torch.float32, device=device)
        flag = torch.zeros(1, dtype=torch.int32, device=device) if check_range else None
        with torch.cuda.device(device):
            rc = N.triplane_forward(n, Cn, resX, resY, resZ, planes[0].data_ptr(), planes[1].data_ptr(),
                                    planes[2].data_ptr(), N.ptr(pts), float(center), float(scale), EPS,
                                    N.ptr(out), N.ptr(flag), _stream_ptr(device))
        N.check(rc, "triplane_forward")
        if check_range and n > 0 and int(flag.item()) != 0:
            t = (pts - center) / scale + 0.5
            raise AssertionError(f"x must be in [0, 1], got {t.min()} and {t.max()}")
        ctx.save_for_backward(pts, *planes)
        ctx.meta = (n, Cn, resX, resY, resZ, float(center), float(scale))
        return out

    @staticmethod
    def backward(ctx, g_out):
        pts, pxy, pxz, pyz = ctx.saved_tensors
        n, Cn, resX, resY, resZ, center, scale = ctx.meta
        device = pts.device
        need_x = ctx.needs_input_grad[0]
        need_p = any(ctx.needs_input_grad[1:4])
        g = g_out.detach().float().contiguous()
        gx = torch.empty_like(pts) if need_x else None
        if need_p:      # one allocation and one memset for the three plane gradients
            sizes = [p.numel() for p in (pxy, pxz, pyz)]
            flat = torch.zeros(sum(sizes), dtype=torch.float32, device=device)
            gp = [c.view(p.shape) for c, p in zip(flat.split(sizes), (pxy, pxz, pyz))]
        else:
            gp = [None, None, None]
        with torch.cuda.device(device):
            rc = N.triplane_backward(n, Cn, resX, resY, resZ, pxy.data_ptr(), pxz.data_ptr(), pyz.data_ptr(),
                                     N.ptr(pts), center, scale, N.ptr(g), N.ptr(gp[0]), N.ptr(gp[1]), N.ptr(gp[2]),
                                     N.ptr(gx), _stream_ptr(device))
        N.check(rc, "triplane_backward")
        # [H,W,C] -> logical [1,C,H,W]: a channels_last tensor, no copy
        grads = [t.permute(2, 0, 1).unsqueeze(0) if t is not None else None for t in gp]
        grads = [t if need else None for t, need in zip(grads, ctx.needs_input_grad[1:4])]
        return gx, grads[0], grads[1], grads[2], None, None, None


def sample_triplane(x, plane_xy, plane_xz, plane_yz, center=0.0, scale=2.0, check_range=True):
    """Functional form: x [...,3] -> [..., 3C]."""
    shape = x.shape
    feat = _TriPlaneSample.apply(x.reshape(-1, 3), plane_xy, plane_xz, plane_yz, center, scale, check_range)
    return feat.reshape(*shape[:-1], feat.shape[-1])


class TriPlane(nn.Module):
    """hugs/models/modules/triplane.py:14-39.  ``check_range=False`` skips the reference's range
    assertion (and the 4-byte device->host read it needs) for asynchronous use."""

    def __init__(self, features=32, resX=256, resY=256, resZ=256, check_range=True):
        super().__init__()
        cl = torch.channels_last
        self.plane_xy = nn.Parameter(torch.randn(1, features, resX, resY).contiguous(memory_format=cl))
        self.plane_xz = nn.Parameter(torch.randn(1, features, resX, resZ).contiguous(memory_format=cl))
        self.plane_yz = nn.Parameter(torch.randn(1, features, resY, resZ).contiguous(memory_format=cl))
        self.dim = features
        self.n_input_dims = 3
        self.n_output_dims = 3 * features
        self.center = 0.0
        self.scale = 2.0
        self.check_range = check_range

    def forward(self, x):
        return sample_triplane(x, self.plane_xy, self.plane_xz, self.plane_yz, self.center, self.scale,
                               self.check_range)
