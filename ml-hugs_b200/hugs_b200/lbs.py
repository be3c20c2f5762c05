"""SMPL linear-blend-skinning deformation of the human Gaussians on the fused CUDA kernel.

Mirrors, for the hot path only:
  lbs_extra               hugs/models/modules/lbs.py:19-73   (same name, arguments, 5-tuple return)
  deformation tail        hugs/models/hugs_trimlp.py:510-518 (scale, translate, R' = T[:3,:3] R, quat)
  matrix_to_quaternion    hugs/utils/rotations.py:105-156

``deform_gaussians`` is the fused replacement for the whole chain (one kernel forward, one
backward); ``lbs_extra`` keeps the reference's signature for callers that want the 4x4 per-Gaussian
transforms materialised.
"""
import ctypes as C

import torch

from . import _native as N


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _c(t):
    return t.contiguous() if t.dtype == torch.float32 else t.float().contiguous()


class _LbsDeform(torch.autograd.Function):
    """(A[F,J,4,4], xyz[N,3], weights[N,J], rotmat[N,3,3], scales[N,3]|None, transl[F,3]|None,
    smpl_scale tensor[1]|float) -> (xyz'[F,N,3], rotq[F,N,4], scales'[F,N,3]|None)"""

    @staticmethod
    def forward(ctx, A, xyz, weights, rotmat, scales, transl, smpl_scale):
        if not xyz.is_cuda:
            raise RuntimeError("hugs_b200 has no CPU path: LBS inputs must be CUDA tensors")
        device = xyz.device
        A, xyz, weights, rotmat = _c(A), _c(xyz), _c(weights), _c(rotmat)
        F, J = int(A.shape[0]), int(A.shape[1])
        Np = int(xyz.shape[0])
        if A.shape[-2:] != (4, 4) or weights.shape != (Np, J) or rotmat.shape != (Np, 3, 3):
            raise RuntimeError("lbs: expected A[F,J,4,4], xyz[N,3], weights[N,J], rotmat[N,3,3]")
        scales_c = _c(scales) if scales is not None else None
        transl_c = _c(transl).reshape(F, 3) if transl is not None else None
        s_val = float(smpl_scale) if not isinstance(smpl_scale, torch.Tensor) else float(smpl_scale.reshape(-1)[0])
        out_xyz = torch.empty(F, Np, 3, dtype=torch.float32, device=device)
        out_rotq = torch.empty(F, Np, 4, dtype=torch.float32, device=device)
        out_scales = torch.empty(F, Np, 3, dtype=torch.float32, device=device) if scales is not None else None
        with torch.cuda.device(device):
            rc = N.lbs_forward(F, Np, J, A.data_ptr(), N.ptr(xyz), N.ptr(weights), N.ptr(rotmat), N.ptr(scales_c),
                               N.ptr(transl_c), s_val, N.ptr(out_xyz), N.ptr(out_rotq), N.ptr(out_scales), None,
                               _stream_ptr(device))
        N.check(rc, "lbs_forward")
        ctx.save_for_backward(A, xyz, weights, rotmat, scales_c, transl_c)
        ctx.s_val = s_val
        ctx.scale_is_tensor = isinstance(smpl_scale, torch.Tensor)
        ctx.scale_shape = smpl_scale.shape if ctx.scale_is_tensor else None
        ctx.transl_shape = transl.shape if transl is not None else None
        if out_scales is None:
            out_scales = torch.empty(0, device=device)
            ctx.mark_non_differentiable(out_scales)
        return out_xyz, out_rotq, out_scales

    @staticmethod
    def backward(ctx, g_xyz, g_rotq, g_scales):
        A, xyz, weights, rotmat, scales, transl = ctx.saved_tensors
        device = xyz.device
        F, J = int(A.shape[0]), int(A.shape[1])
        Np = int(xyz.shape[0])
        zeros = lambda *s: torch.zeros(*s, dtype=torch.float32, device=device)  # noqa: E731
        g_xyz = _c(g_xyz) if g_xyz is not None else zeros(F, Np, 3)
        g_rotq = _c(g_rotq) if g_rotq is not None else zeros(F, Np, 4)
        has_scales = scales is not None
        if has_scales:
            g_scales = _c(g_scales) if g_scales is not None else zeros(F, Np, 3)
        else:
            g_scales = None
        dA = torch.empty(F, J, 4, 4, dtype=torch.float32, device=device)
        d_xyz = torch.empty(Np, 3, dtype=torch.float32, device=device)
        d_w = torch.empty(Np, J, dtype=torch.float32, device=device)
        d_rot = torch.empty(Np, 3, 3, dtype=torch.float32, device=device)
        d_scales = torch.empty(Np, 3, dtype=torch.float32, device=device) if has_scales else None
        d_transl = torch.empty(F, 3, dtype=torch.float32, device=device)
        d_s = torch.empty(1, dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            rc = N.lbs_backward(F, Np, J, A.data_ptr(), N.ptr(xyz), N.ptr(weights), N.ptr(rotmat), N.ptr(scales),
                                N.ptr(transl), ctx.s_val, N.ptr(g_xyz), N.ptr(g_rotq), N.ptr(g_scales),
                                dA.data_ptr(), N.ptr(d_xyz), N.ptr(d_w), N.ptr(d_rot), N.ptr(d_scales),
                                d_transl.data_ptr(), d_s.data_ptr(), _stream_ptr(device))
        N.check(rc, "lbs_backward")
        g_transl = d_transl.reshape(ctx.transl_shape) if ctx.transl_shape is not None else None
        g_s = d_s.reshape(ctx.scale_shape) if ctx.scale_is_tensor else None
        return dA, d_xyz, d_w, d_rot, d_scales, g_transl, g_s


def lbs_deform_frames(A, xyz, lbs_weights, rotmat, scales=None, transl=None, smpl_scale=1.0):
    """Batched-over-pose-frames deformation: A[F,J,4,4] -> (xyz'[F,N,3], rotq[F,N,4], scales'[F,N,3]).
    The animation path (hugs/trainer/gs_trainer.py:686-747: canonical attributes once, then LBS +
    render per pose frame)."""
    out_xyz, out_rotq, out_scales = _LbsDeform.apply(A, xyz, lbs_weights, rotmat, scales, transl, smpl_scale)
    return out_xyz, out_rotq, (out_scales if scales is not None else None)


def deform_gaussians(A, gs_xyz, lbs_weights, gs_rotmat, gs_scales=None, smpl_scale=None, transl=None):
    """Fused replacement of hugs/models/hugs_trimlp.py:470-518 for one pose:

        deformed_xyz, _, lbs_T, _, _ = lbs_extra(A[None], gs_xyz[None], None, lbs_weights, pose,
                                                 disable_posedirs=True)
        deformed_xyz = deformed_xyz * smpl_scale + transl ; gs_scales = gs_scales * smpl_scale
        deformed_gs_rotq = matrix_to_quaternion(lbs_T[:, :3, :3] @ gs_rotmat)

    A: [J,4,4] (A_vitruvian2pose). Returns (deformed_xyz[N,3], deformed_gs_rotq[N,4], gs_scales'[N,3]|None).
    """
    s = 1.0 if smpl_scale is None else smpl_scale
    t = transl.reshape(1, 3) if transl is not None else None
    out_xyz, out_rotq, out_scales = _LbsDeform.apply(A[None], gs_xyz, lbs_weights, gs_rotmat, gs_scales, t, s)
    return out_xyz[0], out_rotq[0], (out_scales[0] if gs_scales is not None else None)


class _LbsT(torch.autograd.Function):
    """verts and materialised per-Gaussian 4x4 transforms, for the reference-shaped lbs_extra."""

    @staticmethod
    def forward(ctx, A, v_posed, weights):
        device = v_posed.device
        A, v_posed, weights = _c(A), _c(v_posed), _c(weights)
        F, J = int(A.shape[0]), int(A.shape[1])
        Np = int(v_posed.shape[0])
        eye = torch.eye(3, dtype=torch.float32, device=device).expand(Np, 3, 3).contiguous()
        out_xyz = torch.empty(F, Np, 3, dtype=torch.float32, device=device)
        out_rotq = torch.empty(F, Np, 4, dtype=torch.float32, device=device)
        out_T = torch.empty(F, Np, 4, 4, dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            rc = N.lbs_forward(F, Np, J, A.data_ptr(), N.ptr(v_posed), N.ptr(weights), N.ptr(eye), None, None, 1.0,
                               N.ptr(out_xyz), N.ptr(out_rotq), None, N.ptr(out_T), _stream_ptr(device))
        N.check(rc, "lbs_forward")
        ctx.save_for_backward(A, v_posed, weights)
        return out_xyz, out_T

    @staticmethod
    def backward(ctx, g_verts, g_T):
        # T = W A and verts = T[:3,:] [v;1] are (bi)linear: plain tensor algebra on the saved inputs.
        A, v, w = ctx.saved_tensors
        F, J = A.shape[0], A.shape[1]
        gT = g_T.clone() if g_T is not None else torch.zeros(F, v.shape[0], 4, 4, device=v.device)
        dv = None
        if g_verts is not None:
            vh = torch.cat([v, torch.ones_like(v[:, :1])], dim=1)  # [N,4]
            gT[:, :, :3, :] += g_verts[..., None] * vh[None, :, None, :]
            T3 = torch.matmul(w, A.reshape(F, J, 16)).reshape(F, -1, 4, 4)[:, :, :3, :3]
            dv = torch.einsum("fnrc,fnr->nc", T3, g_verts)
        gTf = gT.reshape(F, -1, 16)
        dA = torch.einsum("nj,fnk->fjk", w, gTf).reshape(F, J, 4, 4)
        dw = torch.einsum("fnk,fjk->nj", gTf, A.reshape(F, J, 16))
        return dA, dv, dw


def lbs_extra(A, v_shaped, posedirs, lbs_weights, pose, disable_posedirs=False, pose2rot=True):
    """Same signature and return tuple as the reference's lbs_extra (hugs/models/modules/lbs.py:19-73):
    returns (verts[B,N,3], A, T[B,N,4,4], v_posed, v_shaped).  The skinning step (lbs.py:59-72) runs on the
    CUDA kernel; the optional pose-corrective blend shapes (lbs.py:31-52, off in the release config:
    disable_posedirs=True) stay a small torch matmul ahead of it."""
    batch_size = A.shape[0]
    if not disable_posedirs:
        ident = torch.eye(3, dtype=A.dtype, device=A.device)
        if pose2rot:
            rot_mats = batch_rodrigues(pose.view(-1, 3)).view(batch_size, -1, 3, 3)
            pose_feature = (rot_mats[:, 1:, :, :] - ident).view(batch_size, -1)
        else:
            pose_feature = (pose[:, 1:].view(batch_size, -1, 3, 3) - ident).view(batch_size, -1)
        pose_offsets = torch.matmul(pose_feature, posedirs).view(batch_size, -1, 3)
        v_posed = pose_offsets + v_shaped
    else:
        v_posed = v_shaped
    verts_list, T_list = [], []
    for b in range(batch_size):
        vb = v_posed[b if v_posed.shape[0] == batch_size else 0]
        verts_b, T_b = _LbsT.apply(A[b:b + 1], vb, lbs_weights)
        verts_list.append(verts_b[0])
        T_list.append(T_b[0])
    verts = torch.stack(verts_list, dim=0)
    T = torch.stack(T_list, dim=0)
    return verts, A, T, v_posed, v_shaped


def batch_rodrigues(rot_vecs, epsilon=1e-8):
    """Axis-angle -> rotation matrices, the published smplx 0.1.28 formula (smplx/lbs.py
    batch_rodrigues; the reference imports it at hugs/models/modules/lbs.py:9-14 and vendors an
    equivalent at hugs/datasets/neuman_utils/smpl.py:363-395).  Only reached when posedirs are on."""
    batch_size = rot_vecs.shape[0]
    angle = torch.norm(rot_vecs + epsilon, dim=1, keepdim=True)
    rot_dir = rot_vecs / angle
    cos = torch.unsqueeze(torch.cos(angle), dim=1)
    sin = torch.unsqueeze(torch.sin(angle), dim=1)
    rx, ry, rz = torch.split(rot_dir, 1, dim=1)
    zeros = torch.zeros((batch_size, 1), dtype=rot_vecs.dtype, device=rot_vecs.device)
    K = torch.cat([zeros, -rz, ry, rz, zeros, -rx, -ry, rx, zeros], dim=1).view((batch_size, 3, 3))
    ident = torch.eye(3, dtype=rot_vecs.dtype, device=rot_vecs.device).unsqueeze(dim=0)
    return ident + sin * K + (1 - cos) * torch.bmm(K, K)


def matrix_to_quaternion(matrix):
    """hugs/utils/rotations.py:105-156 on the CUDA kernel (identity joint transform): [N,3,3] -> [N,4],
    real part first, not normalised."""
    lead = matrix.shape[:-2]
    m = matrix.reshape(-1, 3, 3)
    n = m.shape[0]
    A = torch.eye(4, dtype=torch.float32, device=m.device).reshape(1, 1, 4, 4)
    w = torch.ones(n, 1, dtype=torch.float32, device=m.device)
    xyz = torch.zeros(n, 3, dtype=torch.float32, device=m.device)
    _, q, _ = _LbsDeform.apply(A, xyz, w, m, None, None, 1.0)
    return q[0].reshape(lead + (4,))
