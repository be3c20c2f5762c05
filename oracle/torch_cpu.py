"""Pure-PyTorch CPU evaluation of the reference's rasterizer math + SMPL LBS.  TEST INFRASTRUCTURE ONLY.

BASELINE.json configs[0] ("CPU: SMPL LBS of 110k synthetic human Gaussians + pure-torch tile
alpha-composite at 256x256, reference math, no GPU") and north_star's "pure-PyTorch CPU evaluation of
the reference's rasterizer math plus LBS, timed on the box's host cores".  Only tests/ and bench.py's
cpu_baseline leg import this.  Where the reference ships the math as Python it is the reference's own
function that runs (oracle/_ref/hugs_ref, copied unmodified by oracle/build_ref.py):
    lbs_extra / matrix_to_quaternion   hugs/models/modules/lbs.py:19-73, hugs/utils/rotations.py:94-156
    eval_sh                            hugs/utils/spherical_harmonics.py:61-125
The CUDA-only parts are restated with torch ops, differentiable through autograd:
    preprocess      cuda_rasterizer/forward.cu:155-256 (+ computeCov3D :117-152, computeCov2D :73-115)
    tile composite  cuda_rasterizer/forward.cu:261-374 (alpha = min(0.99, o exp(power)), skip alpha < 1/255
                    and power > 0, stop when T (1 - alpha) < 1e-4), one 16x16 tile at a time
Pinned by tests/test_torch_cpu_oracle.py against the C oracle (itself pinned on the reference build's
golden vectors): image <= 1e-4, gradients <= 1e-3.
"""
import contextlib
import importlib.util
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF = os.path.join(_HERE, "_ref", "hugs_ref")
TILE = 16


@contextlib.contextmanager
def _cuda_is_identity():
    """hugs/utils/spherical_harmonics.py builds its constant tables with `.float().cuda()` at import;
    on the host that call is made a no-op for the duration of the import (the function itself takes the
    tables as arguments, so it runs on CPU tensors unchanged)."""
    orig = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        yield
    finally:
        torch.Tensor.cuda = orig


_sh_mod = None


def reference_sh_module():
    """The reference's own spherical_harmonics.py, or None when oracle/_ref has not been built."""
    global _sh_mod
    path = os.path.join(_REF, "spherical_harmonics.py")
    if _sh_mod is None and os.path.exists(path):
        spec = importlib.util.spec_from_file_location("hugs_ref_sh", path)
        mod = importlib.util.module_from_spec(spec)
        with _cuda_is_identity():
            spec.loader.exec_module(mod)
        _sh_mod = mod
    return _sh_mod


# real SH constants (cuda_rasterizer/auxiliary.h:19-37), used only when the reference module is absent
_C0 = 0.28209479177387814
_C1 = 0.4886025119029199
_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
       1.445305721320277, -0.5900435899266435]


def _eval_sh_restated(deg, sh, dirs):
    """forward.cu:20-70 with torch ops; sh [P, 3, M], dirs [P, 3] -> [P, 3]."""
    res = _C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[:, 0:1], dirs[:, 1:2], dirs[:, 2:3]
        res = res - _C1 * y * sh[..., 1] + _C1 * z * sh[..., 2] - _C1 * x * sh[..., 3]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            res = (res + _C2[0] * xy * sh[..., 4] + _C2[1] * yz * sh[..., 5] + _C2[2] * (2.0 * zz - xx - yy) * sh[..., 6]
                   + _C2[3] * xz * sh[..., 7] + _C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                res = (res + _C3[0] * y * (3 * xx - yy) * sh[..., 9] + _C3[1] * xy * z * sh[..., 10]
                       + _C3[2] * y * (4 * zz - xx - yy) * sh[..., 11] + _C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12]
                       + _C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + _C3[5] * z * (xx - yy) * sh[..., 14]
                       + _C3[6] * x * (xx - 3 * yy) * sh[..., 15])
    return res


def sh_to_rgb(deg, shs, means3D, campos):
    """computeColorFromSH (forward.cu:20-70): SH in the reference's [P, M, 3] layout -> clamped RGB [P, 3]."""
    dirs = means3D - campos[None]
    dirs = dirs / dirs.norm(dim=1, keepdim=True)
    sh = shs.transpose(1, 2)  # [P, 3, M], what eval_sh wants
    mod = reference_sh_module()
    if mod is not None:
        res = mod.eval_sh(deg, sh, dirs, mod.C0.cpu(), mod.C1.cpu(), mod.C2.cpu(), mod.C3.cpu(),
                          mod.C4.cpu())
    else:
        res = _eval_sh_restated(deg, sh, dirs)
    return torch.clamp_min(res + 0.5, 0.0)


def quat_to_rotmat(q):
    r, x, y, z = q.unbind(-1)
    return torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).reshape(-1, 3, 3)


def preprocess(means3D, scales, rotations, opacity, viewmatrix, projmatrix, W, H, tanfovx, tanfovy,
               scale_modifier=1.0):
    """forward.cu:155-256 -> dict(visible mask, xy [P,2], conic [P,3], opacity [P], depth [P], rect lo/hi)."""
    P = means3D.shape[0]
    ones = torch.ones(P, 1, dtype=means3D.dtype)
    hom = torch.cat([means3D, ones], 1)
    p_view = hom @ viewmatrix            # viewmatrix is W2C transposed: row vector times matrix
    p_hom = hom @ projmatrix
    p_w = 1.0 / (p_hom[:, 3] + 0.0000001)
    p_proj = p_hom[:, :3] * p_w[:, None]
    depth = p_view[:, 2]
    in_front = depth > 0.2
    # computeCov3D: Sigma = R S S^T R^T
    R = quat_to_rotmat(rotations)
    M = R * (scale_modifier * scales)[:, None, :]
    cov3 = M @ M.transpose(1, 2)
    # computeCov2D (EWA splatting, eq. 29/31 of Zwicker et al.)
    focal_x, focal_y = W / (2.0 * tanfovx), H / (2.0 * tanfovy)
    limx, limy = 1.3 * tanfovx, 1.3 * tanfovy
    tz = p_view[:, 2]
    tx = torch.clamp(p_view[:, 0] / tz, -limx, limx) * tz
    ty = torch.clamp(p_view[:, 1] / tz, -limy, limy) * tz
    zero = torch.zeros_like(tz)
    J = torch.stack([focal_x / tz, zero, -(focal_x * tx) / (tz * tz),
                     zero, focal_y / tz, -(focal_y * ty) / (tz * tz)], -1).reshape(P, 2, 3)
    Wm = viewmatrix[:3, :3].T            # world -> camera rotation
    T = J @ Wm[None]
    cov2 = T @ cov3 @ T.transpose(1, 2)
    a, b, c = cov2[:, 0, 0] + 0.3, cov2[:, 0, 1], cov2[:, 1, 1] + 0.3
    det = a * c - b * b
    ok = in_front & (det != 0.0)
    det_safe = torch.where(ok, det, torch.ones_like(det))
    conic = torch.stack([c / det_safe, -b / det_safe, a / det_safe], -1)
    mid = 0.5 * (a + c)
    lam = mid + torch.sqrt(torch.clamp_min(mid * mid - det, 0.1))
    radius = torch.ceil(3.0 * torch.sqrt(lam.detach()))
    xy = torch.stack([((p_proj[:, 0] + 1.0) * W - 1.0) * 0.5, ((p_proj[:, 1] + 1.0) * H - 1.0) * 0.5], -1)
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    xyd = xy.detach()
    lo = torch.stack([((xyd[:, 0] - radius) / TILE).to(torch.int64).clamp(0, gx),
                      ((xyd[:, 1] - radius) / TILE).to(torch.int64).clamp(0, gy)], -1)
    hi = torch.stack([((xyd[:, 0] + radius + TILE - 1) / TILE).to(torch.int64).clamp(0, gx),
                      ((xyd[:, 1] + radius + TILE - 1) / TILE).to(torch.int64).clamp(0, gy)], -1)
    visible = ok & ((hi[:, 0] - lo[:, 0]) * (hi[:, 1] - lo[:, 1]) > 0)
    return dict(visible=visible, xy=xy, conic=conic, opacity=opacity.reshape(-1), depth=depth.detach(), lo=lo, hi=hi)


def composite(pre, rgb, bg, W, H):
    """forward.cu:261-374, tile by tile: [3, H, W]."""
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    order = torch.argsort(pre["depth"], stable=True)
    order = order[pre["visible"][order]]
    lo, hi = pre["lo"][order], pre["hi"][order]
    rows = []
    py, px = torch.meshgrid(torch.arange(TILE), torch.arange(TILE), indexing="ij")
    for ty in range(gy):
        row = []
        in_row = (lo[:, 1] <= ty) & (hi[:, 1] > ty)
        for tx in range(gx):
            sel = order[in_row & (lo[:, 0] <= tx) & (hi[:, 0] > tx)]
            pix_x = (tx * TILE + px).reshape(-1).to(rgb.dtype)
            pix_y = (ty * TILE + py).reshape(-1).to(rgb.dtype)
            if sel.numel() == 0:
                row.append(bg[:, None].expand(3, TILE * TILE).reshape(3, TILE, TILE))
                continue
            xy, con, o, col = pre["xy"][sel], pre["conic"][sel], pre["opacity"][sel], rgb[sel]
            dx = xy[None, :, 0] - pix_x[:, None]           # [256, n]
            dy = xy[None, :, 1] - pix_y[:, None]
            power = -0.5 * (con[None, :, 0] * dx * dx + con[None, :, 2] * dy * dy) - con[None, :, 1] * dx * dy
            alpha = torch.clamp_max(o[None] * torch.exp(power), 0.99)
            valid = (power <= 0.0) & (alpha >= 1.0 / 255.0)
            a_eff = torch.where(valid, alpha, torch.zeros_like(alpha))
            one_minus = 1.0 - a_eff
            T_incl = torch.cumprod(one_minus, dim=1)
            T_excl = torch.cat([torch.ones_like(T_incl[:, :1]), T_incl[:, :-1]], 1)
            alive = torch.cumprod((T_incl.detach() >= 0.0001).to(rgb.dtype), dim=1)   # done once T(1-alpha) < 1e-4
            w = alive * a_eff * T_excl
            C = w @ col                                       # [256, 3]
            final_T = torch.prod(1.0 - alive * a_eff, dim=1)
            row.append((C + final_T[:, None] * bg[None]).T.reshape(3, TILE, TILE))
        rows.append(torch.cat(row, 2))
    return torch.cat(rows, 1)[:, :H, :W]


def render(means3D, scales, rotations, opacity, shs, degree, viewmatrix, projmatrix, campos, bg, W, H, tanfovx,
           tanfovy):
    """GaussianRasterizer forward on CPU tensors (autograd gives the backward)."""
    pre = preprocess(means3D, scales, rotations, opacity, viewmatrix, projmatrix, W, H, tanfovx, tanfovy)
    rgb = sh_to_rgb(degree, shs, means3D, campos)
    return composite(pre, rgb, bg, W, H)


def reference_lbs():
    """(lbs_extra, matrix_to_quaternion) of the reference when oracle/_ref is present, else None."""
    if not os.path.exists(os.path.join(_REF, "lbs.py")):
        return None
    import sys
    tests = os.path.join(os.path.dirname(_HERE), "tests")
    if tests not in sys.path:
        sys.path.insert(0, tests)
    import refutil
    lbs, rot = refutil.load_ref_lbs()
    return lbs.lbs_extra, rot.matrix_to_quaternion


def deform(A, xyz, weights, rotmat, scales, transl):
    """hugs/models/hugs_trimlp.py:470-518 with smpl_scale = 1: LBS of the human Gaussians."""
    ref = reference_lbs()
    if ref is not None:
        lbs_extra, m2q = ref
        pose = torch.zeros(1, 72)
        verts, _, T, _, _ = lbs_extra(A[None], xyz[None], None, weights, pose, disable_posedirs=True, pose2rot=True)
        T = T.squeeze(0)
        return verts.squeeze(0) + transl[None], m2q(T[:, :3, :3] @ rotmat), scales
    T = torch.einsum("nj,jab->nab", weights, A)
    out = (T[:, :3, :3] @ xyz[..., None])[..., 0] + T[:, :3, 3] + transl[None]
    import sys
    tests = os.path.join(os.path.dirname(_HERE), "tests")
    if tests not in sys.path:
        sys.path.insert(0, tests)
    import scenes
    return out, scenes._mat2quat(T[:, :3, :3] @ rotmat), scales
