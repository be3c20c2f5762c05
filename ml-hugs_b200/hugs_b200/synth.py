"""Seeded synthetic inputs for the hot path.

Everything is generated on the CPU with ``torch.Generator().manual_seed(seed)`` and moved to the
requested device afterwards, so the same tensors feed the CUDA path, the reference build and the
CPU oracle.  Camera conventions are the reference's (hugs/datasets/neuman.py:345-362,
hugs/utils/graphics.py:76-96): ``viewmatrix = W2C.T``, ``projmatrix = (P @ W2C).T``,
``campos = inverse(viewmatrix)[3, :3]``.
"""
import math

import torch

SMPL_PARENTS = [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21]
N_HUMAN = 110210  # SMPL 6890 vertices, two midpoint subdivisions (hugs/utils/subdivide_smpl.py:126-133)


def projection_matrix(znear, zfar, fovx, fovy):
    """hugs/utils/graphics.py:76-96."""
    tan_y, tan_x = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = tan_y * znear, tan_x * znear
    bottom, left = -top, -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(W=1920, H=1080, focal=1400.0, view=0, n_views=1, seed=1000, device="cpu", orbit_deg=15.0):
    """Pinhole camera ``view`` of ``n_views`` on a +-orbit_deg orbit about (0,0,6); view 0 of 1 is
    the identity pose."""
    fovx = 2 * math.atan(W / (2 * focal))
    fovy = 2 * math.atan(H / (2 * focal))
    if n_views > 1:
        ang = math.radians(orbit_deg) * (2.0 * view / (n_views - 1) - 1.0)
    else:
        ang = 0.0
    c, s = math.cos(ang), math.sin(ang)
    Rm = torch.tensor([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=torch.float32)
    centre = torch.tensor([0.0, 0.0, 6.0])
    w2c = torch.eye(4)
    w2c[:3, :3] = Rm
    w2c[:3, 3] = centre - Rm @ centre
    viewmatrix = w2c.T.contiguous()
    P = projection_matrix(0.01, 100.0, fovx, fovy)
    projmatrix = (viewmatrix @ P.T).contiguous()
    campos = viewmatrix.inverse()[3, :3].contiguous()
    g = torch.Generator().manual_seed(seed)
    bg = torch.rand(3, generator=g)
    dev = torch.device(device)
    return dict(
        image_width=W, image_height=H, tanfovx=math.tan(fovx * 0.5), tanfovy=math.tan(fovy * 0.5),
        fovx=fovx, fovy=fovy, viewmatrix=viewmatrix.to(dev), projmatrix=projmatrix.to(dev),
        campos=campos.to(dev), bg=bg.to(dev))


def make_scene_gaussians(P, seed, tanfovx, tanfovy, sh_coeffs=16, device="cpu", near_frac=0.02):
    """Scene Gaussians in a widened view frustum (about 30 % off-screen, 2 % behind the near plane)."""
    g = torch.Generator().manual_seed(seed)
    z = 2.0 + 28.0 * torch.rand(P, generator=g)
    n_near = int(P * near_frac)
    if n_near > 0:
        z[:n_near] = -1.0 + 1.2 * torch.rand(n_near, generator=g)
    zz = z.abs().clamp_min(0.5)
    x = zz * tanfovx * (2.4 * torch.rand(P, generator=g) - 1.2)
    y = zz * tanfovy * (2.4 * torch.rand(P, generator=g) - 1.2)
    means = torch.stack([x, y, z], dim=1)
    perm = torch.randperm(P, generator=g)
    means = means[perm].contiguous()
    scales = torch.exp(math.log(0.03) + 0.5 * torch.randn(P, 3, generator=g))
    rot = torch.randn(P, 4, generator=g)
    rot = rot / rot.norm(dim=1, keepdim=True)
    opacity = torch.sigmoid(1.5 * torch.randn(P, 1, generator=g))
    shs = 0.1 * torch.randn(P, sh_coeffs, 3, generator=g)
    shs[:, 0, :] = 0.5 * torch.randn(P, 3, generator=g) / 0.28209479177387814
    dev = torch.device(device)
    return dict(means3D=means.to(dev), scales=scales.to(dev), rotations=rot.to(dev), opacity=opacity.to(dev),
                shs=shs.to(dev))


def rotation_6d_to_matrix(d6):
    """Gram-Schmidt 6D -> rotation matrix (rows), as hugs/utils/rotations.py:552-573."""
    a1, a2 = d6[..., :3], d6[..., 3:]
    b1 = torch.nn.functional.normalize(a1, dim=-1)
    b2 = a2 - (b1 * a2).sum(-1, keepdim=True) * b1
    b2 = torch.nn.functional.normalize(b2, dim=-1)
    b3 = torch.cross(b1, b2, dim=-1)
    return torch.stack((b1, b2, b3), dim=-2)


def rodrigues(rvec):
    """Axis-angle [J,3] -> rotation matrices [J,3,3]."""
    ang = rvec.norm(dim=1, keepdim=True).clamp_min(1e-8)
    k = rvec / ang
    K = torch.zeros(rvec.shape[0], 3, 3)
    K[:, 0, 1], K[:, 0, 2] = -k[:, 2], k[:, 1]
    K[:, 1, 0], K[:, 1, 2] = k[:, 2], -k[:, 0]
    K[:, 2, 0], K[:, 2, 1] = -k[:, 1], k[:, 0]
    s, c = torch.sin(ang)[..., None], torch.cos(ang)[..., None]
    return torch.eye(3)[None] + s * K + (1 - c) * (K @ K)


def make_joint_transforms(n_frames, seed, n_joints=24, walk=0.02):
    """A[F,J,4,4]: random rigid joint transforms (rotation rodrigues(0.3 N(0,1)), translation
    0.1 N(0,1)); frames f > 0 follow a smooth random walk of the pose."""
    g = torch.Generator().manual_seed(seed)
    rvec = 0.3 * torch.randn(n_joints, 3, generator=g)
    trans = 0.1 * torch.randn(n_joints, 3, generator=g)
    out = torch.zeros(n_frames, n_joints, 4, 4)
    for f in range(n_frames):
        if f > 0:
            rvec = rvec + walk * torch.randn(n_joints, 3, generator=g)
            trans = trans + 0.1 * walk * torch.randn(n_joints, 3, generator=g)
        out[f, :, :3, :3] = rodrigues(rvec)
        out[f, :, :3, 3] = trans
        out[f, :, 3, 3] = 1.0
    return out


def make_human_gaussians(N=N_HUMAN, seed=1001, n_joints=24, sh_coeffs=16, n_frames=1, device="cpu"):
    """Canonical human Gaussians + skinning inputs (mirrors what hugs_trimlp.forward hands to
    lbs_extra: softmax(./0.1) weights, rot6d-derived rotation matrices)."""
    g = torch.Generator().manual_seed(seed)
    box = torch.tensor([0.6, 1.8, 0.4])
    xyz = (torch.rand(N, 3, generator=g) - 0.5) * box
    rotmat = rotation_6d_to_matrix(torch.randn(N, 6, generator=g)).contiguous()
    scales = torch.exp(math.log(0.008) + 0.3 * torch.randn(N, 3, generator=g))
    weights = torch.softmax(torch.randn(N, n_joints, generator=g) / 0.1, dim=-1)
    opacity = torch.sigmoid(1.5 * torch.randn(N, 1, generator=g))
    shs = 0.1 * torch.randn(N, sh_coeffs, 3, generator=g)
    shs[:, 0, :] = 0.5 * torch.randn(N, 3, generator=g) / 0.28209479177387814
    A = make_joint_transforms(n_frames, seed + 7, n_joints)
    transl = torch.tensor([0.0, 0.0, 4.0]).expand(n_frames, 3).contiguous()
    dev = torch.device(device)
    return dict(xyz=xyz.to(dev), rotmat=rotmat.to(dev), scales=scales.to(dev), lbs_weights=weights.to(dev),
                opacity=opacity.to(dev), shs=shs.to(dev), A=A.to(dev), transl=transl.to(dev), smpl_scale=1.0)


def make_dL_dcolor(W, H, seed, device="cpu"):
    g = torch.Generator().manual_seed(seed + 99)
    return (2.0 * torch.rand(3, H, W, generator=g) - 1.0).to(torch.device(device))


# BASELINE.json configs (index == position in BASELINE.json "configs").
CONFIGS = {
    0: dict(name="cpu_256", W=256, H=256, focal=187.0, n_human=N_HUMAN, n_scene=0, backward=True),
    1: dict(name="human_only_1080p", W=1920, H=1080, focal=1400.0, n_human=N_HUMAN, n_scene=0, backward=True),
    2: dict(name="human_scene_1080p", W=1920, H=1080, focal=1400.0, n_human=N_HUMAN, n_scene=1_000_000, backward=True),
    3: dict(name="scene_3M_fwd", W=1920, H=1080, focal=1400.0, n_human=0, n_scene=3_000_000, backward=False),
    4: dict(name="anim_300_frames", W=1920, H=1080, focal=1400.0, n_human=N_HUMAN, n_scene=0, backward=False, n_frames=300),
}
