"""CPU oracle of the tri-plane sampler (TEST INFRASTRUCTURE: imported only by tests/ -- never by the
product).

Restates TriPlane.forward (hugs/models/modules/triplane.py:25-39) in numpy: the float32 normalisation
g = ((x - center)/scale + 0.5)*2 - 1, then F.grid_sample(mode='bilinear', padding_mode='zeros',
align_corners=True) of plane_xy at (g.x, g.y), plane_xz at (g.x, g.z), plane_yz at (g.y, g.z), the
first grid coordinate indexing the LAST plane axis, concatenated xy | xz | yz.  grid_sample is restated
from its published definition (ATen grid_sampler_2d: ix = ((g+1)/2)*(W-1), corners floor/floor+1,
weights (x1-ix)(y1-iy)..., out-of-range corners contribute zero); values in float32 as the kernel
computes them, gradients (what autograd gives through the reference) in float64.
Pinned against tests/golden/triplane_golden.npz, which tests/golden/make_golden_triplane.py generated
by importing the reference's own module.
"""
import numpy as np

F32 = np.float32


def grid_of(x, center=0.0, scale=2.0):
    x = np.asarray(x, F32)
    t = ((x - F32(center)) / F32(scale) + F32(0.5)).astype(F32)
    return (t * F32(2) - F32(1)).astype(F32)


def _corners(u, v, H, W):
    ix = (((u + F32(1)) / F32(2)) * F32(W - 1)).astype(F32)
    iy = (((v + F32(1)) / F32(2)) * F32(H - 1)).astype(F32)
    x0 = np.floor(ix).astype(np.int64)
    y0 = np.floor(iy).astype(np.int64)
    return ix, iy, x0, y0


def _sample(plane, u, v):
    """plane [C,H,W] float32; u indexes W, v indexes H.  Returns [N,C] float32."""
    C, H, W = plane.shape
    ix, iy, x0, y0 = _corners(u, v, H, W)
    wx1, wx0 = (ix - x0.astype(F32)).astype(F32), ((x0 + 1).astype(F32) - ix).astype(F32)
    wy1, wy0 = (iy - y0.astype(F32)).astype(F32), ((y0 + 1).astype(F32) - iy).astype(F32)
    out = np.zeros((len(u), C), F32)
    for yy, xx, w in ((y0, x0, wx0 * wy0), (y0, x0 + 1, wx1 * wy0), (y0 + 1, x0, wx0 * wy1), (y0 + 1, x0 + 1, wx1 * wy1)):
        ok = (yy >= 0) & (yy < H) & (xx >= 0) & (xx < W)
        val = plane[:, np.clip(yy, 0, H - 1), np.clip(xx, 0, W - 1)].T          # [N,C]
        out = np.where(ok[:, None], (out.astype(np.float64) + val.astype(np.float64) * w.astype(F32)[:, None]), out).astype(F32)
    return out


PLANE_AXES = ((0, 1), (0, 2), (1, 2))       # (u, v) coordinate index for xy, xz, yz


def forward(x, plane_xy, plane_xz, plane_yz, center=0.0, scale=2.0):
    """x [N,3]; planes [1,C,A,B].  Returns [N,3C] float32."""
    g = grid_of(x, center, scale)
    return np.concatenate([_sample(np.asarray(p, F32)[0], g[:, a], g[:, b])
                           for p, (a, b) in zip((plane_xy, plane_xz, plane_yz), PLANE_AXES)], axis=1)


def backward(x, plane_xy, plane_xz, plane_yz, g_out, center=0.0, scale=2.0):
    """Returns (dL/dx [N,3], dL/dplane_xy, dL/dplane_xz, dL/dplane_yz) in float64."""
    g = grid_of(x, center, scale)
    N = g.shape[0]
    gx = np.zeros((N, 3), np.float64)
    gplanes = []
    for k, (p, (a, b)) in enumerate(zip((plane_xy, plane_xz, plane_yz), PLANE_AXES)):
        plane = np.asarray(p, np.float64)[0]
        C, H, W = plane.shape
        go = np.asarray(g_out, np.float64)[:, k * C:(k + 1) * C]        # [N,C]
        ix, iy, x0, y0 = _corners(g[:, a], g[:, b], H, W)
        ix, iy = ix.astype(np.float64), iy.astype(np.float64)
        wx1, wx0, wy1, wy0 = ix - x0, x0 + 1 - ix, iy - y0, y0 + 1 - iy
        gp = np.zeros_like(plane)
        vals = {}
        for name, yy, xx, w in (("00", y0, x0, wx0 * wy0), ("01", y0, x0 + 1, wx1 * wy0),
                                ("10", y0 + 1, x0, wx0 * wy1), ("11", y0 + 1, x0 + 1, wx1 * wy1)):
            ok = (yy >= 0) & (yy < H) & (xx >= 0) & (xx < W)
            yc, xc = np.clip(yy, 0, H - 1), np.clip(xx, 0, W - 1)
            vals[name] = np.where(ok[:, None], plane[:, yc, xc].T, 0.0)
            contrib = (go * (w * ok)[:, None]).T                      # [C,N]
            for c in range(C):
                np.add.at(gp[c], (yc, xc), contrib[c])
        gix = (go * ((vals["01"] - vals["00"]) * wy0[:, None] + (vals["11"] - vals["10"]) * wy1[:, None])).sum(1)
        giy = (go * ((vals["10"] - vals["00"]) * wx0[:, None] + (vals["11"] - vals["01"]) * wx1[:, None])).sum(1)
        gx[:, a] += gix * 0.5 * (W - 1) * 2.0 / scale
        gx[:, b] += giy * 0.5 * (H - 1) * 2.0 / scale
        gplanes.append(gp[None])
    return (gx, *gplanes)
