"""Python face of the CPU oracle (oracle/splat_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs import this.
It restates the reference's forward / backward orchestration
(cuda_rasterizer/rasterizer_impl.cu:198-336, 340-434) on numpy arrays, one C call per stage.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")


def build(force=False):
    src = os.path.join(_HERE, "splat_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.oracle_higher_msb.restype = C.c_uint32
        _lib.oracle_higher_msb.argtypes = [C.c_uint32]
        _lib.oracle_inclusive_scan.restype = C.c_uint32
    return _lib


def _p(a):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"], "oracle arrays must be C-contiguous"
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    if a is None:
        return None
    a = np.asarray(a)
    if a.size == 0:
        return None
    return np.ascontiguousarray(a, dtype=np.float32)


def sort_bits(W, H):
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    return 32 + int(lib().oracle_higher_msb(C.c_uint32(tiles)))


def preprocess(means3D, scales, scale_modifier, rotations, opacities, shs, degree, cov3D_precomp,
               colors_precomp, viewmatrix, projmatrix, campos, W, H, tan_fovx, tan_fovy):
    means3D = _f32(means3D)
    P = means3D.shape[0]
    scales, rotations, opacities = _f32(scales), _f32(rotations), _f32(opacities)
    shs, cov3D_precomp, colors_precomp = _f32(shs), _f32(cov3D_precomp), _f32(colors_precomp)
    V, Pm, cam = _f32(viewmatrix).reshape(16), _f32(projmatrix).reshape(16), _f32(campos).reshape(3)
    M = shs.shape[1] if shs is not None else 0
    out = dict(
        radii=np.zeros(P, np.int32), means2D=np.zeros((P, 2), np.float32), depths=np.zeros(P, np.float32),
        cov3D=np.zeros((P, 6), np.float32), conic_opacity=np.zeros((P, 4), np.float32),
        rgb=np.zeros((P, 3), np.float32), clamped=np.zeros((P, 3), np.uint8),
        tiles_touched=np.zeros(P, np.uint32))
    lib().oracle_preprocess(
        C.c_int(P), C.c_int(int(degree)), C.c_int(M), _p(means3D), _p(scales), C.c_float(scale_modifier),
        _p(rotations), _p(opacities), _p(shs), _p(cov3D_precomp), _p(colors_precomp), _p(V), _p(Pm), _p(cam),
        C.c_int(W), C.c_int(H), C.c_float(tan_fovx), C.c_float(tan_fovy), _p(out["radii"]), _p(out["means2D"]),
        _p(out["depths"]), _p(out["cov3D"]), _p(out["conic_opacity"]), _p(out["rgb"]), _p(out["clamped"]),
        _p(out["tiles_touched"]))
    if cov3D_precomp is not None:
        out["cov3D"] = cov3D_precomp.reshape(P, 6).copy()
    return out


def mark_visible(means3D, viewmatrix):
    means3D = _f32(means3D)
    P = means3D.shape[0]
    present = np.zeros(P, np.uint8)
    lib().oracle_mark_visible(C.c_int(P), _p(means3D), _p(_f32(viewmatrix).reshape(16)), _p(present))
    return present.astype(bool)


def inclusive_scan(x):
    x = np.ascontiguousarray(x, dtype=np.uint32)
    out = np.zeros_like(x)
    lib().oracle_inclusive_scan(C.c_int(x.shape[0]), _p(x), _p(out))
    return out


def sort_pairs(keys, vals, end_bit):
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    vals = np.ascontiguousarray(vals, dtype=np.uint32)
    ko, vo = np.zeros_like(keys), np.zeros_like(vals)
    lib().oracle_sort_pairs(C.c_int64(keys.shape[0]), _p(keys), _p(vals), _p(ko), _p(vo), C.c_int(end_bit))
    return ko, vo


def bin_and_sort(pre, W, H):
    """scan + duplicateWithKeys + sort + identifyTileRanges (rasterizer_impl.cu:277-318)."""
    P = pre["radii"].shape[0]
    offsets = inclusive_scan(pre["tiles_touched"])
    R = int(offsets[-1]) if P > 0 else 0
    keys_u = np.zeros(R, np.uint64)
    vals_u = np.zeros(R, np.uint32)
    if R > 0:
        lib().oracle_duplicate(C.c_int(P), _p(pre["means2D"]), _p(pre["depths"]), _p(offsets), _p(pre["radii"]),
                               C.c_int(W), C.c_int(H), _p(keys_u), _p(vals_u))
    keys, vals = sort_pairs(keys_u, vals_u, sort_bits(W, H)) if R > 0 else (keys_u, vals_u)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    ranges = np.zeros((tiles, 2), np.uint32)
    lib().oracle_tile_ranges(C.c_int64(R), _p(keys), C.c_int(tiles), _p(ranges))
    return dict(point_offsets=offsets, num_rendered=R, keys_unsorted=keys_u, list_unsorted=vals_u,
                keys=keys, point_list=vals, ranges=ranges)


def render_forward(binning, pre, features, bg, W, H):
    final_T = np.zeros(H * W, np.float32)
    n_contrib = np.zeros(H * W, np.uint32)
    out_color = np.zeros((3, H, W), np.float32)
    features = _f32(features)
    lib().oracle_render_forward(_p(binning["ranges"]), _p(binning["point_list"]), C.c_int(W), C.c_int(H),
                                _p(pre["means2D"]), _p(features), _p(pre["conic_opacity"]), _p(final_T),
                                _p(n_contrib), _p(_f32(bg).reshape(3)), _p(out_color))
    return dict(color=out_color, final_T=final_T, n_contrib=n_contrib)


def forward(means3D, opacities, viewmatrix, projmatrix, campos, bg, W, H, tan_fovx, tan_fovy, shs=None, degree=0,
            colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0):
    """Rasterizer::forward restated (rasterizer_impl.cu:198-336).  Returns every intermediate."""
    pre = preprocess(means3D, scales, scale_modifier, rotations, opacities, shs, degree, cov3D_precomp,
                     colors_precomp, viewmatrix, projmatrix, campos, W, H, tan_fovx, tan_fovy)
    binning = bin_and_sort(pre, W, H)
    feats = _f32(colors_precomp) if colors_precomp is not None and np.asarray(colors_precomp).size else pre["rgb"]
    img = render_forward(binning, pre, feats, bg, W, H)
    res = dict(pre=pre, binning=binning, features=feats)
    res.update(img)
    return res


def backward(fwd, dL_dpix, means3D, viewmatrix, projmatrix, campos, bg, W, H, tan_fovx, tan_fovy, shs=None,
             degree=0, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0):
    """Rasterizer::backward restated (rasterizer_impl.cu:340-434).  Returns the reference's 8 gradient
    tensors (+ dL_dconic) as numpy arrays."""
    means3D = _f32(means3D)
    P = means3D.shape[0]
    pre, binning = fwd["pre"], fwd["binning"]
    shs, scales, rotations = _f32(shs), _f32(scales), _f32(rotations)
    has_colors = colors_precomp is not None and np.asarray(colors_precomp).size > 0
    M = shs.shape[1] if shs is not None else 0
    g = dict(
        dL_dmean2D=np.zeros((P, 3), np.float32), dL_dconic=np.zeros((P, 4), np.float32),
        dL_dopacity=np.zeros((P, 1), np.float32), dL_dcolor=np.zeros((P, 3), np.float32),
        dL_dmean3D=np.zeros((P, 3), np.float32), dL_dcov3D=np.zeros((P, 6), np.float32),
        dL_dsh=np.zeros((P, M, 3), np.float32), dL_dscale=np.zeros((P, 3), np.float32),
        dL_drot=np.zeros((P, 4), np.float32))
    dL_dpix = _f32(dL_dpix)
    lib().oracle_render_backward(
        _p(binning["ranges"]), _p(binning["point_list"]), C.c_int(W), C.c_int(H), _p(_f32(bg).reshape(3)),
        _p(pre["means2D"]), _p(pre["conic_opacity"]), _p(fwd["features"]), _p(fwd["final_T"]), _p(fwd["n_contrib"]),
        _p(dL_dpix), C.c_int(P), _p(g["dL_dmean2D"]), _p(g["dL_dconic"]), _p(g["dL_dopacity"]), _p(g["dL_dcolor"]))
    cov3D = _f32(cov3D_precomp).reshape(P, 6) if (cov3D_precomp is not None and np.asarray(cov3D_precomp).size) else pre["cov3D"]
    lib().oracle_preprocess_backward(
        C.c_int(P), C.c_int(int(degree)), C.c_int(M), _p(means3D), _p(pre["radii"]),
        _p(None if has_colors else shs), _p(pre["clamped"]), _p(scales), _p(rotations), C.c_float(scale_modifier),
        _p(np.ascontiguousarray(cov3D)), _p(_f32(viewmatrix).reshape(16)), _p(_f32(projmatrix).reshape(16)),
        _p(_f32(campos).reshape(3)), C.c_int(W), C.c_int(H), C.c_float(tan_fovx), C.c_float(tan_fovy),
        _p(g["dL_dmean2D"]), _p(g["dL_dconic"]), _p(g["dL_dmean3D"]), _p(g["dL_dcolor"]), _p(g["dL_dcov3D"]),
        _p(g["dL_dsh"]), _p(g["dL_dscale"]), _p(g["dL_drot"]))
    return g


def lbs_forward(A, xyz, weights, rotmat, scales=None, transl=None, smpl_scale=1.0, want_T=False):
    """One pose frame: A[J,4,4] -> (xyz'[N,3], rotq[N,4], scales'[N,3]|None, T[N,4,4]|None)."""
    A, xyz, weights, rotmat = _f32(A), _f32(xyz), _f32(weights), _f32(rotmat)
    Np, J = xyz.shape[0], A.shape[0]
    scales = _f32(scales)
    transl = _f32(transl)
    out_xyz = np.zeros((Np, 3), np.float32)
    out_q = np.zeros((Np, 4), np.float32)
    out_s = np.zeros((Np, 3), np.float32) if scales is not None else None
    out_T = np.zeros((Np, 4, 4), np.float32) if want_T else None
    lib().oracle_lbs_forward(C.c_int(Np), C.c_int(J), _p(A.reshape(J, 16)), _p(xyz), _p(weights),
                             _p(rotmat.reshape(Np, 9)), _p(scales), _p(transl), C.c_float(smpl_scale), _p(out_xyz),
                             _p(out_q), _p(out_s), _p(out_T))
    return out_xyz, out_q, out_s, out_T
