"""Helpers shared by the parity tests: loading the reference build (oracle/_ref), decoding its
opaque buffers, decoding ours, and comparison metrics.  Test infrastructure only."""
import importlib
import importlib.util
import ctypes
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def ref_available():
    pkg = os.path.join(REF_DIR, "diff_gaussian_rasterization")
    return os.path.isdir(pkg) and any(f.startswith("_C") and f.endswith(".so") for f in os.listdir(pkg))


def load_ref_rasterizer():
    """The unmodified reference package, imported under the alias ref_dgr so that it does not
    collide with our drop-in ``diff_gaussian_rasterization``."""
    if "ref_dgr" in sys.modules:
        return sys.modules["ref_dgr"]
    pkg_dir = os.path.join(REF_DIR, "diff_gaussian_rasterization")
    spec = importlib.util.spec_from_file_location(
        "ref_dgr", os.path.join(pkg_dir, "__init__.py"), submodule_search_locations=[pkg_dir])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_dgr"] = mod
    spec.loader.exec_module(mod)
    return mod


def ref_knn_available():
    pkg = os.path.join(REF_DIR, "simple_knn")
    return os.path.isdir(pkg) and any(f.startswith("_C") and f.endswith(".so") for f in os.listdir(pkg))


def load_ref_knn():
    """The unmodified reference simple-knn extension (oracle/_ref/simple_knn/_C*.so) -> its distCUDA2."""
    if "ref_simple_knn_C" not in sys.modules:
        import torch  # noqa: F401  (the extension links against libtorch)
        pkg = os.path.join(REF_DIR, "simple_knn")
        so = [f for f in os.listdir(pkg) if f.startswith("_C") and f.endswith(".so")][0]
        spec = importlib.util.spec_from_file_location("_C", os.path.join(pkg, so))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        sys.modules["ref_simple_knn_C"] = mod
    return sys.modules["ref_simple_knn_C"].distCUDA2


def load_ref_lbs(ref_root=None):
    """The reference's lbs.py + rotations.py, with `smplx.lbs` shimmed by the reference's own
    vendored copy (hugs/datasets/neuman_utils/smpl.py:363-505) because smplx is not installable."""
    d = os.path.join(REF_DIR, "hugs_ref")
    if ref_root is not None:
        paths = dict(lbs=os.path.join(ref_root, "hugs/models/modules/lbs.py"),
                     rotations=os.path.join(ref_root, "hugs/utils/rotations.py"),
                     smpl=os.path.join(ref_root, "hugs/datasets/neuman_utils/smpl.py"))
    else:
        paths = dict(lbs=os.path.join(d, "lbs.py"), rotations=os.path.join(d, "rotations.py"),
                     smpl=os.path.join(d, "smpl_vendored.py"))

    def _load(name, path):
        spec = importlib.util.spec_from_file_location(name, path)
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        return m

    for dep in ("chumpy", "cv2_dummy"):
        pass
    # smpl_vendored imports a few heavy things at module import; provide light stand-ins if missing
    for missing in ("chumpy",):
        if missing not in sys.modules:
            try:
                importlib.import_module(missing)
            except Exception:
                sys.modules[missing] = types.ModuleType(missing)
    smpl = _load("ref_smpl_vendored", paths["smpl"])
    shim = types.ModuleType("smplx.lbs")
    for fn in ("batch_rodrigues", "blend_shapes", "vertices2joints", "batch_rigid_transform"):
        setattr(shim, fn, getattr(smpl, fn))
    pkg = types.ModuleType("smplx")
    pkg.lbs = shim
    sys.modules.setdefault("smplx", pkg)
    sys.modules["smplx.lbs"] = shim
    lbs = _load("ref_lbs", paths["lbs"])
    rot = _load("ref_rotations", paths["rotations"])
    return lbs, rot


def _align(off, a=128):
    return (off + a - 1) // a * a


def decode_ref_buffers(geom, binning, img, P, R, W, H):
    """Slice the reference's three byte buffers by its fromChunk order
    (cuda_rasterizer/rasterizer_impl.cu:155-194; 128-byte aligned slices; torch allocations are
    512-byte aligned so offsets from the tensor base are deterministic)."""
    out = {}

    def take(buf, off, count, dtype, itemsize):
        off = _align(off)
        nbytes = count * itemsize
        out_t = buf[off:off + nbytes].view(dtype)
        return out_t, off + nbytes

    off = 0
    out["depths"], off = take(geom, off, P, torch.float32, 4)
    out["clamped"], off = take(geom, off, 3 * P, torch.uint8, 1)
    out["internal_radii"], off = take(geom, off, P, torch.int32, 4)
    out["means2D"], off = take(geom, off, 2 * P, torch.float32, 4)
    out["cov3D"], off = take(geom, off, 6 * P, torch.float32, 4)
    out["conic_opacity"], off = take(geom, off, 4 * P, torch.float32, 4)
    out["rgb"], off = take(geom, off, 3 * P, torch.float32, 4)
    out["tiles_touched"], off = take(geom, off, P, torch.int32, 4)
    out["means2D"] = out["means2D"].reshape(P, 2)
    out["cov3D"] = out["cov3D"].reshape(P, 6)
    out["conic_opacity"] = out["conic_opacity"].reshape(P, 4)
    out["rgb"] = out["rgb"].reshape(P, 3)
    out["clamped"] = out["clamped"].reshape(P, 3)

    N = W * H
    off = 0
    out["accum_alpha"], off = take(img, off, N, torch.float32, 4)
    out["n_contrib"], off = take(img, off, N, torch.int32, 4)
    ranges, off = take(img, off, 2 * N, torch.int32, 4)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out["ranges"] = ranges.reshape(N, 2)[:tiles]

    if R > 0:
        off = 0
        out["point_list"], off = take(binning, off, R, torch.int32, 4)
        out["point_list_unsorted"], off = take(binning, off, R, torch.int32, 4)
        out["point_list_keys"], off = take(binning, off, R, torch.int64, 8)
        out["point_list_keys_unsorted"], off = take(binning, off, R, torch.int64, 8)
    return out


def decode_our_buffers(geom, binning, img, P, R, W, H):
    """Slice our three byte buffers with the layout the C ABI reports (hugs_b200_*_layout_of)."""
    from hugs_b200 import _native as N

    out = {}
    g, b, im = N.GeomLayout(), N.BinningLayout(), N.ImageLayout()
    N.geom_layout_of(P, g)
    # the binning workspace is laid out for a capacity >= R (recovered from the buffer size)
    N.binning_layout_of(int(N.binning_capacity_of(binning.numel(), W, H)), W, H, b)
    N.image_layout_of(W, H, im)

    def take(buf, off, count, dtype, itemsize):
        return buf[off:off + count * itemsize].view(dtype)

    if P > 0 and geom.numel() > 0:
        out["depths"] = take(geom, g.depths, P, torch.float32, 4)
        out["means2D"] = take(geom, g.means2D, 2 * P, torch.float32, 4).reshape(P, 2)
        out["conic_opacity"] = take(geom, g.conic_opacity, 4 * P, torch.float32, 4).reshape(P, 4)
        out["rgb"] = take(geom, g.rgb, 4 * P, torch.float32, 4).reshape(P, 4)[:, :3]
        out["clamped_bits"] = take(geom, g.clamped, P, torch.uint8, 1)
        out["tiles_touched"] = take(geom, g.tiles_touched, P, torch.int32, 4)
        out["point_offsets"] = take(geom, g.point_offsets, P, torch.int32, 4)
        out["sorted_depths"] = take(geom, g.sorted_depths, P, torch.int32, 4)
        out["perm"] = take(geom, g.perm, P, torch.int32, 4)
    Npix = W * H
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out["accum_alpha"] = take(img, im.accum_alpha, Npix, torch.float32, 4)
    out["n_contrib"] = take(img, im.n_contrib, Npix, torch.int32, 4)
    out["ranges"] = take(img, im.ranges, 2 * tiles, torch.int32, 4).reshape(tiles, 2)
    if R > 0:
        # the two-level binning modes write the sorted 64-bit keys only on demand
        N.check(N.materialize_sorted_keys(P, R, W, H, N.ptr(geom), N.ptr(binning), int(binning.numel()),
                                          ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)),
                "materialize_sorted_keys")
        torch.cuda.synchronize()
        out["point_list_keys"] = take(binning, b.point_list_keys, R, torch.int64, 8)
        out["point_list"] = take(binning, b.point_list, R, torch.int32, 4)
    return out


def rel_err(a, b):
    """Relative error in the sense the acceptance thresholds use: max |a-b| / max |b| (per tensor)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    denom = max(np.abs(b).max(), 1e-30) if b.size else 1.0
    return float(np.abs(a - b).max() / denom) if b.size else 0.0


def robust_rel_err(a, b, q=0.999):
    """Element-wise |a-b| / (|b| + 1e-3 max|b|), q-quantile (ignores the few elements that sit on a
    non-differentiable threshold)."""
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    if b.size == 0:
        return 0.0
    scale = np.abs(b).max() * 1e-3 + 1e-30
    e = np.abs(a - b) / (np.abs(b) + scale)
    return float(np.quantile(e, q))


# ---- gradient acceptance (BASELINE.json: gradients within 1e-3 relative) ---------------------------
# Three views of the same bar; a gradient tensor passes only if all three hold.
#   max-norm      max|a-b| <= tol * max|b|                          (one number per tensor)
#   element-wise  |a-b| <= tol * |b| + floor * max|b|               for a fraction >= q of the elements
#   row-wise      ||a_i - b_i|| <= tol * ||b_i|| + floor * max_j ||b_j||   for a fraction >= q of the rows
# The floor keeps elements that are the difference of large cancelling terms (fp32 summation order:
# both sides accumulate with atomics) from being held to 1e-3 of a value far below the summation
# noise; q < 1 leaves room for the few elements sitting on a non-differentiable threshold
# (alpha == 1/255, T == 1e-4).
GRAD_TOL = 1e-3
GRAD_FLOOR = 1e-5
GRAD_Q = 0.999


def elementwise_violations(a, b, tol=GRAD_TOL, floor=GRAD_FLOOR):
    """Fraction of elements with |a-b| > tol |b| + floor max|b|."""
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    if b.size == 0:
        return 0.0
    bound = tol * np.abs(b) + floor * np.abs(b).max()
    return float(np.mean(np.abs(a - b) > bound))


def rowwise_violations(a, b, tol=GRAD_TOL, floor=GRAD_FLOOR):
    """Fraction of rows (first axis = Gaussian) with ||a_i-b_i||_2 > tol ||b_i||_2 + floor max_j ||b_j||_2."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if b.size == 0:
        return 0.0
    a = a.reshape(a.shape[0], -1)
    b = b.reshape(b.shape[0], -1)
    nb = np.linalg.norm(b, axis=1)
    bound = tol * nb + floor * nb.max()
    return float(np.mean(np.linalg.norm(a - b, axis=1) > bound))


def grad_report(a, b):
    return dict(max_norm=rel_err(a, b), elem_viol=elementwise_violations(a, b), row_viol=rowwise_violations(a, b),
                robust_q999=robust_rel_err(a, b))


def assert_grad_close(a, b, name="", tol=GRAD_TOL, q=GRAD_Q):
    """All three acceptance views (see above); the message carries every figure."""
    rep = grad_report(a, b)
    ok = rep["max_norm"] <= tol and rep["elem_viol"] <= 1.0 - q and rep["row_viol"] <= 1.0 - q
    assert ok, "gradient %s out of tolerance: %s" % (name, rep)
    return rep

