"""ctypes binding of libhugs_b200.so (C ABI declared in include/hugs_b200.h).

This is the only place the shared library is loaded.  There is NO fallback: if the CUDA library
is missing the import fails loudly, and nothing here (or anywhere in the product) imports
``oracle/``.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_CANDIDATES = [
    os.environ.get("HUGS_B200_LIB", ""),
    os.path.join(_HERE, "..", "lib", "libhugs_b200.so"),
]


class NativeLibraryMissing(ImportError):
    pass


def _load():
    for p in _LIB_CANDIDATES:
        if p and os.path.exists(p):
            return C.CDLL(os.path.abspath(p)), os.path.abspath(p)
    raise NativeLibraryMissing(
        "libhugs_b200.so not found (looked in %s). Build it with `make -C ml-hugs_b200/csrc` or "
        "`python -c 'import __graft_entry__ as g; g.build()'`. There is no CPU fallback."
        % [p for p in _LIB_CANDIDATES if p]
    )


lib, LIB_PATH = _load()

RESIZE_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)
fp = C.c_void_p  # device pointers are passed as integers / None
i32 = C.c_int


class GeomLayout(C.Structure):
    _fields_ = [(n, C.c_size_t) for n in (
        "depths", "means2D", "conic_opacity", "rgb", "clamped", "tiles_touched", "point_offsets",
        "scan_temp", "sorted_depths", "perm", "total")]


class ImageLayout(C.Structure):
    _fields_ = [(n, C.c_size_t) for n in ("accum_alpha", "n_contrib", "ranges", "total")]


class BinningLayout(C.Structure):
    _fields_ = [(n, C.c_size_t) for n in (
        "point_list_keys", "point_list", "point_list_keys_unsorted", "point_list_unsorted", "sort_temp",
        "hit_words", "total")]


def _sig(name, restype, argtypes):
    f = getattr(lib, name)
    f.restype = restype
    f.argtypes = argtypes
    return f


abi_version = _sig("hugs_b200_abi_version", i32, [])
last_error = _sig("hugs_b200_last_error", C.c_char_p, [])
launch_count = _sig("hugs_b200_launch_count", C.c_longlong, [i32])

rasterize_forward = _sig("hugs_b200_rasterize_forward", i32, [
    RESIZE_FN, C.c_void_p, RESIZE_FN, C.c_void_p, RESIZE_FN, C.c_void_p,
    i32, i32, i32,            # P, D, M
    fp, i32, i32,             # background, width, height
    fp, fp, fp, fp,           # means3D, shs, colors_precomp, opacities
    fp, C.c_float, fp, fp,    # scales, scale_modifier, rotations, cov3D_precomp
    fp, fp, fp,               # viewmatrix, projmatrix, cam_pos
    C.c_float, C.c_float, i32,  # tan_fovx, tan_fovy, prefiltered
    fp, fp, i32, fp,          # out_color, radii, debug, stream
    C.POINTER(i32),
])

rasterize_backward = _sig("hugs_b200_rasterize_backward", i32, [
    i32, i32, i32, i32,       # P, D, M, R
    fp, i32, i32,             # background, width, height
    fp, fp, fp,               # means3D, shs, colors_precomp
    fp, C.c_float, fp, fp,    # scales, scale_modifier, rotations, cov3D_precomp
    fp, fp, fp,               # viewmatrix, projmatrix, cam_pos
    C.c_float, C.c_float,     # tan_fovx, tan_fovy
    fp, fp, fp, C.c_size_t, fp,  # radii, geom, binning, binning_bytes, image
    fp,                       # dL_dpix
    fp, fp, fp, fp, fp, fp, fp, fp, fp,  # 9 gradient outputs
    i32, fp,                  # debug, stream
])

rasterize_backward_acc = _sig("hugs_b200_rasterize_backward_acc", i32, [
    i32, i32, i32, i32,       # P, D, M, R
    fp, i32, i32,             # background, width, height
    fp, fp, fp,               # means3D, shs, colors_precomp
    fp, C.c_float, fp, fp,    # scales, scale_modifier, rotations, cov3D_precomp
    fp, fp, fp,               # viewmatrix, projmatrix, cam_pos
    C.c_float, C.c_float,     # tan_fovx, tan_fovy
    fp, fp, fp, C.c_size_t, fp,  # radii, geom, binning, binning_bytes, image
    fp,                       # dL_dpix
    fp, fp, fp, fp, fp, fp, fp, fp, fp,  # 9 gradient outputs
    i32, fp,                  # debug, stream
    i32,                      # dL_dsh_accumulate
])

rasterize_forward_seg2 = _sig("hugs_b200_rasterize_forward_seg2", i32, [
    RESIZE_FN, C.c_void_p, RESIZE_FN, C.c_void_p, RESIZE_FN, C.c_void_p,
    i32, i32, fp, i32, i32,   # D, M, background, width, height
    i32, fp, fp, fp, fp, fp,  # P0, means3D0, shs0, opacities0, scales0, rotations0
    i32, fp, fp, fp, fp, fp,  # P1, means3D1, shs1, opacities1, scales1, rotations1
    C.c_float, fp, fp, fp,    # scale_modifier, viewmatrix, projmatrix, cam_pos
    C.c_float, C.c_float, i32,  # tan_fovx, tan_fovy, prefiltered
    fp, fp, i32, fp,          # out_color, radii, debug, stream
    C.POINTER(i32),
])

rasterize_backward_seg2 = _sig("hugs_b200_rasterize_backward_seg2", i32, [
    i32, i32, i32, fp, i32, i32,   # D, M, R, background, width, height
    i32, fp, fp, fp, fp,      # P0, means3D0, shs0, scales0, rotations0
    i32, fp, fp, fp, fp,      # P1, means3D1, shs1, scales1, rotations1
    C.c_float, fp, fp, fp,    # scale_modifier, viewmatrix, projmatrix, cam_pos
    C.c_float, C.c_float,     # tan_fovx, tan_fovy
    fp, fp, fp, C.c_size_t, fp,  # radii, geom, binning, binning_bytes, image
    fp,                       # dL_dpix
    fp, fp, fp, fp, fp, fp, fp, fp, fp,  # 9 gradient outputs
    i32, fp,                  # debug, stream
])

layer_bytes = _sig("hugs_b200_layer_bytes", C.c_size_t, [i32, C.c_size_t, i32, i32])
layer_layout_of = _sig("hugs_b200_layer_layout_of", i32, [i32, C.c_size_t, i32, i32, C.POINTER(C.c_size_t)])
layer_forward = _sig("hugs_b200_layer_forward", i32, [
    i32, i32, fp, i32, i32,   # P, n_layer, background, width, height
    fp, fp, fp, C.c_size_t, fp,  # colors_precomp, geom, binning, binning_bytes, image
    fp, C.c_size_t, fp, i32, fp,  # layer buffer, layer bytes, out_color, debug, stream
])
_LAYER_TAIL = [i32, fp, fp, C.c_size_t, fp]   # n_layer, background_layer, layer buffer, layer bytes, dL_dpix_layer
rasterize_backward_layer = _sig("hugs_b200_rasterize_backward_layer", i32,
                                list(rasterize_backward_acc.argtypes) + _LAYER_TAIL)
rasterize_backward_seg2_layer = _sig("hugs_b200_rasterize_backward_seg2_layer", i32,
                                     list(rasterize_backward_seg2.argtypes) + _LAYER_TAIL)



class GradSinks(C.Structure):
    """hugs_b200_grad_sinks: per segment, the accumulation buffer a gradient is ADDED to (NULL: none)."""
    _fields_ = [(n, C.c_void_p * 2) for n in ("means3D", "sh", "opacity", "scales", "rotations")]


rasterize_backward_seg2_ex = _sig("hugs_b200_rasterize_backward_seg2_ex", i32,
                                  list(rasterize_backward_seg2.argtypes) + _LAYER_TAIL + [C.POINTER(GradSinks)])


class SceneView(C.Structure):
    """hugs_b200_scene_view: one view of hugs_b200_scene_backward_views."""
    _fields_ = [("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("cam_pos", C.c_void_p),
                ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("radii", C.c_void_p),
                ("geom_buffer", C.c_void_p)]


preprocess_scene_views = _sig("hugs_b200_preprocess_scene_views", i32,
                              [i32, C.POINTER(SceneView), i32, i32, i32, i32, i32, i32, fp, fp, fp, fp, C.c_float, fp,
                               i32, fp])
scene_backward_views = _sig("hugs_b200_scene_backward_views", i32,
                            [i32, C.POINTER(SceneView), i32, i32, i32, i32, i32, i32, fp, fp, fp, C.c_float, fp,
                             C.POINTER(GradSinks), fp])



class MlpOut(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("ld", i32), ("col0", i32), ("ncols", i32), ("act", i32)]


class MlpLayer(C.Structure):
    _fields_ = [("K", i32), ("N", i32), ("act", i32), ("packed", C.c_void_p), ("z_out", C.c_void_p)]


class MlpChain(C.Structure):
    _fields_ = [("n_layers", i32), ("n_outs", i32), ("layer", MlpLayer * 4), ("out", MlpOut * 3)]


class MlpBwdOut(C.Structure):
    _fields_ = [("dptr", C.c_void_p), ("ld", i32), ("col0", i32), ("ncols", i32), ("act", i32)]


class MlpBwdLayer(C.Structure):
    _fields_ = [("K", i32), ("N", i32), ("packed_t", C.c_void_p), ("z_prev", C.c_void_p), ("dWt", C.c_void_p),
                ("db", C.c_void_p)]


class MlpBwdChain(C.Structure):
    _fields_ = [("n_layers", i32), ("n_outs", i32), ("layer", MlpBwdLayer * 4), ("out", MlpBwdOut * 3),
                ("z_head", C.c_void_p), ("dx", C.c_void_p)]


MLP_ACT_NONE, MLP_ACT_GELU, MLP_ACT_SIGMOID = 0, 1, 2
mlp_packed_floats = _sig("hugs_b200_mlp_packed_floats", C.c_size_t, [i32, i32])
mlp_pack = _sig("hugs_b200_mlp_pack", i32, [fp, fp, i32, i32, i32, fp, fp])
mlp_pack_t = _sig("hugs_b200_mlp_pack_t", i32, [fp, i32, i32, i32, fp, fp])
mlp_chains_backward = _sig("hugs_b200_mlp_chains_backward", i32, [i32, fp, i32, i32, i32, C.POINTER(MlpBwdChain), fp])
mlp_chains_forward = _sig("hugs_b200_mlp_chains_forward", i32, [i32, fp, i32, i32, i32, C.POINTER(MlpChain), fp])

l1_ssim_forward = _sig("hugs_b200_l1_ssim_forward", i32, [i32, i32, i32, fp, fp, fp, fp, fp])
l1_ssim_backward = _sig("hugs_b200_l1_ssim_backward", i32, [i32, i32, i32, fp, fp, fp, fp, fp, fp])

knn_temp_bytes = _sig("hugs_b200_knn_temp_bytes", C.c_size_t, [i32])
knn_mean_dist2 = _sig("hugs_b200_knn_mean_dist2", i32, [i32, fp, fp, fp, C.c_size_t, fp])

densification_stats = _sig("hugs_b200_densification_stats", i32, [i32, fp, fp, i32, fp, fp, fp, fp])

f32 = C.c_float
triplane_forward = _sig("hugs_b200_triplane_forward", i32,
                        [i32, i32, i32, i32, i32, fp, fp, fp, fp, f32, f32, f32, fp, fp, fp])
triplane_backward = _sig("hugs_b200_triplane_backward", i32,
                         [i32, i32, i32, i32, i32, fp, fp, fp, fp, f32, f32, fp, fp, fp, fp, fp, fp])

mark_visible = _sig("hugs_b200_mark_visible", i32, [i32, fp, fp, fp, fp, fp])

geom_layout_of = _sig("hugs_b200_geom_layout_of", i32, [i32, C.POINTER(GeomLayout)])
image_layout_of = _sig("hugs_b200_image_layout_of", i32, [i32, i32, C.POINTER(ImageLayout)])
binning_layout_of = _sig("hugs_b200_binning_layout_of", i32, [i32, i32, i32, C.POINTER(BinningLayout)])

materialize_sorted_keys = _sig("hugs_b200_materialize_sorted_keys", i32,
                               [i32, i32, i32, i32, fp, fp, C.c_size_t, fp])
binning_capacity_of = _sig("hugs_b200_binning_capacity_of", C.c_size_t, [C.c_size_t, i32, i32])
BINNING_ASYNC, BINNING_SYNC, BINNING_REFERENCE, BINNING_TILESORT = 0, 1, 2, 3
FLAG_DEBUG, FLAG_CLAMP01, FLAG_DEFER_SEG1, FLAG_SEG1_PREPROCESSED = 1, 2, 4, 8   # the flag word passed in the entry points' `debug` argument
set_binning_mode = _sig("hugs_b200_set_binning_mode", i32, [i32])
set_blend_priority = _sig("hugs_b200_set_blend_priority", i32, [i32, i32])
_last_pair_counts = _sig("hugs_b200_last_pair_counts", i32, [C.POINTER(i32), C.POINTER(i32)])


def last_pair_counts():
    """(pairs, level-1 entries) of this thread's last uncaptured forward call."""
    r, r1 = i32(0), i32(0)
    check(_last_pair_counts(C.byref(r), C.byref(r1)), "last_pair_counts")
    return int(r.value), int(r1.value)

sort_bits = _sig("hugs_b200_sort_bits", i32, [i32, i32])
sort_temp_bytes = _sig("hugs_b200_sort_temp_bytes", C.c_size_t, [i32])
scan_temp_bytes = _sig("hugs_b200_scan_temp_bytes", C.c_size_t, [i32])
sort_pairs = _sig("hugs_b200_sort_pairs", i32, [fp, fp, fp, fp, i32, i32, fp, C.c_size_t, fp])
sort_pairs_range = _sig("hugs_b200_sort_pairs_range", i32, [fp, fp, fp, fp, i32, i32, i32, fp, C.c_size_t, fp])
inclusive_scan = _sig("hugs_b200_inclusive_scan", i32, [fp, fp, i32, fp, C.c_size_t, fp])
inclusive_scan_static = _sig("hugs_b200_inclusive_scan_static", i32, [fp, fp, i32, fp])
inclusive_scan_static_max_items = _sig("hugs_b200_inclusive_scan_static_max_items", i32, [])

lbs_forward = _sig("hugs_b200_lbs_forward", i32, [
    i32, i32, i32, fp, fp, fp, fp, fp, fp, C.c_float, fp, fp, fp, fp, fp])
lbs_backward = _sig("hugs_b200_lbs_backward", i32, [
    i32, i32, i32, fp, fp, fp, fp, fp, fp, C.c_float, fp, fp, fp, fp, fp, fp, fp, fp, fp, fp, fp])

NUM_STAGES = 11
stage_timing_enable = _sig("hugs_b200_stage_timing_enable", i32, [i32])
stage_timing_read = _sig("hugs_b200_stage_timing_read", i32, [C.POINTER(C.c_double), C.POINTER(C.c_longlong), i32])
stage_name = _sig("hugs_b200_stage_name", C.c_char_p, [i32])


def read_stage_times(reset=True):
    """{stage name: (total ms, calls)} accumulated while stage timing was enabled."""
    ms = (C.c_double * NUM_STAGES)()
    calls = (C.c_longlong * NUM_STAGES)()
    stage_timing_read(ms, calls, int(bool(reset)))
    return {stage_name(i).decode(): (float(ms[i]), int(calls[i])) for i in range(NUM_STAGES)}


EXPORTED_SYMBOLS = [
    "hugs_b200_binning_capacity_of", "hugs_b200_set_binning_mode", "hugs_b200_set_blend_priority", "hugs_b200_last_pair_counts", "hugs_b200_materialize_sorted_keys",
    "hugs_b200_stage_timing_enable", "hugs_b200_stage_timing_read", "hugs_b200_stage_name",
    "hugs_b200_abi_version", "hugs_b200_last_error", "hugs_b200_launch_count",
    "hugs_b200_rasterize_forward", "hugs_b200_rasterize_backward", "hugs_b200_rasterize_backward_acc",
    "hugs_b200_mark_visible",
    "hugs_b200_geom_layout_of", "hugs_b200_image_layout_of", "hugs_b200_binning_layout_of",
    "hugs_b200_sort_bits", "hugs_b200_sort_temp_bytes", "hugs_b200_sort_pairs", "hugs_b200_sort_pairs_range",
    "hugs_b200_scan_temp_bytes", "hugs_b200_inclusive_scan",
    "hugs_b200_inclusive_scan_static", "hugs_b200_inclusive_scan_static_max_items",
    "hugs_b200_scene_backward_views", "hugs_b200_preprocess_scene_views",
    "hugs_b200_lbs_forward", "hugs_b200_lbs_backward",
    "hugs_b200_l1_ssim_forward", "hugs_b200_l1_ssim_backward",
    "hugs_b200_knn_temp_bytes", "hugs_b200_knn_mean_dist2",
    "hugs_b200_densification_stats",
    "hugs_b200_triplane_forward", "hugs_b200_triplane_backward",
    "hugs_b200_rasterize_forward_seg2", "hugs_b200_rasterize_backward_seg2",
    "hugs_b200_layer_bytes", "hugs_b200_layer_layout_of", "hugs_b200_layer_forward", "hugs_b200_rasterize_backward_layer",
    "hugs_b200_rasterize_backward_seg2_layer", "hugs_b200_rasterize_backward_seg2_ex",
    "hugs_b200_mlp_packed_floats", "hugs_b200_mlp_pack", "hugs_b200_mlp_chains_forward",
    "hugs_b200_mlp_pack_t", "hugs_b200_mlp_chains_backward",
]


def check(rc, what):
    """Raise RuntimeError(last_error) for a non-zero return code (reference: C++ exceptions
    surfacing as Python RuntimeError, cuda_rasterizer/auxiliary.h:166-173)."""
    if rc != 0:
        msg = last_error().decode("utf-8", "replace")
        raise RuntimeError("%s failed (code %d): %s" % (what, rc, msg))


def ptr(t):
    """Device pointer of a tensor, or None for an absent optional (numel()==0), mirroring the
    reference's convention of passing torch.Tensor([]) for unused inputs
    (diff_gaussian_rasterization/__init__.py:197-207)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()
