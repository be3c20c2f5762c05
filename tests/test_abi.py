"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/hugs_b200.h declares, and its host-only entry points behave (no GPU compute calls)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "hugs_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hugs_b200_[a-z0-9_]+)\s*\(", src)) - {"hugs_b200_resize_fn"})


def test_header_declares_the_path():
    names = declared_symbols()
    for must in ("hugs_b200_rasterize_forward", "hugs_b200_rasterize_backward", "hugs_b200_mark_visible",
                 "hugs_b200_sort_pairs", "hugs_b200_inclusive_scan", "hugs_b200_lbs_forward", "hugs_b200_lbs_backward"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from hugs_b200 import _native as N

    lib = C.CDLL(N.LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, "declared in include/hugs_b200.h but not exported: %s" % missing
    assert set(N.EXPORTED_SYMBOLS) <= set(declared_symbols())
    assert N.abi_version() == 10


def test_sort_bits_match_reference_rule():
    # 32 + getHigherMsb(tiles) (cuda_rasterizer/rasterizer_impl.cu:35-50,300): 8160 tiles -> 13, 256 -> 9
    from hugs_b200 import _native as N

    assert N.sort_bits(1920, 1080) == 45
    assert N.sort_bits(256, 256) == 41
    assert N.sort_bits(16, 16) == 33
    assert N.sort_bits(200, 120) == 32 + 7  # 13*8 = 104 tiles


def test_layouts_are_aligned_and_disjoint():
    from hugs_b200 import _native as N

    g, b, im = N.GeomLayout(), N.BinningLayout(), N.ImageLayout()
    assert N.geom_layout_of(1000, g) > 0
    offs = [g.depths, g.means2D, g.conic_opacity, g.rgb, g.clamped, g.tiles_touched, g.point_offsets, g.scan_temp]
    assert offs == sorted(offs) and all(o % 256 == 0 for o in offs) and g.total >= offs[-1]
    assert g.means2D - g.depths >= 4000 and g.conic_opacity - g.means2D >= 8000
    assert N.binning_layout_of(12345, 1920, 1080, b) > 0
    assert b.point_list == 0 and b.point_list_keys >= 4 * 12345 and b.total > b.hit_words > b.sort_temp
    # the capacity recovered from the buffer size describes the same layout (slices are 256-B aligned)
    cap = N.binning_capacity_of(b.total, 1920, 1080)
    b2 = N.BinningLayout()
    N.binning_layout_of(cap, 1920, 1080, b2)
    assert 12345 <= cap < 12345 + 64 and bytes(b2) == bytes(b)
    assert N.binning_capacity_of(b.total + 100000, 1920, 1080) > cap
    assert N.binning_capacity_of(0, 1920, 1080) == 0
    assert N.set_binning_mode(7) == -1 and N.set_binning_mode(N.BINNING_ASYNC) == N.BINNING_ASYNC
    assert N.image_layout_of(1920, 1080, im) > 0
    assert im.n_contrib - im.accum_alpha >= 4 * 1920 * 1080 and im.total - im.ranges >= 8 * 8160
    assert N.geom_layout_of(-1, g) < 0
    assert N.image_layout_of(0, 10, im) < 0
    assert N.geom_layout_of(0, g) > 0  # P = 0 is legal


def test_argument_errors_are_reported_without_a_gpu():
    from hugs_b200 import _native as N

    nr = C.c_int(7)
    null_cb = C.cast(None, N.RESIZE_FN)
    rc = N.rasterize_forward(null_cb, None, null_cb, None, null_cb, None, 0, 0, 0, None, 16, 16, None, None, None,
                             None, None, 1.0, None, None, None, None, None, 1.0, 1.0, 0, None, None, 0, None,
                             C.byref(nr))
    assert rc == -1 and b"null" in N.last_error()
    with pytest.raises(RuntimeError, match="rasterize_forward"):
        N.check(rc, "x")
    assert N.sort_pairs(None, None, None, None, -1, 45, None, 0, None) == -1
    assert N.sort_pairs(None, None, None, None, 5, 45, None, 0, None) == -1
    assert N.inclusive_scan(None, None, 3, None, 0, None) == -1
    assert N.rasterize_backward(-1, 0, 0, 0, None, 16, 16, None, None, None, None, 1.0, None, None, None, None, None,
                                1.0, 1.0, None, None, None, 0, None, None, None, None, None, None, None, None, None,
                                None, None, 0, None) == -1
    assert N.mark_visible(-3, None, None, None, None, None) == -1
    assert N.stage_name(0) == b"preprocess" and N.stage_name(99) is None
    assert N.sort_temp_bytes(0) > 0 and N.scan_temp_bytes(10**6) > 8 * (10**6 // 8192)


def test_product_never_touches_the_oracle():
    """The product path must not import, link or execute anything under oracle/."""
    bad = []
    for base, _, files in os.walk(os.path.join(ROOT, "ml-hugs_b200")):
        if os.sep + "build" in base:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", "Makefile")):
                txt = open(os.path.join(base, f), errors="replace").read()
                if re.search(r"(^|\n)\s*(from|import)\s+oracle|liboracle|oracle/|splat_oracle", txt):
                    if f == "_native.py" and "oracle/" in txt and "import oracle" not in txt:
                        continue  # the docstring that states the rule
                    bad.append(os.path.join(base, f))
    assert not bad, bad


def test_ctypes_structures_match_the_header():
    """The ctypes mirrors of the header's structs: sizes of an LP64 C compiler for the declared fields, and the
    argument checks of the view-batched entry points answer without a GPU."""
    import ctypes as C

    from hugs_b200 import _native as N
    assert C.sizeof(N.GradSinks) == 5 * 2 * 8                 # hugs_b200_grad_sinks: five float*[2]
    assert C.sizeof(N.SceneView) == 3 * 8 + 2 * 4 + 2 * 8     # hugs_b200_scene_view: 3 pointers, 2 floats, 2 pointers
    header = open(os.path.join(ROOT, "include", "hugs_b200.h")).read()
    body = header[header.index("typedef struct hugs_b200_scene_view {"):header.index("} hugs_b200_scene_view;")]
    fields = [ln.split(";")[0].split()[-1].lstrip("*") for ln in body.splitlines()[1:] if ";" in ln]
    assert fields == ["viewmatrix", "projmatrix", "cam_pos", "tan_fovy", "radii", "geom_buffer"]   # tan_fovx shares a line
    assert [n for n, _ in N.SceneView._fields_] == ["viewmatrix", "projmatrix", "cam_pos", "tan_fovx", "tan_fovy", "radii",
                                                    "geom_buffer"]
    assert N.scene_backward_views(-1, None, 0, 0, 3, 16, 16, 16, None, None, None, 1.0, None, None, None) == -1
    assert N.preprocess_scene_views(2, None, 0, 8, 3, 16, 16, 16, None, None, None, None, 1.0, None, 0, None) == -1
    assert N.scene_backward_views(0, None, 0, 8, 3, 16, 16, 16, None, None, None, 1.0, None, None, None) == 0
    assert N.FLAG_DEFER_SEG1 == 4 and N.FLAG_SEG1_PREPROCESSED == 8
    assert "#define HUGS_B200_FLAG_DEFER_SEG1 4" in header and "#define HUGS_B200_FLAG_SEG1_PREPROCESSED 8" in header
