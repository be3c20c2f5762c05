#!/usr/bin/env python3
"""Build the UNMODIFIED reference rasterizer into oracle/_ref/ (test infrastructure only).

This compiles the reference's own sources *where they lie* under /root/reference
(submodules/diff-gaussian-rasterization: cuda_rasterizer/{rasterizer_impl,forward,backward}.cu,
rasterize_points.cu, ext.cpp) with plain nvcc/g++ commands -- the reference's setup.py / CMake
are not run -- and writes every output under oracle/_ref/ (git-ignored, NOT gpurun-ignored,
so the built extension travels to the GPU box).

The only deviation from a stock build is `-include cstdint`: cuda_rasterizer/rasterizer_impl.h:24,40-61
uses std::uintptr_t / uint32_t / uint64_t without including <cstdint>, which GCC 13 rejects.

It also installs (copies, into the git-ignored oracle/_ref/ only) the three pure-Python reference
modules the oracle harness imports: the rasterizer's diff_gaussian_rasterization/__init__.py and the
LBS path (hugs/models/modules/lbs.py, hugs/utils/rotations.py, and the vendored smplx helpers in
hugs/datasets/neuman_utils/smpl.py that stand in for the un-installable smplx dependency).

Nothing under oracle/ is imported by the product; only tests/, __graft_entry__.smoke() and
bench.py's reference / cpu_baseline legs use it.
"""
import os
import shutil
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

REF = os.environ.get("HUGS_REFERENCE", "/root/reference")
DGR = os.path.join(REF, "submodules", "diff-gaussian-rasterization")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
OBJ = os.path.join(OUT, "obj")
PKG = os.path.join(OUT, "diff_gaussian_rasterization")
HUGS = os.path.join(OUT, "hugs_ref")

CU_SOURCES = [
    "cuda_rasterizer/rasterizer_impl.cu",
    "cuda_rasterizer/forward.cu",
    "cuda_rasterizer/backward.cu",
    "rasterize_points.cu",
]
CPP_SOURCES = ["ext.cpp"]


def run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def build_simple_knn():
    """The reference's simple-knn extension (distCUDA2), same recipe: its own sources where they lie,
    plain nvcc / g++, output under oracle/_ref/simple_knn/.  `-include cfloat`: simple_knn.cu uses
    FLT_MAX without including <cfloat>."""
    from torch.utils import cpp_extension as ce
    knn = os.path.join(REF, "submodules", "simple-knn")
    if not os.path.isdir(knn):
        return
    pkg = os.path.join(OUT, "simple_knn")
    os.makedirs(pkg, exist_ok=True)
    os.makedirs(OBJ, exist_ok=True)
    target = os.path.join(pkg, "_C" + sysconfig.get_config_var("EXT_SUFFIX"))
    srcs = [os.path.join(knn, f) for f in ("simple_knn.cu", "spatial.cu", "ext.cpp")]
    if os.path.exists(target) and os.path.getmtime(target) > max(os.path.getmtime(f) for f in srcs) \
            and "--force" not in sys.argv:
        print(f"{target} is up to date")
        return
    incs = ["-I" + q for q in ce.include_paths()] + ["-I" + sysconfig.get_paths()["include"], "-I" + knn]
    defs = ["-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=1"]
    cuda_home = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    nvcc = os.path.join(cuda_home, "bin", "nvcc")
    objs, jobs = [], []
    for f in ("simple_knn.cu", "spatial.cu"):
        o = os.path.join(OBJ, "knn_" + f + ".o")
        objs.append(o)
        jobs.append([nvcc, "-c", os.path.join(knn, f), "-o", o, "-std=c++17", "-O3", "-gencode",
                     "arch=compute_100,code=sm_100", "--expt-relaxed-constexpr", "-include", "cfloat",
                     "--compiler-options", "-fPIC", "-w"] + incs + defs)
    o = os.path.join(OBJ, "knn_ext.cpp.o")
    objs.append(o)
    jobs.append(["g++", "-c", os.path.join(knn, "ext.cpp"), "-o", o, "-std=c++17", "-O2", "-fPIC", "-w",
                 "-I" + os.path.join(cuda_home, "include")] + incs + defs)
    with ThreadPoolExecutor(max_workers=3) as ex:
        list(ex.map(run, jobs))
    link = ["g++", "-shared", "-o", target] + objs
    for d in ce.library_paths():
        link += ["-L" + d, "-Wl,-rpath," + d]
    link += ["-L" + os.path.join(cuda_home, "lib64"), "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python",
             "-lc10_cuda", "-ltorch_cuda", "-lcudart"]
    run(link)
    with open(os.path.join(pkg, "__init__.py"), "w") as f:
        f.write("")
    print("oracle/_ref ready:", target)


def main():
    if not os.path.isdir(DGR):
        print(f"reference not present at {DGR}; keeping any prebuilt oracle/_ref", file=sys.stderr)
        return 0
    import torch  # noqa: F401  (only for include/library paths)
    from torch.utils import cpp_extension as ce

    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(PKG, exist_ok=True)
    os.makedirs(HUGS, exist_ok=True)
    ext_suffix = sysconfig.get_config_var("EXT_SUFFIX")
    target = os.path.join(PKG, "_C" + ext_suffix)

    srcs = [os.path.join(DGR, s) for s in CU_SOURCES + CPP_SOURCES]
    newest_src = max(os.path.getmtime(s) for s in srcs)
    if os.path.exists(target) and os.path.getmtime(target) > newest_src and "--force" not in sys.argv:
        print(f"{target} is up to date")
    else:
        incs = ["-I" + p for p in ce.include_paths()] + [
            "-I" + sysconfig.get_paths()["include"],
            "-I" + os.path.join(DGR, "third_party", "glm"),
            "-I" + DGR,
        ]
        defs = [
            "-DTORCH_EXTENSION_NAME=_C",
            "-DTORCH_API_INCLUDE_EXTENSION_H",
            "-D_GLIBCXX_USE_CXX11_ABI=1",
        ]
        nvcc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "bin", "nvcc")
        jobs = []
        objs = []
        for s in CU_SOURCES:
            o = os.path.join(OBJ, os.path.basename(s) + ".o")
            objs.append(o)
            jobs.append([
                nvcc, "-c", os.path.join(DGR, s), "-o", o, "-std=c++17", "-O3",
                "-gencode", "arch=compute_100,code=sm_100",
                "--expt-relaxed-constexpr", "-include", "cstdint",
                "--compiler-options", "-fPIC", "-w",
            ] + incs + defs)
        for s in CPP_SOURCES:
            o = os.path.join(OBJ, os.path.basename(s) + ".o")
            objs.append(o)
            jobs.append(["g++", "-c", os.path.join(DGR, s), "-o", o, "-std=c++17", "-O2", "-fPIC",
                         "-include", "cstdint", "-w",
                         "-I" + os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")]
                        + incs + defs)
        with ThreadPoolExecutor(max_workers=5) as ex:
            list(ex.map(run, jobs))
        libdirs = ce.library_paths()
        link = ["g++", "-shared", "-o", target] + objs
        for d in libdirs:
            link += ["-L" + d, "-Wl,-rpath," + d]
        cuda_lib = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "lib64")
        link += ["-L" + cuda_lib, "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python",
                 "-lc10_cuda", "-ltorch_cuda", "-lcudart"]
        run(link)

    # pure-Python reference modules (install step; git-ignored destination)
    shutil.copyfile(os.path.join(DGR, "diff_gaussian_rasterization", "__init__.py"),
                    os.path.join(PKG, "__init__.py"))
    for rel, name in [
        ("hugs/models/modules/lbs.py", "lbs.py"),
        ("hugs/utils/rotations.py", "rotations.py"),
        ("hugs/datasets/neuman_utils/smpl.py", "smpl_vendored.py"),
        ("hugs/utils/graphics.py", "graphics.py"),
        ("hugs/utils/spherical_harmonics.py", "spherical_harmonics.py"),
    ]:
        shutil.copyfile(os.path.join(REF, rel), os.path.join(HUGS, name))
    print("oracle/_ref ready:", target)
    try:
        build_simple_knn()
    except Exception as e:  # noqa: BLE001  (the rasterizer build above is what the parity suite needs)
        print("simple-knn reference build failed (distCUDA2 parity stays unpinned):", e, file=sys.stderr)
    return 0


if __name__ == "__main__":
    sys.exit(main())
