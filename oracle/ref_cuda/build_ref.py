#!/usr/bin/env python
"""TEST INFRASTRUCTURE: build the UNMODIFIED reference rasterizer for sm_100 into oracle/_ref/.

The reference (`thirdparty/diff-gaussian-rasterization-w-pose`, setup.py:30-38) ships gencode
flags for sm_60..sm_90 + sm_120 only, so as shipped it has no image for B200.  This recipe
compiles the reference sources *where they lie* under /root/reference (nothing is copied into
the repo history) with `-gencode arch=compute_100,code=sm_100 -include cstdint` (gcc 13 needs
<cstdint> for rasterizer_impl.h:22-61), default -fmad, no fast-math — i.e. the reference's
own flags apart from the arch.  Outputs (git-ignored, shipped to the GPU box by gpurun):

  oracle/_ref/libref_rasterizer.so                 reference kernels + oracle/ref_cuda/ref_shim.cu
                                                   (raw-pointer C ABI, state export)
  oracle/_ref/diff_gaussian_rasterization/         "installed" reference package: its own
      __init__.py + _C.so                          __init__.py + pybind module built from
                                                   rasterize_points.cu / ext.cpp (needs torch
                                                   headers, ~4 min; --no-torch-ext skips it)

Usage: python oracle/ref_cuda/build_ref.py [--no-torch-ext] [--force]
"""
import argparse
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("WGS_REFERENCE_ROOT", "/root/reference")
RAST = os.path.join(REF, "thirdparty", "diff-gaussian-rasterization-w-pose")
OUT = os.path.join(ROOT, "oracle", "_ref")
ARCH = ["-gencode", "arch=compute_100,code=sm_100"]


def run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def build_shim(force=False):
    os.makedirs(os.path.join(OUT, "obj"), exist_ok=True)
    cr = os.path.join(RAST, "cuda_rasterizer")
    srcs = [os.path.join(cr, f) for f in ("forward.cu", "backward.cu", "rasterizer_impl.cu")]
    srcs.append(os.path.join(HERE, "ref_shim.cu"))
    target = os.path.join(OUT, "libref_rasterizer.so")
    if not force and newer(target, srcs):
        print("up to date:", target)
        return target
    objs = []
    for s in srcs:
        o = os.path.join(OUT, "obj", os.path.basename(s).replace(".cu", ".o"))
        run(["nvcc", "-c", s, "-o", o, *ARCH, "-include", "cstdint", "-std=c++17",
             "-I", os.path.join(RAST, "third_party", "glm"),
             '-DREF_RASTERIZER_H="%s"' % os.path.join(cr, "rasterizer.h"),
             '-DREF_RASTERIZER_IMPL_H="%s"' % os.path.join(cr, "rasterizer_impl.h"),
             "-Xcompiler", "-fPIC", "-lineinfo"])
        objs.append(o)
    run(["nvcc", "-shared", "-o", target, *objs, *ARCH, "-Xcompiler", "-fPIC", "-lcudart"])
    return target


def build_torch_ext(force=False):
    pkg = os.path.join(OUT, "diff_gaussian_rasterization")
    os.makedirs(pkg, exist_ok=True)
    target = os.path.join(pkg, "_C.so")
    cr = os.path.join(RAST, "cuda_rasterizer")
    srcs = [os.path.join(cr, "rasterizer_impl.cu"), os.path.join(cr, "forward.cu"),
            os.path.join(cr, "backward.cu"), os.path.join(RAST, "rasterize_points.cu"),
            os.path.join(RAST, "ext.cpp")]
    if not force and newer(target, srcs):
        print("up to date:", target)
        return target
    from torch.utils.cpp_extension import load
    bdir = os.path.join(OUT, "torch_build")
    os.makedirs(bdir, exist_ok=True)
    load(name="_C", sources=srcs, build_directory=bdir, is_python_module=False, verbose=True,
         extra_include_paths=[RAST, os.path.join(RAST, "third_party", "glm")],
         extra_cflags=["-include", "cstdint"],
         extra_cuda_cflags=["-gencode=arch=compute_100,code=sm_100", "-include", "cstdint"])
    shutil.copy2(os.path.join(bdir, "_C.so"), target)
    # "install" the reference's Python front-end next to it (like pip install --target would)
    shutil.copy2(os.path.join(RAST, "diff_gaussian_rasterization", "__init__.py"),
                 os.path.join(pkg, "__init__.py"))
    return target


def build_simple_knn(force=False):
    """The unmodified reference simple-knn (thirdparty/simple-knn/simple_knn.cu) + ref_knn_shim.cu
    -> oracle/_ref/libref_simple_knn.so (raw-pointer C entry ref_dist_cuda2)."""
    knn = os.path.join(REF, "thirdparty", "simple-knn")
    srcs = [os.path.join(knn, "simple_knn.cu"), os.path.join(HERE, "ref_knn_shim.cu")]
    target = os.path.join(OUT, "libref_simple_knn.so")
    if not force and newer(target, srcs):
        print("up to date:", target)
        return target
    os.makedirs(OUT, exist_ok=True)
    run(["nvcc", "-shared", "-o", target, *srcs, *ARCH, "-include", "cstdint", "-include", "cfloat", "-std=c++17",
         '-DREF_SIMPLE_KNN_H="%s"' % os.path.join(knn, "simple_knn.h"), "-Xcompiler", "-fPIC", "-lcudart", "-w"])
    return target


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--no-torch-ext", action="store_true")
    ap.add_argument("--force", action="store_true")
    a = ap.parse_args()
    if not os.path.isdir(RAST):
        print("reference tree not present at", RAST, "- nothing to do")
        return 0
    build_shim(a.force)
    build_simple_knn(a.force)
    if not a.no_torch_ext:
        build_torch_ext(a.force)
    return 0


if __name__ == "__main__":
    sys.exit(main())
