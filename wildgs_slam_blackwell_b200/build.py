"""Build libwgs_rasterizer.so (hand-written sm_100a CUDA + C ABI) in-tree with nvcc.

    python -m wildgs_slam_blackwell_b200.build [--force] [--verbose-ptxas]

nvcc cross-compiles without a GPU.  No fast-math: the forward path must round exactly like the
reference (see csrc/common.cuh).  -lineinfo keeps ncu's source page usable.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_build")
LIB = os.path.join(HERE, "libwgs_rasterizer.so")
SOURCES = ["api.cu", "preprocess_fwd.cu", "binning.cu", "render_fwd.cu", "render_bwd.cu", "preprocess_bwd.cu",
           "loss.cu", "camera.cu", "peer_allreduce.cu", "knn.cu", "adam.cu", "pose.cu", "compact.cu", "ssim.cu", "uncertainty.cu"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr"]


def _deps():
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    inc = os.path.join(os.path.dirname(HERE), "include")
    hdrs += [os.path.join(inc, f) for f in os.listdir(inc) if f.endswith(".h")]
    return hdrs


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose_ptxas=False, quiet=False):
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _deps()
    jobs = []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        if force or _stale(o, [s] + hdrs):
            cmd = ["nvcc", "-c", s, "-o", o, *ARCH, *FLAGS]
            if verbose_ptxas:
                cmd += ["-Xptxas", "-v"]
            jobs.append(cmd)

    def run(cmd):
        if not quiet:
            print("+", " ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0 or verbose_ptxas:
            sys.stderr.write(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))

    with ThreadPoolExecutor(max_workers=6) as ex:
        list(ex.map(run, jobs))
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in SOURCES]
    if force or jobs or _stale(LIB, objs):
        run(["nvcc", "-shared", "-o", LIB, *objs, *ARCH, "-Xcompiler", "-fPIC", "-lcudart"])
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose_ptxas="--verbose-ptxas" in sys.argv)
    print(LIB)
