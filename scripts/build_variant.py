#!/usr/bin/env python
"""Kernel A/B runs: build a copy of libwgs_rasterizer.so in which some csrc/*.cu files are replaced by variants.

    python scripts/build_variant.py NAME render_fwd.cu=scripts/variants/render_fwd_c.cu [more=...]
      -> scripts/variants/libwgs_NAME.so      (run anything with WGS_RASTERIZER_LIB=<that path>)

The other objects are the in-tree build's (python -m wildgs_slam_blackwell_b200.build first)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wildgs_slam_blackwell_b200 import build as B  # noqa: E402


def main():
    name, repl = sys.argv[1], dict(a.split("=") for a in sys.argv[2:])
    B.build(quiet=True)
    out_dir = os.path.join(ROOT, "scripts", "variants")
    os.makedirs(out_dir, exist_ok=True)
    objs = []
    for src in B.SOURCES:
        if src in repl:
            o = os.path.join(out_dir, f"{name}_{src.replace('.cu', '.o')}")
            cmd = ["nvcc", "-c", os.path.join(ROOT, repl[src]), "-o", o, "-I", B.CSRC, "-I", os.path.join(ROOT, "include"),
                   *B.ARCH, *B.FLAGS]
            subprocess.check_call(cmd)
        else:
            o = os.path.join(B.OBJ, src.replace(".cu", ".o"))
        objs.append(o)
    lib = os.path.join(out_dir, f"libwgs_{name}.so")
    subprocess.check_call(["nvcc", "-shared", "-o", lib, *objs, *B.ARCH, "-Xcompiler", "-fPIC", "-lcudart"])
    print(lib)


if __name__ == "__main__":
    main()
