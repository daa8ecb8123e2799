"""TEST INFRASTRUCTURE: ctypes front-end to oracle/cpu_rasterizer.c (the CPU restatement of the
reference rasterizer).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module; the product (wildgs_slam_blackwell_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "cpu_rasterizer.c")
LIB = os.path.join(HERE, "_build", "liborc.so")
_lib = None
_fp = C.c_void_p


class Scene(C.Structure):
    _fields_ = [("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
                ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
                ("bg", _fp), ("means3D", _fp), ("sh", _fp), ("colors_precomp", _fp), ("opacities", _fp),
                ("scales", _fp), ("rotations", _fp), ("cov3D_precomp", _fp), ("viewmatrix", _fp),
                ("projmatrix", _fp), ("projmatrix_raw", _fp), ("campos", _fp)]


class Grads(C.Structure):
    _fields_ = [(n, _fp) for n in ("dL_dmeans3D", "dL_dmeans2D", "dL_dsh", "dL_dcolors", "dL_dopacity", "dL_dscales",
                                   "dL_drotations", "dL_dcov3D", "dL_dconic", "dL_ddepths", "dL_dtau")]


def build(force=False, quiet=False):
    """gcc -O2 -ffp-contract=off (the explicit fmaf() sequence must not be re-contracted)."""
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-shared",
           "-fvisibility=hidden", "-o", LIB, SRC, "-lm"]
    if not quiet:
        print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return LIB


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build(quiet=True)
        lib = C.CDLL(LIB)
        lib.orc_forward.restype = C.c_void_p
        lib.orc_forward.argtypes = [C.POINTER(Scene), _fp, _fp, _fp, _fp, _fp, C.c_int]
        lib.orc_num_rendered.restype = C.c_int64
        lib.orc_num_rendered.argtypes = [C.c_void_p]
        lib.orc_export_state.argtypes = [C.c_void_p] + [_fp] * 11
        lib.orc_backward.argtypes = [C.c_void_p, _fp, _fp, C.POINTER(Grads), C.c_int]
        lib.orc_free.argtypes = [C.c_void_p]
        lib.orc_mark_visible.argtypes = [C.c_int, _fp, _fp, _fp]
        _lib = lib
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def forward_backward(g, cam, cot=None, colors=None, cov=None, threads=0, export=False, scale_modifier=1.0):
    """g / cam: dicts as made by wildgs_slam_blackwell_b200.scenes; cot = (dL_dcolor, dL_ddepth) or None
    (forward only).  Returns a dict keyed like tests/golden/*.npz (outputs, st_* state, g_* gradients)."""
    lib = load()
    means3D = _f32(g["means3D"])
    P = means3D.shape[0]
    W, H = int(cam["width"]), int(cam["height"])
    sh = None if colors is not None else _f32(g["sh"])
    M = 0 if sh is None else sh.shape[1]
    keep = dict(bg=_f32(cam["bg"]), means3D=means3D, sh=sh, colors=_f32(colors), opac=_f32(g["opacities"]),
                scales=None if cov is not None else _f32(g["scales"]),
                rots=None if cov is not None else _f32(g["rotations"]), cov=_f32(cov),
                view=_f32(cam["viewmatrix"]), proj=_f32(cam["projmatrix"]), praw=_f32(cam["projmatrix_raw"]),
                campos=_f32(cam["campos"]))
    sc = Scene(P, int(g["sh_degree"]), M, W, H, float(cam["tanfovx"]), float(cam["tanfovy"]), float(scale_modifier),
               _p(keep["bg"]), _p(keep["means3D"]), _p(keep["sh"]), _p(keep["colors"]), _p(keep["opac"]),
               _p(keep["scales"]), _p(keep["rots"]), _p(keep["cov"]), _p(keep["view"]), _p(keep["proj"]),
               _p(keep["praw"]), _p(keep["campos"]))
    out = dict(color=np.zeros((3, H, W), np.float32), depth=np.zeros((1, H, W), np.float32),
               opacity=np.zeros((1, H, W), np.float32), radii=np.zeros(P, np.int32), n_touched=np.zeros(P, np.int32))
    h = lib.orc_forward(C.byref(sc), _p(out["color"]), _p(out["depth"]), _p(out["opacity"]), _p(out["radii"]),
                        _p(out["n_touched"]), int(threads))
    if not h:
        raise RuntimeError("orc_forward: invalid arguments")
    try:
        R = int(lib.orc_num_rendered(h))
        out["num_rendered"] = np.int64(R)
        if export:
            tiles = ((W + 15) // 16) * ((H + 15) // 16)
            st = dict(st_depths=np.zeros(P, np.float32), st_means2D=np.zeros((P, 2), np.float32),
                      st_conic_opacity=np.zeros((P, 4), np.float32), st_rgb=np.zeros((P, 3), np.float32),
                      st_clamped=np.zeros((P, 3), np.uint8), st_cov3D=np.zeros((P, 6), np.float32),
                      st_point_list_keys=np.zeros(R, np.uint64), st_point_list=np.zeros(R, np.uint32),
                      st_ranges=np.zeros((tiles, 2), np.uint32), st_n_contrib=np.zeros(H * W, np.uint32),
                      st_final_T=np.zeros(H * W, np.float32))
            lib.orc_export_state(h, *[_p(st[k]) for k in ("st_depths", "st_means2D", "st_conic_opacity", "st_rgb",
                                                           "st_clamped", "st_cov3D", "st_point_list_keys",
                                                           "st_point_list", "st_ranges", "st_n_contrib", "st_final_T")])
            out.update(st)
        if cot is not None:
            dc, dd = _f32(cot[0]), _f32(cot[1])
            gr = dict(dL_dmeans3D=np.zeros((P, 3), np.float32), dL_dmeans2D=np.zeros((P, 3), np.float32),
                      dL_dsh=np.zeros((P, M, 3), np.float32) if sh is not None else None,
                      dL_dcolors=np.zeros((P, 3), np.float32), dL_dopacity=np.zeros((P, 1), np.float32),
                      dL_dscales=np.zeros((P, 3), np.float32) if cov is None else None,
                      dL_drotations=np.zeros((P, 4), np.float32) if cov is None else None,
                      dL_dcov3D=np.zeros((P, 6), np.float32), dL_dconic=np.zeros((P, 2, 2), np.float32),
                      dL_ddepths=np.zeros((P, 1), np.float32), dL_dtau=np.zeros(6, np.float32))
            gs = Grads(*[_p(gr[n]) for n, _ in Grads._fields_])
            rc = lib.orc_backward(h, _p(dc), _p(dd), C.byref(gs), int(threads))
            if rc != 0:
                raise RuntimeError("orc_backward failed")
            for k, v in gr.items():
                out["g_" + k] = v
            out["dL_dtau"] = gr["dL_dtau"]
            out["g_dL_dtau_sum"] = gr["dL_dtau"]
    finally:
        lib.orc_free(h)
    return out


def mark_visible(means3D, viewmatrix):
    m, v = _f32(means3D), _f32(viewmatrix)
    present = np.zeros(m.shape[0], np.uint8)
    load().orc_mark_visible(m.shape[0], _p(m), _p(v), _p(present))
    return present.astype(bool)
