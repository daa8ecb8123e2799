"""TEST INFRASTRUCTURE: ctypes front-end to oracle/_ref/libref_rasterizer.so, the UNMODIFIED
reference rasterizer compiled for sm_100 by build_ref.py (+ ref_shim.cu's extern "C" entry
points).  Only tests/, __graft_entry__.smoke(), tests/golden/make_golden.py and the
reference arm of bench.py may import this.  Needs a GPU.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "_ref", "libref_rasterizer.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB_PATH)
        _lib.ref_create.restype = C.c_void_p
        _lib.ref_destroy.argtypes = [C.c_void_p]
        _lib.ref_forward.restype = C.c_int
        _lib.ref_backward.restype = C.c_int
        _lib.ref_export_state.restype = C.c_int
    return _lib


def _p(t):
    if t is None or t.numel() == 0:
        return C.c_void_p(0)
    assert t.is_cuda and t.is_contiguous()
    return C.c_void_p(t.data_ptr())


class RefRasterizer:
    """One reference context (owns the reference's geom/binning/image chunks)."""

    def __init__(self):
        self.h = C.c_void_p(lib().ref_create())

    def __del__(self):
        try:
            lib().ref_destroy(self.h)
        except Exception:
            pass

    def forward(self, g, cam, colors_precomp=None, cov3D_precomp=None, scale_modifier=1.0,
                debug=False):
        """g: dict of cuda tensors (means3D, sh, opacities, scales, rotations, sh_degree);
        cam: dict of cuda tensors/scalars.  Returns dict of outputs."""
        dev = g["means3D"].device
        P = g["means3D"].shape[0]
        W, H = cam["width"], cam["height"]
        sh = None if colors_precomp is not None else g["sh"]
        M = 0 if sh is None else sh.shape[1]
        out = dict(color=torch.zeros(3, H, W, device=dev), depth=torch.zeros(1, H, W, device=dev),
                   opacity=torch.zeros(1, H, W, device=dev),
                   radii=torch.zeros(P, dtype=torch.int32, device=dev),
                   n_touched=torch.zeros(P, dtype=torch.int32, device=dev))
        scales = None if cov3D_precomp is not None else g["scales"]
        rots = None if cov3D_precomp is not None else g["rotations"]
        torch.cuda.synchronize()
        R = lib().ref_forward(
            self.h, C.c_int(P), C.c_int(g["sh_degree"]), C.c_int(M), _p(cam["bg"]), C.c_int(W),
            C.c_int(H), _p(g["means3D"]), _p(sh), _p(colors_precomp), _p(g["opacities"]),
            _p(scales), C.c_float(scale_modifier), _p(rots), _p(cov3D_precomp),
            _p(cam["viewmatrix"]), _p(cam["projmatrix"]), _p(cam["campos"]),
            C.c_float(cam["tanfovx"]), C.c_float(cam["tanfovy"]), C.c_int(0), _p(out["color"]),
            _p(out["depth"]), _p(out["opacity"]), _p(out["radii"]), _p(out["n_touched"]),
            C.c_int(int(debug)))
        torch.cuda.synchronize()
        if R < 0:
            raise RuntimeError("ref_forward failed")
        out["num_rendered"] = R
        self._last = (g, cam, colors_precomp, cov3D_precomp, scale_modifier, out, M)
        return out

    def backward(self, dL_dcolor, dL_ddepth, debug=False):
        g, cam, colors_precomp, cov3D_precomp, scale_modifier, out, M = self._last
        dev = g["means3D"].device
        P = g["means3D"].shape[0]
        W, H = cam["width"], cam["height"]
        z = lambda *s: torch.zeros(*s, device=dev)
        gr = dict(dL_dmeans2D=z(P, 3), dL_dconic=z(P, 2, 2), dL_dopacity=z(P, 1),
                  dL_dcolors=z(P, 3), dL_ddepths=z(P, 1), dL_dmeans3D=z(P, 3), dL_dcov3D=z(P, 6),
                  dL_dsh=z(P, max(M, 0), 3), dL_dscales=z(P, 3), dL_drotations=z(P, 4),
                  dL_dtau=z(P, 6))
        sh = None if colors_precomp is not None else g["sh"]
        scales = None if cov3D_precomp is not None else g["scales"]
        rots = None if cov3D_precomp is not None else g["rotations"]
        torch.cuda.synchronize()
        rc = lib().ref_backward(
            self.h, C.c_int(P), C.c_int(g["sh_degree"]), C.c_int(M), C.c_int(out["num_rendered"]),
            _p(cam["bg"]), C.c_int(W), C.c_int(H), _p(g["means3D"]), _p(sh), _p(colors_precomp),
            _p(scales), C.c_float(scale_modifier), _p(rots), _p(cov3D_precomp),
            _p(cam["viewmatrix"]), _p(cam["projmatrix"]), _p(cam["projmatrix_raw"]),
            _p(cam["campos"]), C.c_float(cam["tanfovx"]), C.c_float(cam["tanfovy"]),
            _p(out["radii"]), _p(dL_dcolor.contiguous()), _p(dL_ddepth.contiguous()),
            _p(gr["dL_dmeans2D"]), _p(gr["dL_dconic"]), _p(gr["dL_dopacity"]), _p(gr["dL_dcolors"]),
            _p(gr["dL_ddepths"]), _p(gr["dL_dmeans3D"]), _p(gr["dL_dcov3D"]), _p(gr["dL_dsh"]),
            _p(gr["dL_dscales"]), _p(gr["dL_drotations"]), _p(gr["dL_dtau"]), C.c_int(int(debug)))
        torch.cuda.synchronize()
        if rc != 0:
            raise RuntimeError("ref_backward failed")
        gr["dL_dtau_sum"] = gr["dL_dtau"].double().sum(0).float()
        return gr

    def export_state(self):
        g, cam, _, _, _, out, _ = self._last
        dev = g["means3D"].device
        P = g["means3D"].shape[0]
        R = out["num_rendered"]
        W, H = cam["width"], cam["height"]
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        f = lambda *s: torch.zeros(*s, device=dev)
        u32 = lambda *s: torch.zeros(*s, dtype=torch.int32, device=dev)
        st = dict(depths=f(P), means2D=f(P, 2), cov3D=f(P, 6), conic_opacity=f(P, 4), rgb=f(P, 3),
                  tiles_touched=u32(P), point_offsets=u32(P),
                  clamped=torch.zeros(P, 3, dtype=torch.uint8, device=dev),
                  point_list_keys=torch.zeros(max(R, 1), dtype=torch.int64, device=dev),
                  point_list=u32(max(R, 1)), ranges=u32(tiles, 2), n_contrib=u32(H * W),
                  final_T=f(H * W))
        rc = lib().ref_export_state(
            self.h, _p(st["depths"]), _p(st["means2D"]), _p(st["cov3D"]), _p(st["conic_opacity"]),
            _p(st["rgb"]), _p(st["tiles_touched"]), _p(st["point_offsets"]), _p(st["clamped"]),
            _p(st["point_list_keys"]), _p(st["point_list"]), C.c_void_p(0), C.c_void_p(0),
            _p(st["ranges"]), _p(st["n_contrib"]), _p(st["final_T"]))
        if rc != 0:
            raise RuntimeError("ref_export_state failed")
        st["point_list_keys"] = st["point_list_keys"][:R]
        st["point_list"] = st["point_list"][:R]
        return st


def to_cuda(d, device="cuda:0"):
    """numpy scene dict -> cuda tensors (scalars kept)."""
    import numpy as np
    out = {}
    for k, v in d.items():
        if isinstance(v, np.ndarray) and v.dtype == np.float32:
            out[k] = torch.from_numpy(v).to(device)
        else:
            out[k] = v
    return out
