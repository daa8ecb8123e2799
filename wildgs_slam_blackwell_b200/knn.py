"""distCUDA2 (include/wgs_knn.h): host-side mirror of the reference's
simple_knn._C.distCUDA2 (thirdparty/simple-knn/spatial.cu:14-35) -- same name, argument and result:
points (P,3) float CUDA tensor -> (P,) float32, mean squared distance to the 3 nearest other points.
The kernels live in libwgs_rasterizer.so; there is no CPU fallback."""
import ctypes as C

import torch

from . import _lib

EXPORTS = ("wgs_knn_workspace_bytes", "wgs_dist_cuda2")
_proto_done = False


def _lib_knn():
    global _proto_done
    lib = _lib.load()
    if not _proto_done:
        lib.wgs_knn_workspace_bytes.restype = C.c_size_t
        lib.wgs_knn_workspace_bytes.argtypes = [C.c_int32]
        lib.wgs_dist_cuda2.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _proto_done = True
    return lib


def distCUDA2(points):
    if not points.is_cuda:
        raise RuntimeError("distCUDA2 runs on a CUDA (sm_100) device only; there is no CPU fallback")
    if points.dim() != 2 or points.shape[1] != 3:
        raise RuntimeError("points must have dimensions (num_points, 3)")
    device = points.device
    pts = points.detach().to(torch.float32).contiguous()
    P = pts.shape[0]
    means = torch.full((P,), 0.0, dtype=torch.float32, device=device)   # spatial.cu:28
    if P == 0:
        return means
    lib = _lib_knn()
    ws = torch.empty(int(lib.wgs_knn_workspace_bytes(P)), dtype=torch.uint8, device=device)
    with torch.cuda.device(device):
        rc = lib.wgs_dist_cuda2(P, pts.data_ptr(), means.data_ptr(), ws.data_ptr(),
                                C.c_void_p(torch.cuda.current_stream(device).cuda_stream))
    if rc != _lib.WGS_OK:
        raise RuntimeError(f"wgs_dist_cuda2 failed ({rc}): {_lib.last_error()}")
    return means
