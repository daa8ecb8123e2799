"""ctypes binding of libwgs_rasterizer.so (include/wgs_rasterizer.h).  No CPU fallback: if the
library is missing or the device is not sm_100 this raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# WGS_RASTERIZER_LIB: load another build of the same ABI (kernel A/B runs: scripts/build_variant.py)
LIB_PATH = os.environ.get("WGS_RASTERIZER_LIB") or os.path.join(_HERE, "libwgs_rasterizer.so")
ABI_VERSION = 11

WGS_OK = 0
WGS_ERR_INVALID_ARGUMENT = -1
WGS_ERR_CUDA = -2
WGS_ERR_CAPACITY = -3
WGS_ERR_NO_COLOR_SOURCE = -4
WGS_ERR_ARCH = -5

_fp = C.c_void_p


class WorkspaceSizes(C.Structure):
    _fields_ = [(n, C.c_size_t) for n in ("geom_bytes", "tile_bytes", "binning_bytes", "image_bytes", "grad_bytes")]


class _SizedStructure(C.Structure):
    """Fills the leading `struct_size` member (ABI v10: the library rejects a struct of another size)."""

    def __init__(self, *args, **kw):
        super().__init__(*args, **kw)
        self.struct_size = C.sizeof(self)


class ForwardArgs(_SizedStructure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("P", C.c_int32), ("sh_degree", C.c_int32), ("M", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
        ("prefiltered", C.c_int32), ("debug", C.c_int32),
        ("background", _fp), ("means3D", _fp), ("sh", _fp), ("colors_precomp", _fp), ("opacities", _fp),
        ("scales", _fp), ("rotations", _fp), ("cov3D_precomp", _fp), ("viewmatrix", _fp), ("projmatrix", _fp),
        ("campos", _fp),
        ("out_color", _fp), ("out_depth", _fp), ("out_opacity", _fp), ("radii", _fp), ("n_touched", _fp),
        ("geom", _fp), ("tile", _fp), ("binning", _fp), ("image", _fp),
        ("capacity", C.c_int64), ("tile_row_begin", C.c_int32), ("tile_row_end", C.c_int32),
        ("activations", C.c_int32), ("sh_rest", _fp),
    ]


class BackwardArgs(_SizedStructure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("P", C.c_int32), ("sh_degree", C.c_int32), ("M", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float), ("debug", C.c_int32),
        ("accumulate", C.c_int32),
        ("background", _fp), ("means3D", _fp), ("sh", _fp), ("colors_precomp", _fp), ("scales", _fp),
        ("rotations", _fp), ("cov3D_precomp", _fp), ("viewmatrix", _fp), ("projmatrix", _fp),
        ("projmatrix_raw", _fp), ("campos", _fp), ("radii", _fp),
        ("dL_dout_color", _fp), ("dL_dout_depth", _fp),
        ("dL_dmeans3D", _fp), ("dL_dmeans2D", _fp), ("dL_dsh", _fp), ("dL_dcolors", _fp), ("dL_dopacity", _fp),
        ("dL_dscales", _fp), ("dL_drotations", _fp), ("dL_dcov3D", _fp), ("dL_dtau", _fp),
        ("geom", _fp), ("tile", _fp), ("binning", _fp), ("image", _fp), ("grad", _fp),
        ("capacity", C.c_int64), ("tile_row_begin", C.c_int32), ("tile_row_end", C.c_int32),
        ("activations", C.c_int32), ("sh_rest", _fp), ("dL_dsh_rest", _fp),
        ("stat_grad_accum", _fp), ("stat_denom", _fp), ("stat_max_radii", _fp),
    ]


EXPORTS = ("wgs_abi_version", "wgs_last_error", "wgs_context_create", "wgs_context_destroy",
           "wgs_workspace_sizes", "wgs_forward", "wgs_backward", "wgs_mark_visible", "wgs_export_binning",
           "wgs_export_geometry", "wgs_stage_timing", "wgs_stage_times", "wgs_camera_matrices", "wgs_set_async",
           "wgs_last_num_rendered", "wgs_set_chained_launch", "wgs_overflow_status", "wgs_count_pairs", "wgs_peer_flags_bytes",
           "wgs_peer_allreduce", "wgs_peer_exchange", "wgs_peer_configure")
NUM_STAGES = 8
STAGE_NAMES = ("preprocess_fwd", "tile_scan", "scatter", "tile_sort", "render_fwd", "render_bwd", "preprocess_bwd",
               "reserved")

_lib = None
_contexts = {}


def load():
    """Load the shared library (once) and declare prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m wildgs_slam_blackwell_b200.build` "
            "(there is no CPU or PyTorch fallback for the rasterizer)")
    lib = C.CDLL(LIB_PATH)
    lib.wgs_abi_version.restype = C.c_int
    lib.wgs_last_error.restype = C.c_char_p
    lib.wgs_context_create.restype = C.c_void_p
    lib.wgs_context_create.argtypes = [C.c_int]
    lib.wgs_context_destroy.argtypes = [C.c_void_p]
    lib.wgs_workspace_sizes.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.POINTER(WorkspaceSizes)]
    lib.wgs_forward.argtypes = [C.c_void_p, C.POINTER(ForwardArgs), C.c_void_p, C.POINTER(C.c_int64)]
    lib.wgs_backward.argtypes = [C.c_void_p, C.POINTER(BackwardArgs), C.c_void_p]
    lib.wgs_mark_visible.argtypes = [C.c_int32, _fp, _fp, _fp, _fp, C.c_void_p]
    lib.wgs_export_binning.argtypes = [C.POINTER(ForwardArgs), C.c_int64, _fp, _fp, _fp, _fp, _fp, C.c_void_p]
    lib.wgs_export_geometry.argtypes = [C.POINTER(ForwardArgs), _fp, _fp, _fp, _fp, _fp, C.c_void_p]
    lib.wgs_camera_matrices.argtypes = [_fp, _fp, _fp, _fp, _fp, _fp, C.c_void_p]
    lib.wgs_peer_flags_bytes.restype = C.c_size_t
    lib.wgs_peer_allreduce.argtypes = [C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_uint64, C.c_int32, C.c_int32,
                                       C.c_uint64, C.c_void_p]
    lib.wgs_peer_exchange.argtypes = [C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_uint64, C.c_int32, C.c_int32,
                                      C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_int32, C.c_int32, C.c_void_p]
    lib.wgs_peer_configure.argtypes = [C.c_int32, C.c_int32]
    lib.wgs_set_async.argtypes = [C.c_void_p, C.c_int]
    lib.wgs_set_chained_launch.argtypes = [C.c_int]
    lib.wgs_last_num_rendered.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lib.wgs_overflow_status.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.c_int]
    lib.wgs_count_pairs.argtypes = [C.POINTER(ForwardArgs), _fp, C.c_void_p]
    lib.wgs_stage_timing.argtypes = [C.c_void_p, C.c_int]
    lib.wgs_stage_times.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
    if lib.wgs_abi_version() != ABI_VERSION:
        raise RuntimeError(f"libwgs_rasterizer.so ABI {lib.wgs_abi_version()} != expected {ABI_VERSION}; rebuild")
    _lib = lib
    return lib


def last_error():
    return load().wgs_last_error().decode()


def context(device_index):
    """Per-device WgsContext (created lazily; raises on a non-sm_100 device)."""
    ctx = _contexts.get(device_index)
    if ctx is None:
        h = load().wgs_context_create(int(device_index))
        if not h:
            raise RuntimeError(last_error())
        ctx = C.c_void_p(h)
        _contexts[device_index] = ctx
    return ctx


def workspace_sizes(P, W, H, capacity):
    ws = WorkspaceSizes()
    rc = load().wgs_workspace_sizes(P, W, H, capacity, C.byref(ws))
    if rc != WGS_OK:
        raise RuntimeError(last_error())
    return ws


def stage_timing(device_index, enable):
    rc = load().wgs_stage_timing(context(device_index), int(bool(enable)))
    if rc != WGS_OK:
        raise RuntimeError(last_error())


def stage_times(device_index):
    """dict stage name -> ms of the most recent forward(+backward); synchronises."""
    buf = (C.c_float * NUM_STAGES)()
    rc = load().wgs_stage_times(context(device_index), buf)
    if rc != WGS_OK:
        raise RuntimeError(last_error())
    return {n: float(buf[i]) for i, n in enumerate(STAGE_NAMES) if n != "reserved"}
