"""Fused row compaction for prune / densify (include/wgs_compact.h).

`compact_rows(keep, tensors)` == `[t[keep] for t in tensors]`, bit for bit, with one scan of the mask and one
gather launch for all tensors.  `prune_points(gaussians, mask)` mirrors GaussianModel.prune_points +
_prune_optimizer (thirdparty/gaussian_splatting/scene/gaussian_model.py:526-564): same attribute names, same
optimizer-state surgery, the 21 device tensors compacted together.  No CPU fallback for the device tensors."""
import ctypes as C

import torch
import torch.nn as nn

from . import _lib

EXPORTS = ("wgs_compact_workspace_bytes", "wgs_compact_plan", "wgs_compact_plan_static", "wgs_compact_rows")
STATIC_MAX_ROWS = 1 << 24      # WGS_COMPACT_STATIC_MAX_ROWS
MAX_TENSORS = 32


class CompactTensor(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("row_bytes", C.c_uint32)]


_proto_done = False


def _lib_compact():
    global _proto_done
    lib = _lib.load()
    if not _proto_done:
        lib.wgs_compact_workspace_bytes.restype = C.c_size_t
        lib.wgs_compact_workspace_bytes.argtypes = [C.c_int32]
        lib.wgs_compact_plan.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.wgs_compact_plan_static.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.wgs_compact_rows.argtypes = [C.c_int32, C.c_void_p, C.POINTER(CompactTensor), C.c_int32, C.c_void_p]
        _proto_done = True
    return lib


def compact_rows(keep, tensors):
    """keep: (P,) bool CUDA tensor; tensors: CUDA tensors with leading dimension P whose rows are a multiple of
    4 bytes.  Returns the list of compacted tensors (one host synchronisation, to learn the surviving count --
    boolean indexing in torch has the same one, per tensor)."""
    if not keep.is_cuda:
        raise RuntimeError("compact_rows runs on a CUDA (sm_100) device only; there is no CPU fallback")
    device = keep.device
    P = keep.shape[0]
    keep_u8 = keep.to(torch.uint8).contiguous() if keep.dtype != torch.uint8 else keep.contiguous()
    lib = _lib_compact()
    srcs = []
    for t in tensors:
        if t.device != device or t.shape[0] != P:
            raise RuntimeError("compact_rows: every tensor must live on the mask's device and have P rows")
        t = t.detach().contiguous()
        row_bytes = (t.numel() // max(P, 1)) * t.element_size() if P else 0
        if row_bytes % 4:
            raise RuntimeError("compact_rows: row size must be a multiple of 4 bytes")
        srcs.append((t, row_bytes))
    stream = C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
    dst_index = torch.empty(max(P, 1), dtype=torch.int32, device=device)
    count = torch.empty(1, dtype=torch.int32, device=device)
    ws = torch.empty(int(lib.wgs_compact_workspace_bytes(P)), dtype=torch.uint8, device=device)
    with torch.cuda.device(device):
        rc = lib.wgs_compact_plan(P, keep_u8.data_ptr(), dst_index.data_ptr(), count.data_ptr(), ws.data_ptr(), stream)
        if rc != _lib.WGS_OK:
            raise RuntimeError(f"wgs_compact_plan failed ({rc}): {_lib.last_error()}")
        n_keep = int(count.item())
        outs = [torch.empty((n_keep,) + tuple(t.shape[1:]), dtype=t.dtype, device=device) for t, _ in srcs]
        for i in range(0, len(srcs), MAX_TENSORS):
            chunk = [(s, o) for s, o in zip(srcs[i:i + MAX_TENSORS], outs[i:i + MAX_TENSORS]) if s[1] and n_keep]
            if not chunk:
                continue
            arr = (CompactTensor * len(chunk))()
            for k, ((t, rb), o) in enumerate(chunk):
                arr[k].src, arr[k].dst, arr[k].row_bytes = t.data_ptr(), o.data_ptr(), rb
            rc = lib.wgs_compact_rows(P, dst_index.data_ptr(), arr, len(chunk), stream)
            if rc != _lib.WGS_OK:
                raise RuntimeError(f"wgs_compact_rows failed ({rc}): {_lib.last_error()}")
    return outs


_PARAM_ATTR = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity",
               "scaling": "_scaling", "rotation": "_rotation"}


def prune_points(gaussians, mask):
    """GaussianModel.prune_points(mask) (gaussian_model.py:548-564): `mask` marks the points to REMOVE."""
    keep = ~mask
    opt = gaussians.optimizer
    todo, slots = [], []
    for group in opt.param_groups:
        p = group["params"][0]
        st = opt.state.get(p, None)
        todo.append(p)
        slots.append((group, None))
        if st is not None:
            todo += [st["exp_avg"], st["exp_avg_sq"]]
            slots += [(group, "exp_avg"), (group, "exp_avg_sq")]
    stats = [n for n in ("xyz_gradient_accum", "denom", "max_radii2D") if torch.is_tensor(getattr(gaussians, n, None))
             and getattr(gaussians, n).is_cuda and getattr(gaussians, n).shape[:1] == keep.shape]
    todo += [getattr(gaussians, n) for n in stats]
    outs = compact_rows(keep, todo)
    it = iter(outs)
    for group in opt.param_groups:
        old = group["params"][0]
        st = opt.state.get(old, None)
        new = nn.Parameter(next(it).requires_grad_(True))
        if st is not None:
            st["exp_avg"], st["exp_avg_sq"] = next(it), next(it)
            del opt.state[old]
            opt.state[new] = st
        group["params"][0] = new
        attr = _PARAM_ATTR.get(group.get("name"))
        if attr is not None:
            setattr(gaussians, attr, new)
    for n in stats:
        setattr(gaussians, n, next(it))
    keep_cpu = keep.cpu()
    for n in ("unique_kfIDs", "n_obs"):          # CPU bookkeeping tensors stay torch (gaussian_model.py:563-564)
        t = getattr(gaussians, n, None)
        if torch.is_tensor(t) and t.shape[:1] == keep_cpu.shape:
            setattr(gaussians, n, t[keep_cpu])
