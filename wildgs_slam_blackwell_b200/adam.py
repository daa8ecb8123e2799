"""Fused Adam for the Gaussian model (include/wgs_adam.h).

`FusedAdam` IS a torch.optim.Adam (same constructor, param_groups and per-parameter state entries
'step', 'exp_avg', 'exp_avg_sq'), so the reference's densification code that edits optimizer.state directly
(gaussian_model.py replace_tensor_to_optimizer / cat_tensors_to_optimizer / _prune_optimizer) keeps working;
only step() differs: one launch of wgs_adam_step over all parameters instead of torch's per-op kernels.
amsgrad / weight_decay / maximize are not part of the fused kernel and raise.  CUDA float32 parameters only."""
import ctypes as C
import math

import torch

from . import _lib

EXPORTS = ("wgs_adam_step",)
MAX_TENSORS = 16


class AdamTensor(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("n", C.c_uint64), ("step_size", C.c_float), ("bias_correction2_sqrt", C.c_float),
                ("step_dev", C.c_void_p), ("lr_dev", C.c_void_p)]


_proto_done = False


def _lib_adam():
    global _proto_done
    lib = _lib.load()
    if not _proto_done:
        lib.wgs_adam_step.argtypes = [C.POINTER(AdamTensor), C.c_int32, C.c_double, C.c_double, C.c_double, C.c_void_p]
        _proto_done = True
    return lib


class FusedAdam(torch.optim.Adam):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, **kw):
        for k in ("amsgrad", "maximize"):
            if kw.get(k):
                raise NotImplementedError(f"FusedAdam: {k} is not fused")
        if kw.get("weight_decay", 0):
            raise NotImplementedError("FusedAdam: weight_decay is not fused")
        # capturable=True (torch's meaning): state['step'] is a float tensor on the device, advanced by the launch
        # itself, and a group's lr may be a device tensor -- the step then replays from a CUDA graph (graph.py)
        super().__init__(params, lr=lr, betas=betas, eps=eps, foreach=False, fused=False,
                         capturable=bool(kw.get("capturable", False)))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        batches = {}
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue                                  # like torch: untouched, step not advanced
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("FusedAdam: parameters must be contiguous float32 CUDA tensors")
                st = self.state[p]
                capt = bool(group.get("capturable", False))
                if len(st) == 0 or "exp_avg" not in st:
                    st["step"] = torch.zeros((), dtype=torch.float32, device=p.device) if capt else 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                if capt:
                    if not torch.is_tensor(st["step"]) or not st["step"].is_cuda:   # state made before the switch
                        st["step"] = torch.full((), float(st["step"]), dtype=torch.float32, device=p.device)
                    step = 1   # unused: the kernel reads the device counter
                else:
                    step = int(st.get("step", 0)) + 1
                    st["step"] = step
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                t = AdamTensor()
                t.param, t.grad = p.data_ptr(), g.data_ptr()
                t.exp_avg, t.exp_avg_sq, t.n = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), p.numel()
                lr = group["lr"]
                if capt:
                    t.step_dev = st["step"].data_ptr()
                    if torch.is_tensor(lr):
                        if not lr.is_cuda or lr.dtype != torch.float32:
                            raise RuntimeError("FusedAdam(capturable): a tensor lr must be a float32 CUDA tensor")
                        t.lr_dev, t.step_size = lr.data_ptr(), 0.0
                    else:
                        t.step_size = float(lr)
                else:
                    t.step_size = float(lr) / (1.0 - beta1 ** step)
                    t.bias_correction2_sqrt = math.sqrt(1.0 - beta2 ** step)
                batches.setdefault((p.device, beta1, beta2, group["eps"]), []).append((t, g))
        lib = _lib_adam()
        for (device, beta1, beta2, eps), items in batches.items():
            for i in range(0, len(items), MAX_TENSORS):
                chunk = items[i:i + MAX_TENSORS]
                arr = (AdamTensor * len(chunk))(*[c[0] for c in chunk])
                with torch.cuda.device(device):
                    rc = lib.wgs_adam_step(arr, len(chunk), beta1, beta2, eps,
                                           C.c_void_p(torch.cuda.current_stream(device).cuda_stream))
                if rc != _lib.WGS_OK:
                    raise RuntimeError(f"wgs_adam_step failed ({rc}): {_lib.last_error()}")
        return loss
