"""Drop-in for the SSIM side of the mapping loss (include/wgs_ssim.h):

    ssim(img1, img2, window_size=11, size_average=True)      thirdparty/gaussian_splatting/utils/loss_utils.py:61-101
    compute_ssim_components(img1, img2, window_size=11)      src/utils/dyn_uncertainty/mapping_utils.py:99-204
    l1_loss / l2_loss                                        loss_utils.py:20-21, 38-39 (plain torch)

`ssim` is differentiable w.r.t. img1 (the rendered image; img2 is the ground truth, as in every call site of the
reference: slam_utils.py:130, :200): forward = two launches, backward = one.  CUDA (sm_100) only, no fallback."""
import ctypes as C

import torch

from . import _lib

EXPORTS = ("wgs_ssim_workspace_bytes", "wgs_ssim_forward", "wgs_ssim_backward", "wgs_ssim_components")


class SsimArgs(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("planes", C.c_int32), ("planes_per_mean", C.c_int32),
                ("window", C.c_int32), ("img1", C.c_void_p), ("img2", C.c_void_p), ("mean", C.c_void_p),
                ("maps", C.c_void_p), ("dL_dmean", C.c_void_p), ("dL_dimg1", C.c_void_p), ("workspace", C.c_void_p),
                ("exposure_a", C.c_void_p), ("exposure_b", C.c_void_p), ("dL_dexposure", C.c_void_p)]


_proto_done = False


def _lib_ssim():
    global _proto_done
    lib = _lib.load()
    if not _proto_done:
        lib.wgs_ssim_workspace_bytes.restype = C.c_size_t
        lib.wgs_ssim_workspace_bytes.argtypes = [C.c_int32, C.c_int32, C.c_int32]
        lib.wgs_ssim_forward.argtypes = [C.POINTER(SsimArgs), C.c_void_p]
        lib.wgs_ssim_backward.argtypes = [C.POINTER(SsimArgs), C.c_void_p]
        lib.wgs_ssim_components.argtypes = [C.c_int32] * 5 + [C.c_void_p] * 8
        _proto_done = True
    return lib


def _check(img1, img2):
    if not img1.is_cuda:
        raise RuntimeError("fused SSIM runs on a CUDA (sm_100) device only; there is no CPU fallback")
    if img1.shape != img2.shape or img1.dim() not in (3, 4):
        raise RuntimeError("ssim: img1 and img2 must have the same (C,H,W) or (B,C,H,W) shape")
    f = lambda t: t.detach().to(device=img1.device, dtype=torch.float32).contiguous()
    return f(img1), f(img2)


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _exposure(exposure, dev):
    if exposure is None or exposure[0] is None:
        return None, None
    f = lambda t: t.detach().to(device=dev, dtype=torch.float32).reshape(1).contiguous()
    return f(exposure[0]), f(exposure[1])


class _FusedSSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img1, img2, window_size, per_item, exp_a, exp_b):
        x1, x2 = _check(img1, img2)
        H, W = x1.shape[-2:]
        planes = x1.numel() // (H * W)
        ppm = x1.shape[-3] if per_item else planes
        dev = x1.device
        lib = _lib_ssim()
        ea, eb = _exposure((exp_a, exp_b), dev)
        ctx.exp_grad = ea is not None and (exp_a.requires_grad or exp_b.requires_grad)
        need_grad = img1.requires_grad or ctx.exp_grad
        mean = torch.empty(planes // ppm, device=dev)
        maps = torch.empty((3, planes, H, W), device=dev) if need_grad else None
        ws = torch.empty(int(lib.wgs_ssim_workspace_bytes(W, H, planes)), dtype=torch.uint8, device=dev)
        p = lambda t: None if t is None else t.data_ptr()
        a = SsimArgs(W, H, planes, ppm, int(window_size), x1.data_ptr(), x2.data_ptr(), mean.data_ptr(),
                     p(maps), None, None, ws.data_ptr(), p(ea), p(eb), None)
        with torch.cuda.device(dev):
            rc = lib.wgs_ssim_forward(C.byref(a), _stream(dev))
        if rc != _lib.WGS_OK:
            raise RuntimeError(f"wgs_ssim_forward failed ({rc}): {_lib.last_error()}")
        ctx.geom = (W, H, planes, ppm, int(window_size))
        ctx.in_dtype = img1.dtype
        if need_grad:
            ctx.save_for_backward(x1, x2, maps, ws, *(() if ea is None else (ea, eb)))
        return mean if per_item else mean.reshape(())

    @staticmethod
    def backward(ctx, g):
        x1, x2, maps, ws, *exp = ctx.saved_tensors
        ea, eb = exp if exp else (None, None)
        W, H, planes, ppm, window = ctx.geom
        dev = x1.device
        g = g.detach().to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
        out = torch.empty_like(x1)
        dexp = torch.empty(2, device=dev) if ctx.exp_grad else None
        p = lambda t: None if t is None else t.data_ptr()
        a = SsimArgs(W, H, planes, ppm, window, x1.data_ptr(), x2.data_ptr(), None, maps.data_ptr(), g.data_ptr(),
                     out.data_ptr(), ws.data_ptr(), p(ea), p(eb), p(dexp))
        with torch.cuda.device(dev):
            rc = _lib_ssim().wgs_ssim_backward(C.byref(a), _stream(dev))
        if rc != _lib.WGS_OK:
            raise RuntimeError(f"wgs_ssim_backward failed ({rc}): {_lib.last_error()}")
        ga, gb = (dexp[0:1], dexp[1:2]) if dexp is not None else (None, None)
        return out.to(ctx.in_dtype), None, None, None, ga, gb


def ssim(img1, img2, window_size=11, size_average=True, exposure=None):
    """loss_utils.ssim: mean SSIM (0-dim) or, with size_average=False, one value per batch item (4-D input).
    exposure=(a, b) (extension): evaluates ssim(exp(a) * img1 + b, img2) with the compensation and its gradients
    (img1, a, b) inside the kernels -- what slam_utils.py:111/:185 followed by :130/:200 compute."""
    if not size_average and img1.dim() != 4:
        raise IndexError("ssim(size_average=False) needs (B,C,H,W) input")   # the reference's .mean(1) chain fails too
    ea, eb = (None, None) if exposure is None else exposure
    return _FusedSSIM.apply(img1, img2, window_size, not size_average, ea, eb)


def compute_ssim_components(img1, img2, window_size=11, exposure=None):
    """mapping_utils.compute_ssim_components: (luminance, contrast, structure), each the channel mean of the clipped
    component -- (H,W) for (C,H,W) input, (B,H,W) for (B,C,H,W).  Not differentiable (the reference detaches them).
    exposure=(a, b) (extension) compensates img2 -- the reference passes (gt, rendered)."""
    x1, x2 = _check(img1, img2)
    ea, eb = _exposure(exposure, x1.device)
    pe = lambda t: None if t is None else t.data_ptr()
    H, W = x1.shape[-2:]
    Cn = x1.shape[-3]
    B = x1.shape[0] if x1.dim() == 4 else 1
    dev = x1.device
    out = [torch.empty((B, H, W), device=dev) for _ in range(3)]
    with torch.cuda.device(dev):
        rc = _lib_ssim().wgs_ssim_components(W, H, B, Cn, int(window_size), x1.data_ptr(), x2.data_ptr(), pe(ea), pe(eb),
                                             out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), _stream(dev))
    if rc != _lib.WGS_OK:
        raise RuntimeError(f"wgs_ssim_components failed ({rc}): {_lib.last_error()}")
    if x1.dim() == 3:
        out = [o[0] for o in out]
    return tuple(out)


def l1_loss(network_output, gt):
    return torch.abs((network_output - gt)).mean()


def l2_loss(network_output, gt):
    return ((network_output - gt) ** 2).mean()
