"""Fused photometric loss + cotangent producer (include/wgs_loss.h).

Host-side mirror of the reference's loss entry points for the rasterizer's two callers,
same names and argument meaning (src/utils/slam_utils.py):

    get_loss_tracking(config, image, depth, opacity, viewpoint, monocular=True, uncertainty=None)   (:47-53)
    get_loss_tracking_rgb(config, image, opacity, viewpoint, uncertainty=None)                      (:55-87)
    get_loss_mapping(config, image, depth, viewpoint, initialization=False)                         (:107-113)
    get_loss_mapping_rgbd(config, image, depth, viewpoint)                                          (:116-143)

Each returns a scalar tensor whose `.backward()` feeds dL/dcolor, dL/ddepth and
dL/d(exposure_a, exposure_b) to autograd, like the reference's torch expression, but the forward
is ONE kernel that also produces those gradients (the backward only scales them by the incoming
gradient).  `photometric_loss()` is the raw form: loss and cotangents without autograd, for
callers that drive `torch.autograd.backward([color, depth], [dL_dcolor, dL_ddepth])` themselves.

ssim_loss=True (a convolution chain) is not fused: those configs raise.
No CPU / PyTorch fallback: the kernel lives in libwgs_rasterizer.so.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib

_fp = C.c_void_p


class LossArgs(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32),
        ("rgb_boundary_threshold", C.c_float), ("depth_min", C.c_float),
        ("color_weight", C.c_float), ("depth_weight", C.c_float),
        ("depth_max", C.c_float), ("depth_margin", C.c_float), ("depth_weighted", C.c_int32),
        ("patch_w", C.c_int32), ("patch_h", C.c_int32), ("uncertainty_rate", C.c_float),
        ("color", _fp), ("depth", _fp), ("opacity", _fp), ("gt_color", _fp), ("gt_depth", _fp),
        ("pixel_mask", _fp), ("uncertainty", _fp), ("exposure_a", _fp), ("exposure_b", _fp),
        ("loss", _fp), ("dL_dcolor", _fp), ("dL_ddepth", _fp), ("dL_dexposure", _fp), ("weight_mean", _fp),
        ("workspace", _fp),
    ]


class UncertaintyLossArgs(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("patch_w", C.c_int32), ("patch_h", C.c_int32),
                ("median", C.c_int32), ("ssim_weight", C.c_float), ("depth_threshold", C.c_float),
                ("depth_mult", C.c_float), ("opacity_th", C.c_float), ("weight", C.c_float),
                ("uncertainty", _fp), ("luminance", _fp), ("contrast", _fp), ("structure", _fp), ("opacity", _fp),
                ("depth", _fp), ("ref_depth", _fp), ("loss", _fp), ("dL_duncertainty", _fp), ("workspace", _fp)]


EXPORTS = ("wgs_loss_workspace_bytes", "wgs_photometric_loss")
EXPORTS_UNCERTAINTY = ("wgs_uncertainty_workspace_bytes", "wgs_uncertainty_loss")
_ws = {}
_proto_done = False


def _lib_loss():
    global _proto_done
    lib = _lib.load()
    if not _proto_done:
        lib.wgs_loss_workspace_bytes.restype = C.c_size_t
        lib.wgs_loss_workspace_bytes.argtypes = [C.c_int32, C.c_int32]
        lib.wgs_photometric_loss.argtypes = [C.POINTER(LossArgs), C.c_void_p]
        lib.wgs_uncertainty_workspace_bytes.restype = C.c_size_t
        lib.wgs_uncertainty_workspace_bytes.argtypes = [C.c_int32, C.c_int32]
        lib.wgs_uncertainty_loss.argtypes = [C.POINTER(UncertaintyLossArgs), C.c_void_p]
        _proto_done = True
    return lib


def _workspace(device, W, H):
    # per stream: the block partials / ticket of two losses in flight on two streams must not mix (a mapping
    # window runs its views on two streams)
    key = (device.index, int(torch.cuda.current_stream(device).cuda_stream), W, H)
    ws = _ws.get(key)
    if ws is None:
        ws = torch.zeros(int(_lib_loss().wgs_loss_workspace_bytes(W, H)), dtype=torch.uint8, device=device)
        _ws[key] = ws
    return ws


def _f32(t, device, shape=None):
    if t is None:
        return None
    if t.dtype != torch.float32 or t.device != device:
        t = t.to(device=device, dtype=torch.float32)
    t = t.contiguous()
    if shape is not None and tuple(t.shape) != tuple(shape):
        t = t.reshape(shape)
    return t


def photometric_loss(color, gt_color, *, depth=None, gt_depth=None, opacity=None, pixel_mask=None,
                     uncertainty=None, exposure_a=None, exposure_b=None, rgb_boundary_threshold=0.01,
                     depth_min=0.01, color_weight=1.0, depth_weight=0.0, depth_max=0.0, depth_weighted=False,
                     depth_margin=1.0, uncertainty_rate=None, weight_mean=None, _packed=False):
    """One launch: returns (loss (1,), dL_dcolor (3,H,W), dL_ddepth (1,H,W), dL_dexposure (2,) or None)
    (+ the (4HW,) buffer both cotangents are views of, with _packed).
    See include/wgs_loss.h for the exact expression.  uncertainty_rate given: `uncertainty` is the patch-resolution
    network output, up-sampled inside; weight_mean: a (1,) float tensor that receives mean(w)."""
    if color.dim() != 3 or color.shape[0] != 3:
        raise RuntimeError("color must have dimensions (3, H, W)")
    device = color.device
    if device.type != "cuda":
        raise RuntimeError("photometric_loss runs on a CUDA (sm_100) device only; there is no CPU fallback")
    _, H, W = color.shape
    color = _f32(color.detach(), device)
    gt_color = _f32(gt_color, device, (3, H, W))
    depth = _f32(None if depth is None else depth.detach(), device, (1, H, W))
    gt_depth = _f32(gt_depth, device, (1, H, W))
    opacity = _f32(None if opacity is None else opacity.detach(), device, (1, H, W))
    pixel_mask = _f32(pixel_mask, device, (1, H, W))
    patch = (0, 0)
    if uncertainty_rate is not None:
        patch = (int(uncertainty.shape[-1]), int(uncertainty.shape[-2]))
        uncertainty = _f32(uncertainty.detach(), device)
    else:
        uncertainty = _f32(uncertainty, device, (H, W))
    ea = _f32(None if exposure_a is None else exposure_a.detach(), device, (1,))
    eb = _f32(None if exposure_b is None else exposure_b.detach(), device, (1,))
    if (ea is None) != (eb is None):
        raise RuntimeError("exposure_a and exposure_b go together")
    if depth_weight != 0.0 and (depth is None or gt_depth is None):
        raise RuntimeError("depth_weight != 0 needs depth and gt_depth")
    out = torch.empty(4 * H * W + 4, dtype=torch.float32, device=device)  # one allocation for all outputs
    dL_dcolor = out[:3 * H * W].view(3, H, W)
    dL_ddepth = out[3 * H * W:4 * H * W].view(1, H, W)
    loss = out[4 * H * W:4 * H * W + 1]
    dexp = out[4 * H * W + 1:4 * H * W + 3]
    a = LossArgs()
    a.width, a.height = W, H
    a.rgb_boundary_threshold, a.depth_min = float(rgb_boundary_threshold), float(depth_min)
    a.color_weight, a.depth_weight = float(color_weight), float(depth_weight)
    a.depth_max, a.depth_margin, a.depth_weighted = float(depth_max), float(depth_margin), int(bool(depth_weighted))
    a.patch_w, a.patch_h = patch
    a.uncertainty_rate = float(uncertainty_rate) if uncertainty_rate is not None else 1.0
    a.weight_mean = None if weight_mean is None else weight_mean.data_ptr()
    p = lambda t: None if t is None else t.data_ptr()
    a.color, a.depth, a.opacity, a.gt_color, a.gt_depth = p(color), p(depth), p(opacity), p(gt_color), p(gt_depth)
    a.pixel_mask, a.uncertainty, a.exposure_a, a.exposure_b = p(pixel_mask), p(uncertainty), p(ea), p(eb)
    a.loss, a.dL_dcolor, a.dL_ddepth = p(loss), p(dL_dcolor), p(dL_ddepth)
    a.dL_dexposure = p(dexp) if ea is not None else None
    a.workspace = _workspace(device, W, H).data_ptr()
    with torch.cuda.device(device):
        rc = _lib_loss().wgs_photometric_loss(C.byref(a), C.c_void_p(torch.cuda.current_stream(device).cuda_stream))
    if rc != _lib.WGS_OK:
        raise RuntimeError(f"wgs_photometric_loss failed ({rc}): {_lib.last_error()}")
    if _packed:
        return loss, dL_dcolor, dL_ddepth, (dexp if ea is not None else None), out[:4 * H * W]
    return loss, dL_dcolor, dL_ddepth, (dexp if ea is not None else None)


class _PhotometricLoss(torch.autograd.Function):
    """loss = f(color, depth, exposure_a, exposure_b); the cotangents are computed in forward."""

    @staticmethod
    def forward(ctx, color, depth, exposure_a, exposure_b, kw):
        loss, dc, dd, de, packed = photometric_loss(color, kw["gt_color"], depth=depth, gt_depth=kw.get("gt_depth"),
                                                    opacity=kw.get("opacity"), pixel_mask=kw.get("pixel_mask"),
                                                    uncertainty=kw.get("uncertainty"), exposure_a=exposure_a,
                                                    exposure_b=exposure_b,
                                                    rgb_boundary_threshold=kw["rgb_boundary_threshold"],
                                                    color_weight=kw["color_weight"], depth_weight=kw["depth_weight"],
                                                    depth_max=kw.get("depth_max", 0.0),
                                                    depth_weighted=kw.get("depth_weighted", False),
                                                    uncertainty_rate=kw.get("uncertainty_rate"),
                                                    weight_mean=kw.get("weight_mean"), _packed=True)
        ctx.has_depth = depth is not None and kw["depth_weight"] != 0.0
        ctx.has_exp = de is not None
        ctx.shape = tuple(dc.shape)
        ctx.save_for_backward(packed, de if de is not None else dc.new_empty(0))
        return loss.reshape(())

    @staticmethod
    def backward(ctx, g):
        packed, de = ctx.saved_tensors
        _, H, W = ctx.shape
        ga = gb = None
        if ctx.has_exp:
            ge = de * g
            ga, gb = ge[0:1], ge[1:2]
        gp = packed * g                      # both cotangents in one launch (they share the kernel's output buffer)
        return gp[:3 * H * W].view(3, H, W), (gp[3 * H * W:].view(1, H, W) if ctx.has_depth else None), ga, gb, None


def _gt_image(viewpoint, device):
    return viewpoint.original_image.to(device)


def _gt_depth(viewpoint, device):
    """viewpoint.depth is a numpy (H,W) array (src/utils/camera_utils.py); the reference re-uploads it on
    every call (slam_utils.py:123-127).  Here the device copy is cached on the viewpoint."""
    cache = getattr(viewpoint, "_wgs_depth_dev", None)
    if cache is None or cache[0] is not viewpoint.depth or cache[1].device != device:
        d = viewpoint.depth
        t = torch.from_numpy(np.ascontiguousarray(d, np.float32)) if isinstance(d, np.ndarray) else d.float()
        cache = (viewpoint.depth, t.to(device)[None].contiguous())
        try:
            viewpoint._wgs_depth_dev = cache
        except AttributeError:
            pass
    return cache[1]


def get_loss_tracking_rgb(config, image, opacity, viewpoint, uncertainty=None, _exposure=(None, None)):
    kw = dict(gt_color=_gt_image(viewpoint, image.device), opacity=opacity,
              pixel_mask=getattr(viewpoint, "grad_mask", None), uncertainty=uncertainty,
              rgb_boundary_threshold=config["Training"]["rgb_boundary_threshold"], color_weight=1.0,
              depth_weight=0.0)
    return _PhotometricLoss.apply(image, None, _exposure[0], _exposure[1], kw)


def get_loss_tracking(config, image, depth, opacity, viewpoint, monocular=True, uncertainty=None):
    if not monocular:
        raise NotImplementedError("Only implemented monocular, not rgbd for uncertainty-aware tracking")
    return get_loss_tracking_rgb(config, image, opacity, viewpoint, uncertainty,
                                 _exposure=(viewpoint.exposure_a, viewpoint.exposure_b))


def get_loss_mapping_rgbd(config, image, depth, viewpoint, _exposure=(None, None)):
    """slam_utils.py:116-143.  alpha * mean((1 - l) * l1_rgb + l * (1 - ssim)) + (1 - alpha) * mean(l1_depth): the L1 terms
    are one launch (wgs_photometric_loss, exposure applied inside), the SSIM term two (wgs_ssim_forward/backward)."""
    from . import loss_utils
    alpha = config["Training"]["alpha"] if "alpha" in config["Training"] else 0.95
    lam = config["opt_params"]["lambda_dssim"] if config["Training"].get("ssim_loss", False) else 0.0
    gt = _gt_image(viewpoint, image.device)
    kw = dict(gt_color=gt, gt_depth=_gt_depth(viewpoint, image.device),
              rgb_boundary_threshold=config["Training"]["rgb_boundary_threshold"], color_weight=alpha * (1.0 - lam),
              depth_weight=1.0 - alpha)
    loss = _PhotometricLoss.apply(image, depth, _exposure[0], _exposure[1], kw)
    if lam:
        loss = loss + (alpha * lam) * (1.0 - loss_utils.ssim(image, gt, exposure=_exposure))
    return loss


def get_loss_mapping(config, image, depth, viewpoint, initialization=False):
    exposure = (None, None) if initialization else (viewpoint.exposure_a, viewpoint.exposure_b)
    return get_loss_mapping_rgbd(config, image, depth, viewpoint, _exposure=exposure)


def compute_bias_factor(x, s):
    """mapping_utils.py:45-57 (the nerf-on-the-go schedule)"""
    return x / (1 + (1 - x) * (1 / s - 2))


def _depth_threshold(viewpoint, ref_depth):
    """min(10 * median(ref_depth), 50) (mapping_utils.py:262-263): a property of the keyframe's reference depth, so it is
    computed (one host read) when the depth is first seen instead of on every iteration."""
    cache = getattr(viewpoint, "_wgs_depth_thr", None)
    if cache is None or cache[0] is not viewpoint.depth:
        cache = (viewpoint.depth, min(float(10 * ref_depth.median()), 50.0))   # float32 product, as the reference
        try:
            viewpoint._wgs_depth_thr = cache
        except AttributeError:
            pass
    return cache[1]


class _UncertaintyLoss(torch.autograd.Function):
    """weight * mean(uncertainty loss) of mapping_utils.py:270-321 as a function of the network's uncertainty map; every
    other input is detached in the reference.  Two launches (wgs_uncertainty_loss); the gradient is made in forward."""

    @staticmethod
    def forward(ctx, uncertainty, lum, con, stu, opacity, depth, ref_depth, kw):
        dev = uncertainty.device
        hp, wp = uncertainty.shape
        h, w = lum.shape[-2:]
        f = lambda t: _f32(t.detach(), dev, (h, w))
        unc = _f32(uncertainty.detach(), dev)
        ins = [f(t) for t in (lum, con, stu, opacity, depth, ref_depth)]
        lib = _lib_loss()
        out = torch.empty(hp * wp + 1, device=dev)
        need = uncertainty.requires_grad and not kw["freeze"]
        ws = torch.empty(int(lib.wgs_uncertainty_workspace_bytes(wp, hp)), dtype=torch.uint8, device=dev)
        a = UncertaintyLossArgs(w, h, wp, hp, int(kw["median"]), float(kw["ssim_weight"]), float(kw["depth_threshold"]),
                                float(kw["depth_mult"]), float(kw["opacity_th"]), float(kw["weight"]), unc.data_ptr(),
                                *[t.data_ptr() for t in ins], out[hp * wp:].data_ptr(),
                                out.data_ptr() if need else None, ws.data_ptr())
        with torch.cuda.device(dev):
            rc = lib.wgs_uncertainty_loss(C.byref(a), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
        if rc != _lib.WGS_OK:
            raise RuntimeError(f"wgs_uncertainty_loss failed ({rc}): {_lib.last_error()}")
        ctx.need = need
        if need:
            ctx.save_for_backward(out[:hp * wp].view(hp, wp))
        return out[hp * wp:].reshape(())

    @staticmethod
    def backward(ctx, g):
        return (ctx.saved_tensors[0] * g if ctx.need else None,) + (None,) * 7


def get_loss_mapping_uncertainty(config, rendered_img, rendered_depth, viewpoint, opacity, uncertainty_network,
                                 train_frac, ssim_frac, initialization=False, freeze_uncertainty_loss=False):
    """slam_utils.get_loss_mapping_uncertainty (slam_utils.py:146-258) with compute_mapping_loss_components
    (mapping_utils.py:207-323): same arguments, returns (uncertainty, total_loss).

    The full-resolution work is fused: the uncertainty-weighted L1 colour and depth terms and their cotangents are
    ONE launch (wgs_photometric_loss: exposure, rgb mask, weights 0.5/u^2 with the 0.1 cut, depth mask with the
    10x-median bound, depth weighted where the reference depth lies in front), the SSIM term two launches, the
    luminance/contrast/structure maps of the uncertainty loss one, the patch-resolution uncertainty loss and its
    gradient (resampling, 5x5 median, log) two (wgs_uncertainty_loss); the network's patch-resolution output is
    up-sampled to the per-pixel weights inside the L1 launch.  What stays in torch is the uncertainty MLP."""
    from . import loss_utils
    dev = rendered_img.device
    tr, ucfg = config["Training"], config["uncertainty_params"]
    ea, eb = (None, None) if initialization else (viewpoint.exposure_a, viewpoint.exposure_b)
    alpha = tr.get("alpha", 0.95)
    lam = config["opt_params"]["lambda_dssim"] if tr["ssim_loss"] else 0.0
    gt = _gt_image(viewpoint, dev)
    ref_depth = _gt_depth(viewpoint, dev)
    _, h, w = gt.shape
    if config.get("full_resolution", False) and tuple(rendered_depth.shape[-2:]) != (h, w):
        raise NotImplementedError("full_resolution with a depth map of another size than the image")
    depth_thr = _depth_threshold(viewpoint, ref_depth)

    uncertainty = uncertainty_network(viewpoint.features.to(device=dev))
    with torch.no_grad():
        lum, con, stu = loss_utils.compute_ssim_components(gt, rendered_img, window_size=ucfg["ssim_window_size"],
                                                            exposure=(ea, eb))
    uncer = _UncertaintyLoss.apply(uncertainty, lum, con, stu, opacity, rendered_depth, ref_depth, dict(
        median=ucfg["ssim_median_filter_size"], ssim_weight=100 + 900 * compute_bias_factor(ssim_frac, 0.8),
        depth_threshold=depth_thr, depth_mult=ucfg["uncer_depth_mult"], opacity_th=ucfg["opacity_th_for_uncer_loss"],
        weight=ucfg["ssim_mult"], freeze=freeze_uncertainty_loss))

    w_mean = torch.empty(1, device=dev)     # mean of the per-pixel weights, written by the L1 launch
    kw = dict(gt_color=gt, gt_depth=ref_depth, uncertainty=uncertainty, rgb_boundary_threshold=tr["rgb_boundary_threshold"],
              color_weight=alpha * (1.0 - lam), depth_weight=1.0 - alpha, depth_max=depth_thr, depth_weighted=True,
              uncertainty_rate=1 + compute_bias_factor(train_frac, 0.8), weight_mean=w_mean)
    total = _PhotometricLoss.apply(rendered_img, rendered_depth, ea, eb, kw)
    if lam:
        total = total + (alpha * lam) * w_mean[0] * (1.0 - loss_utils.ssim(rendered_img, gt, exposure=(ea, eb)))
    total = total + uncer
    return uncertainty, total
