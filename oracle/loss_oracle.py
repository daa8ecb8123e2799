"""TEST INFRASTRUCTURE ONLY (imported by tests/ and bench.py's checker legs, never by the product).

numpy restatement of the reference's photometric losses and of their gradients:
  get_loss_tracking / get_loss_tracking_rgb  -- /root/reference/src/utils/slam_utils.py:47-87
  get_loss_mapping  / get_loss_mapping_rgbd  -- /root/reference/src/utils/slam_utils.py:107-143 (SSIM term: ssim_oracle.py)
  get_loss_mapping_uncertainty               -- /root/reference/src/utils/slam_utils.py:146-258 with
      compute_mapping_loss_components        -- /root/reference/src/utils/dyn_uncertainty/mapping_utils.py:207-323
      (mapping_uncertainty() below: float64; the resampling is torch's F.interpolate on the CPU, the library the
      reference itself calls; pinned against tests/golden/mapping_*.npz made by tests/golden/make_mapping_loss_golden.py)
Pinned by tests/test_loss_oracle.py against tests/golden/loss_*.npz, which
tests/golden/make_loss_golden.py produced by importing the reference's own functions (torch CPU,
autograd gradients) in the build container.
"""
import numpy as np


def photometric_loss(color, gt_color, depth=None, gt_depth=None, opacity=None, pixel_mask=None, uncertainty=None,
                     exposure_a=None, exposure_b=None, rgb_boundary_threshold=0.01, depth_min=0.01,
                     color_weight=1.0, depth_weight=0.0, depth_max=0.0, depth_weighted=False, depth_margin=1.0):
    """Returns dict(loss, dL_dcolor, dL_ddepth, dL_da, dL_db); arithmetic in float64."""
    f8 = lambda x: None if x is None else np.asarray(x, np.float64)
    color, gt = f8(color), f8(gt_color)
    _, H, W = color.shape
    ea = 1.0 if exposure_a is None else float(np.exp(np.float64(np.asarray(exposure_a).reshape(-1)[0])))
    eb = 0.0 if exposure_b is None else float(np.asarray(exposure_b).reshape(-1)[0])
    image_ab = ea * color + eb                                                   # :48 / :111
    # the mask compares in float32 like the reference (:72, :132)
    m = (np.asarray(gt_color, np.float32).sum(axis=0, dtype=np.float32) > np.float32(rgb_boundary_threshold))
    m = m.reshape(1, H, W).astype(np.float64)
    if pixel_mask is not None:
        m = m * f8(pixel_mask).reshape(1, H, W)                                   # :73
    r = image_ab * m - gt * m
    l1 = np.abs(r)                                                                # :76 / :133
    scale = np.ones((1, H, W))
    w = np.ones((1, H, W))
    if opacity is not None:
        scale = scale * f8(opacity).reshape(1, H, W)                              # :76
    if uncertainty is not None:
        u = np.asarray(uncertainty, np.float32).reshape(1, H, W)
        w = (np.float32(0.5) / (u * u)).astype(np.float32)                        # :81
        w = np.where(w < np.float32(0.1), np.float32(0.0), w).astype(np.float64)  # :82
        scale = scale * w
    n3 = 3.0 * H * W
    loss = color_weight * float((scale * l1).sum() / n3)                          # :86 / :143
    g_ab = color_weight / n3 * scale * np.sign(r) * m
    out = dict(dL_dcolor=ea * g_ab, dL_da=float((g_ab * ea * color).sum()), dL_db=float(g_ab.sum()),
               dL_ddepth=np.zeros((1, H, W)))
    if depth_weight != 0.0:
        d, gd = f8(depth).reshape(1, H, W), f8(gt_depth).reshape(1, H, W)
        dm = (np.asarray(gt_depth, np.float32).reshape(1, H, W) > np.float32(depth_min)).astype(np.float64)  # :140
        if depth_max > 0:                                                         # mapping_utils.py:262-266
            dm = dm * (np.asarray(gt_depth, np.float32).reshape(1, H, W) < np.float32(depth_max))
        rd = d * dm - gd * dm                                                     # :141
        wd = np.ones((1, H, W))
        if depth_weighted:                                                        # :238-242
            near = np.asarray(gt_depth, np.float32).reshape(1, H, W) < (np.asarray(depth, np.float32).reshape(1, H, W)
                                                                        + np.float32(depth_margin))
            wd = np.where(near, w, 1.0)
        loss += depth_weight * float((wd * np.abs(rd)).sum() / (H * W))
        out["dL_ddepth"] = depth_weight / (H * W) * wd * np.sign(rd) * dm
    out["loss"] = loss
    return out


def mapping_rgbd(image, gt, depth, gt_depth, a=None, b=None, alpha=0.5, lam=0.2, thr=0.01):
    """get_loss_mapping with ssim_loss on: returns loss and gradients (image, depth, a, b), float64."""
    from . import ssim_oracle
    o = photometric_loss(image, gt, depth=depth, gt_depth=gt_depth, exposure_a=a, exposure_b=b,
                         rgb_boundary_threshold=thr, color_weight=alpha * (1 - lam), depth_weight=1 - alpha)
    ea = 1.0 if a is None else float(np.exp(np.float64(np.asarray(a).reshape(-1)[0])))
    eb = 0.0 if b is None else float(np.asarray(b).reshape(-1)[0])
    iab = ea * np.asarray(image, np.float64) + eb
    g = -alpha * lam * ssim_oracle.ssim_grad(iab, gt)          # d(1 - ssim)/d image_ab
    return dict(loss=o["loss"] + alpha * lam * (1 - ssim_oracle.ssim(iab, gt)), g_image=o["dL_dcolor"] + ea * g,
                g_depth=o["dL_ddepth"], g_a=o["dL_da"] + float((g * ea * np.asarray(image, np.float64)).sum()),
                g_b=o["dL_db"] + float(g.sum()))


def mapping_uncertainty(image, gt, depth, gt_depth, opacity, uncertainty, cfg, train_frac, ssim_frac, a=None, b=None):
    """Value of get_loss_mapping_uncertainty and its gradients w.r.t. image, depth, a, b and the (H',W') uncertainty map
    the network produced (dict key g_unc; zero if cfg['freeze']).  cfg: alpha, lam, thr, window, median, opacity_th,
    ssim_mult, depth_mult."""
    import torch
    import torch.nn.functional as F
    from . import ssim_oracle
    t = lambda x: torch.from_numpy(np.asarray(x, np.float64))
    bias = lambda x, s: x / (1 + (1 - x) * (1 / s - 2))
    rs = lambda x, shape, mode="bilinear": F.interpolate(x.reshape((1, 1) + tuple(x.shape[-2:])), size=shape, mode=mode)[0, 0]
    _, H, W = np.asarray(gt).shape
    ea = 1.0 if a is None else float(np.exp(np.float64(np.asarray(a).reshape(-1)[0])))
    eb = 0.0 if b is None else float(np.asarray(b).reshape(-1)[0])
    iab = ea * np.asarray(image, np.float64) + eb
    gd32 = np.asarray(gt_depth, np.float32)
    thr = min(float(np.float32(10) * torch.from_numpy(gd32).median().numpy()), 50.0)
    unc = t(uncertainty).requires_grad_(True)
    proc = torch.clip(unc, min=0.1) + 1e-3
    resized = (rs(proc.detach(), (H, W)) - 0.1) * (1 + bias(train_frac, 0.8)) + 0.1
    wts = 0.5 / resized ** 2
    w_mean = float(torch.where(wts < 0.1, 0.0, wts).mean())
    o = photometric_loss(image, gt, depth=depth, gt_depth=gt_depth, uncertainty=resized.numpy(), exposure_a=a, exposure_b=b,
                         rgb_boundary_threshold=cfg["thr"], color_weight=cfg["alpha"] * (1 - cfg["lam"]),
                         depth_weight=1 - cfg["alpha"], depth_max=thr, depth_weighted=True)
    k = cfg["alpha"] * cfg["lam"] * w_mean
    g = -k * ssim_oracle.ssim_grad(iab, gt)
    # uncertainty loss (patch resolution)
    op = t(opacity).reshape(H, W)
    lum, con, stu = (t(x) for x in ssim_oracle.components(np.asarray(gt, np.float64), iab, cfg["window"]))
    ssim_l = torch.clip(op * (100 + 900 * bias(ssim_frac, 0.8)) * (1 - lum) * (1 - stu) * (1 - con), max=5.0)
    small = rs(ssim_l, tuple(unc.shape))
    kq = cfg["median"]
    pad = F.pad(small[None, None], ((kq - 1) // 2, kq - 1 - (kq - 1) // 2) * 2, mode="reflect")
    filt = pad.unfold(2, kq, 1).unfold(3, kq, 1).reshape(unc.shape + (-1,)).median(dim=-1)[0]
    dm = t((gd32 > np.float32(0.01)) & (gd32 < np.float32(thr))).reshape(1, H, W)
    dl1 = torch.abs(t(depth).reshape(1, H, W) * dm - t(gt_depth).reshape(1, H, W) * dm)
    sdl = rs(torch.clip(dl1[0], max=5.0), tuple(unc.shape), "bicubic")
    sdl = torch.where(rs(t(gt_depth).reshape(H, W), tuple(unc.shape), "bicubic") > thr, 0.0, sdl)
    ul = filt / proc ** 2 + 0.5 * torch.log(proc) + cfg["depth_mult"] * sdl / proc ** 2
    ul = torch.where(rs(op, tuple(unc.shape)) < cfg["opacity_th"], 0.0, ul).mean()
    g_unc = np.zeros(unc.shape)
    if not cfg.get("freeze", False):
        ul.backward()
        g_unc = cfg["ssim_mult"] * unc.grad.numpy()
    return dict(loss=o["loss"] + k * (1 - ssim_oracle.ssim(iab, gt)) + cfg["ssim_mult"] * float(ul.detach()),
                g_image=o["dL_dcolor"] + ea * g, g_depth=o["dL_ddepth"],
                g_a=o["dL_da"] + float((g * ea * np.asarray(image, np.float64)).sum()), g_b=o["dL_db"] + float(g.sum()),
                g_unc=g_unc)
