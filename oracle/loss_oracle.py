"""TEST INFRASTRUCTURE ONLY (imported by tests/ and bench.py's checker legs, never by the product).

numpy restatement of the reference's photometric losses and of their gradients:
  get_loss_tracking / get_loss_tracking_rgb  -- /root/reference/src/utils/slam_utils.py:47-87
  get_loss_mapping  / get_loss_mapping_rgbd  -- /root/reference/src/utils/slam_utils.py:107-143 (ssim_loss off)
Pinned by tests/test_loss_oracle.py against tests/golden/loss_*.npz, which
tests/golden/make_loss_golden.py produced by importing the reference's own functions (torch CPU,
autograd gradients) in the build container.
"""
import numpy as np


def photometric_loss(color, gt_color, depth=None, gt_depth=None, opacity=None, pixel_mask=None, uncertainty=None,
                     exposure_a=None, exposure_b=None, rgb_boundary_threshold=0.01, depth_min=0.01,
                     color_weight=1.0, depth_weight=0.0):
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
        rd = d * dm - gd * dm                                                     # :141
        loss += depth_weight * float(np.abs(rd).sum() / (H * W))
        out["dL_ddepth"] = depth_weight / (H * W) * np.sign(rd) * dm
    out["loss"] = loss
    return out
