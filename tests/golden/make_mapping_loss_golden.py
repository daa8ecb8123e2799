#!/usr/bin/env python
"""Golden vectors for the mapper's real loss functions, produced by the REFERENCE's own Python code (imported from
/root/reference, torch CPU + autograd) on seeded inputs:

    python tests/golden/make_mapping_loss_golden.py      # build container only: /root/reference must exist

  mapping_ssim.npz         get_loss_mapping (ssim_loss on, exposure on)            slam_utils.py:107-143
  mapping_uncertainty.npz  get_loss_mapping_uncertainty                            slam_utils.py:146-258
  mapping_uncertainty_init.npz   ... initialization=True, freeze_uncertainty_loss=True, other fractions
Inputs, loss, gradients w.r.t. image, depth, exposure a/b and the stand-in uncertainty network's parameters."""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

REF = "/root/reference"
sys.path.insert(0, REF)
from src.utils import slam_utils  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
H, W, HP, WP, FD = 84, 112, 6, 8, 12
CONFIG = {"Training": {"rgb_boundary_threshold": 0.01, "alpha": 0.5, "ssim_loss": True},
          "opt_params": {"lambda_dssim": 0.2}, "full_resolution": False,
          "uncertainty_params": {"ssim_window_size": 7, "ssim_median_filter_size": 5, "opacity_th_for_uncer_loss": 0.9,
                                 "ssim_mult": 0.5, "uncer_depth_mult": 0.2}}


class _T(torch.Tensor):
    def cuda(self, *a, **k):
        return self.as_subclass(torch.Tensor)


class _View:
    pass


def make_net(seed):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(FD, 16), torch.nn.ReLU(), torch.nn.Linear(16, 1), torch.nn.Softplus(),
                               torch.nn.Flatten(-2, -1))


def make_inputs(seed):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g)
    k = torch.ones(1, 1, 7, 7) / 49
    lp = lambda t: F.conv2d(t.reshape(-1, 1, H, W), k, padding=3).reshape(t.shape)
    gt = (lp(r(3, H, W)) * 1.6).clamp(0, 1)
    gt[:, : H // 7, : W // 6] = 0.001                       # below rgb_boundary_threshold
    image = (gt + 0.15 * (r(3, H, W) - 0.5)).clamp(0, 1)
    gt_depth = 0.5 + 5.5 * lp(r(1, H, W))[0]
    gt_depth[H // 2:, : W // 5] = 0.0                       # invalid depth
    gt_depth[:6, -9:] = 45.0                                # beyond 10 x median
    depth = gt_depth[None] + 1.5 * (r(1, H, W) - 0.5)       # both sides of the +1 m margin appear
    depth[:, 20:30, 30:50] -= 2.0
    opacity = (0.85 + 0.2 * r(1, H, W)).clamp(0, 1)         # both sides of the 0.9 threshold after resampling
    feats = torch.randn(HP, WP, FD, generator=g) * 1.5
    return dict(image=image, gt=gt, depth=depth, gt_depth=gt_depth, opacity=opacity, features=feats,
                a=torch.tensor([0.07]), b=torch.tensor([-0.02]))


def run(name, d, fn):
    image = d["image"].clone().requires_grad_(True)
    depth = d["depth"].clone().requires_grad_(True)
    vp = _View()
    vp.original_image = d["gt"].as_subclass(_T)
    vp.depth = d["gt_depth"].numpy()
    vp.features = d["features"]
    vp.exposure_a = d["a"].clone().requires_grad_(True)
    vp.exposure_b = d["b"].clone().requires_grad_(True)
    net = make_net(7)
    out = fn(image, depth, vp, net)
    unc, loss = out if isinstance(out, tuple) else (None, out)
    loss.backward()
    z = lambda t, like: (t.grad if t.grad is not None else torch.zeros_like(like)).numpy()
    rec = {k: v.numpy() for k, v in d.items()}
    rec.update(loss=loss.detach().numpy(), g_image=image.grad.numpy(), g_depth=z(depth, depth),
               g_a=z(vp.exposure_a, d["a"]), g_b=z(vp.exposure_b, d["b"]))
    if unc is not None:
        rec["uncertainty"] = unc.detach().numpy()
        for i, p in enumerate(net.parameters()):
            rec[f"g_net{i}"] = (p.grad if p.grad is not None else torch.zeros_like(p)).numpy()
    np.savez_compressed(os.path.join(HERE, name), **rec)
    print(name, float(loss), float(image.grad.abs().max()), float(depth.grad.abs().max()) if depth.grad is not None else None)


d = make_inputs(11)
run("mapping_ssim.npz", d, lambda im, dp, vp, net: slam_utils.get_loss_mapping(CONFIG, im, dp, vp))
run("mapping_uncertainty.npz", d, lambda im, dp, vp, net: slam_utils.get_loss_mapping_uncertainty(
    CONFIG, im, dp, vp, d["opacity"], net, 0.3, 0.6))
run("mapping_uncertainty_init.npz", make_inputs(12), lambda im, dp, vp, net: slam_utils.get_loss_mapping_uncertainty(
    CONFIG, im, dp, vp, make_inputs(12)["opacity"], net, 0.9, 0.1, initialization=True, freeze_uncertainty_loss=True))
