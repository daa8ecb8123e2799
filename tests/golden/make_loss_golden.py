#!/usr/bin/env python
"""Golden vectors for the fused photometric loss, produced by the REFERENCE's own Python functions
(imported from /root/reference, torch CPU + autograd) on seeded inputs:

    python tests/golden/make_loss_golden.py      # build container only: /root/reference must exist

Writes tests/golden/loss_tracking.npz, loss_tracking_unc.npz, loss_mapping.npz, loss_mapping_init.npz
(inputs + loss + gradients w.r.t. image, depth, exposure_a, exposure_b).  The reference calls
`.cuda()` on the ground-truth image; a tensor subclass whose .cuda() is the identity stands in.
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)
from src.utils import slam_utils  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


class _T(torch.Tensor):
    def cuda(self, *a, **k):
        return self.as_subclass(torch.Tensor)


class _View:
    pass


def make_inputs(seed, H, W):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g)
    image = r(3, H, W)
    gt = r(3, H, W)
    gt[:, : H // 6, : W // 5] = 0.001            # below rgb_boundary_threshold
    gt[:, 5, 7] = image[:, 5, 7]                 # exact zeros of the residual (sign(0) = 0)
    depth = 0.5 + 5.5 * r(1, H, W)
    gt_depth = 0.5 + 5.5 * r(H, W)
    gt_depth[H // 2:, : W // 4] = 0.0            # invalid depth
    opacity = r(1, H, W)
    grad_mask = r(1, H, W) > 0.3
    unc = 0.1 + 3.0 * r(H, W)                    # weights 0.5/u^2 fall below 0.1 for u > 2.236
    return dict(image=image, gt=gt, depth=depth, gt_depth=gt_depth, opacity=opacity, grad_mask=grad_mask, unc=unc,
                a=torch.tensor([0.13]), b=torch.tensor([-0.04]))


def run(kind, d, config):
    image = d["image"].clone().requires_grad_(True)
    depth = d["depth"].clone().requires_grad_(True)
    vp = _View()
    vp.original_image = d["gt"].as_subclass(_T)
    vp.grad_mask = d["grad_mask"]
    vp.depth = d["gt_depth"].numpy()
    vp.exposure_a = d["a"].clone().requires_grad_(True)
    vp.exposure_b = d["b"].clone().requires_grad_(True)
    if kind == "tracking":
        loss = slam_utils.get_loss_tracking(config, image, depth, d["opacity"], vp, monocular=True, uncertainty=None)
    elif kind == "tracking_unc":
        loss = slam_utils.get_loss_tracking(config, image, depth, d["opacity"], vp, monocular=True, uncertainty=d["unc"])
    elif kind == "mapping":
        loss = slam_utils.get_loss_mapping(config, image, depth, vp, initialization=False)
    else:
        loss = slam_utils.get_loss_mapping(config, image, depth, vp, initialization=True)
    loss.backward()
    z = lambda t, like: (t.grad if t.grad is not None else torch.zeros_like(like)).numpy()
    return dict(loss=np.float32(loss.item()), g_image=z(image, image), g_depth=z(depth, depth),
                g_a=z(vp.exposure_a, vp.exposure_a), g_b=z(vp.exposure_b, vp.exposure_b))


def main():
    config = {"Training": {"rgb_boundary_threshold": 0.01, "alpha": 0.5, "ssim_loss": False}}
    for i, kind in enumerate(("tracking", "tracking_unc", "mapping", "mapping_init")):
        H, W = (37, 53) if i % 2 == 0 else (48, 64)   # one size that is not a multiple of 4
        d = make_inputs(4000 + i, H, W)
        out = run(kind, d, config)
        np.savez_compressed(os.path.join(HERE, f"loss_{kind}.npz"), kind=kind, alpha=0.5, threshold=0.01,
                            **{k: v.numpy() for k, v in d.items()}, **out)
        print(kind, "loss", out["loss"])


if __name__ == "__main__":
    main()
