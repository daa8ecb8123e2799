#!/usr/bin/env python
"""Golden vectors for the fused SSIM, produced by the REFERENCE's own Python functions (imported from
/root/reference, torch CPU + autograd) on seeded inputs:

    python tests/golden/make_ssim_golden.py      # build container only: /root/reference must exist

Writes tests/golden/ssim_a.npz (3,37,53: ragged tiles, smaller than two windows in places), ssim_b.npz (3,96,128,
smooth images) and ssim_batch.npz (2,3,40,48, size_average False): inputs, loss_utils.ssim value, its gradient
w.r.t. img1, and mapping_utils.compute_ssim_components at window 7 and 11."""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

REF = "/root/reference"
sys.path.insert(0, REF)
from thirdparty.gaussian_splatting.utils import loss_utils  # noqa: E402
from src.utils.dyn_uncertainty import mapping_utils  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def images(seed, shape, smooth):
    g = torch.Generator().manual_seed(seed)
    a, b = torch.rand(*shape, generator=g), torch.rand(*shape, generator=g)
    if smooth:   # image-like: low-pass noise, rendered = ground truth + small error
        k = torch.ones(1, 1, 9, 9) / 81
        lp = lambda t: F.conv2d(t.reshape(-1, 1, *shape[-2:]), k, padding=4).reshape(shape)
        b = lp(b) * 2.0
        a = (b + 0.05 * (a - 0.5)).clamp(0, 1)
        b[..., :10, :12] = 0.0      # flat black corner: variances hit the eps floor of the components
        a[..., :10, :12] = 0.0
    return a, b


def record(name, seed, shape, smooth, size_average=True):
    a, b = images(seed, shape, smooth)
    x = a.clone().requires_grad_(True)
    val = loss_utils.ssim(x, b, size_average=size_average)
    wts = torch.arange(1, val.numel() + 1, dtype=torch.float32).reshape(val.shape)
    (val * wts).sum().backward()
    out = dict(img1=a.numpy(), img2=b.numpy(), ssim=val.detach().numpy(), dmean=wts.numpy(), grad=x.grad.numpy())
    for ws in (7, 11):
        l, c, s = mapping_utils.compute_ssim_components(b, a, window_size=ws)     # the reference passes (gt, rendered)
        out.update({f"lum{ws}": l.numpy(), f"con{ws}": c.numpy(), f"str{ws}": s.numpy()})
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, float(val.sum()), float(x.grad.abs().max()))


record("ssim_a.npz", 1, (3, 37, 53), False)
record("ssim_b.npz", 2, (3, 96, 128), True)
record("ssim_batch.npz", 3, (2, 3, 40, 48), True, size_average=False)
