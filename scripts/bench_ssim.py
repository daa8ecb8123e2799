"""SSIM term of the mapping loss at 640x480: fused (wgs_ssim_forward/backward) vs the reference's formulation
(loss_utils.py:79-101 restated with torch ops: five depthwise conv2d + elementwise, autograd backward), both as
`loss = 1 - ssim(img, gt); loss.backward()`, and compute_ssim_components at window 7.  Prints one JSON line."""
import json
import os
import sys
from math import exp

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wildgs_slam_blackwell_b200 import loss_utils  # noqa: E402

dev = "cuda:0"
H, W = 480, 640
g = torch.Generator().manual_seed(0)
gt = torch.rand(3, H, W, generator=g).to(dev)
img = (gt + 0.05 * torch.randn(3, H, W, generator=g).to(dev)).clamp(0, 1)


def window(ws, ch):
    gs = torch.tensor([exp(-((x - ws // 2) ** 2) / float(2 * 1.5 ** 2)) for x in range(ws)])
    gs = (gs / gs.sum()).unsqueeze(1)
    return gs.mm(gs.t()).float()[None, None].expand(ch, 1, ws, ws).contiguous().to(dev)


def torch_ssim(a, b, ws=11):
    w = window(ws, a.size(-3))
    c = lambda t: F.conv2d(t, w, padding=ws // 2, groups=a.size(-3))
    mu1, mu2 = c(a), c(b)
    s1, s2, s12 = c(a * a) - mu1.pow(2), c(b * b) - mu2.pow(2), c(a * b) - mu1 * mu2
    return (((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1.pow(2) + mu2.pow(2) + 1e-4) * (s1 + s2 + 9e-4))).mean()


def timeit(fn, n=50):
    for _ in range(5):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def step(f):
    x = img.clone().requires_grad_(True)
    (1.0 - f(x, gt)).backward()
    return x.grad


ga, gb = step(loss_utils.ssim), step(torch_ssim)
out = {"what": "1 - ssim(img, gt) forward + backward, 3x480x640, ms", "fused": round(timeit(lambda: step(loss_utils.ssim)), 4),
       "torch_formulation": round(timeit(lambda: step(torch_ssim)), 4),
       "max_abs_grad_diff": float((ga - gb).abs().max()), "max_abs_grad": float(gb.abs().max()),
       "components_w7_fused_ms": round(timeit(lambda: loss_utils.compute_ssim_components(gt, img, 7)), 4)}
print(json.dumps(out))
