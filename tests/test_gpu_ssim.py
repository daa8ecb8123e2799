"""GPU: fused SSIM (wgs_ssim_forward / _backward / _components through loss_utils) against the goldens made by the
reference's loss_utils.ssim + autograd and mapping_utils.compute_ssim_components, against the float64 oracle at
other sizes, and size-independent properties at 640x480."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ssim_oracle  # noqa: E402
from wildgs_slam_blackwell_b200 import loss_utils  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


@pytest.mark.parametrize("name", ["ssim_a", "ssim_b", "ssim_batch"])
def test_matches_reference_golden(name):
    d = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    batched = d["img1"].ndim == 4
    x = t(d["img1"]).requires_grad_(True)
    val = loss_utils.ssim(x, t(d["img2"]), size_average=not batched)
    assert val.shape == d["ssim"].shape
    np.testing.assert_allclose(val.detach().cpu().numpy(), d["ssim"], rtol=2e-5, atol=2e-6)
    (val * t(d["dmean"])).sum().backward()
    want = d["grad"]
    np.testing.assert_allclose(x.grad.cpu().numpy(), want, rtol=2e-3, atol=2e-4 * np.abs(want).max())
    for ws in (7, 11):
        lum, con, stu = loss_utils.compute_ssim_components(t(d["img2"]), t(d["img1"]), window_size=ws)
        assert lum.shape == d[f"lum{ws}"].shape
        np.testing.assert_allclose(lum.cpu().numpy(), d[f"lum{ws}"], rtol=1e-4, atol=1e-5)
        # float32 E[x^2] - mu^2 cancellation noise, see tests/test_loss_oracle.py
        np.testing.assert_allclose(con.cpu().numpy(), d[f"con{ws}"], rtol=1e-4, atol=5e-3)
        assert (np.abs(stu.cpu().numpy() - d[f"str{ws}"]) > 5e-3).mean() < 5e-3


@pytest.mark.parametrize("shape,ws", [((3, 16, 16), 11), ((1, 5, 3), 11), ((3, 33, 17), 7), ((2, 1, 1), 3), ((3, 120, 160), 11)])
def test_matches_float64_oracle(shape, ws):
    g = torch.Generator().manual_seed(sum(shape) + ws)
    a, b = torch.rand(*shape, generator=g), torch.rand(*shape, generator=g)
    x = a.to(DEV).requires_grad_(True)
    val = loss_utils.ssim(x, b.to(DEV), window_size=ws)
    val.backward()
    np.testing.assert_allclose(val.item(), ssim_oracle.ssim(a.numpy(), b.numpy(), ws), rtol=2e-5, atol=2e-6)
    want = ssim_oracle.ssim_grad(a.numpy(), b.numpy(), ws)
    np.testing.assert_allclose(x.grad.cpu().numpy(), want, rtol=2e-3, atol=2e-4 * np.abs(want).max())


def test_full_size_properties():
    """640x480: ssim(x, x) = 1 with zero gradient; deterministic; gradient is linear in dL/dmean; no grad -> no maps."""
    g = torch.Generator().manual_seed(5)
    a, b = torch.rand(3, 480, 640, generator=g).to(DEV), torch.rand(3, 480, 640, generator=g).to(DEV)
    x = a.clone().requires_grad_(True)
    v = loss_utils.ssim(x, a)
    v.backward()
    assert abs(v.item() - 1.0) < 1e-6 and x.grad.abs().max().item() < 1e-6
    outs = []
    for scale in (1.0, 1.0, -2.5):
        x = a.clone().requires_grad_(True)
        v = loss_utils.ssim(x, b)
        (scale * v).backward()
        outs.append((v.item(), x.grad.clone()))
    assert outs[0][0] == outs[1][0] and torch.equal(outs[0][1], outs[1][1])
    torch.testing.assert_close(outs[2][1], -2.5 * outs[0][1], rtol=1e-6, atol=0)
    with torch.no_grad():
        assert loss_utils.ssim(a, b).item() == outs[0][0]
    # finite differences along a random direction
    d = torch.randn(3, 480, 640, generator=g).to(DEV)
    h = 1e-2
    num = (loss_utils.ssim(a + h * d, b).item() - loss_utils.ssim(a - h * d, b).item()) / (2 * h)
    ana = (outs[0][1] * d).sum().item()
    assert abs(num - ana) < 2e-2 * abs(ana) + 1e-6


def test_errors():
    a = torch.rand(3, 8, 8, device=DEV)
    with pytest.raises(RuntimeError, match="window"):
        loss_utils.ssim(a, a, window_size=13)
    with pytest.raises(RuntimeError, match="window"):
        loss_utils.ssim(a, a, window_size=4)
    with pytest.raises(RuntimeError, match="same"):
        loss_utils.ssim(a, a[:, :4])
    with pytest.raises(RuntimeError, match="CUDA"):
        loss_utils.ssim(a.cpu(), a.cpu())
    with pytest.raises(IndexError):
        loss_utils.ssim(a, a, size_average=False)


def test_fused_exposure_matches_the_composition():
    """ssim(img, gt, exposure=(a, b)) == ssim(exp(a) * img + b, gt) incl. the gradients w.r.t. img, a and b (the
    composition runs the affine map and its backward in torch around the same kernels)."""
    g = torch.Generator().manual_seed(9)
    x0, y = torch.rand(3, 70, 90, generator=g).to(DEV), torch.rand(3, 70, 90, generator=g).to(DEV)
    outs = []
    for fused in (False, True):
        x = x0.clone().requires_grad_(True)
        a = torch.tensor([0.11], device=DEV, requires_grad=True)
        b = torch.tensor([-0.03], device=DEV, requires_grad=True)
        v = loss_utils.ssim(x, y, exposure=(a, b)) if fused else loss_utils.ssim(torch.exp(a) * x + b, y)
        (3.0 * v).backward()
        outs.append((v.item(), x.grad.clone(), a.grad.item(), b.grad.item()))
    assert abs(outs[0][0] - outs[1][0]) < 1e-6
    torch.testing.assert_close(outs[1][1], outs[0][1], rtol=1e-4, atol=1e-9)
    assert abs(outs[0][2] - outs[1][2]) < 1e-4 * abs(outs[0][2]) + 1e-7
    assert abs(outs[0][3] - outs[1][3]) < 1e-4 * abs(outs[0][3]) + 1e-7
    lum0, con0, stu0 = loss_utils.compute_ssim_components(y, float(np.exp(0.11)) * x0 - 0.03, 7)
    lum1, con1, stu1 = loss_utils.compute_ssim_components(y, x0, 7, exposure=(torch.tensor([0.11]), torch.tensor([-0.03])))
    torch.testing.assert_close(lum1, lum0, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(con1, con0, rtol=1e-4, atol=1e-4)
    assert ((stu1 - stu0).abs() > 1e-3).float().mean().item() < 1e-3
