"""GPU: the fused photometric loss kernel (through the Python mirror of the reference's loss functions and
the C ABI underneath) against the reference's golden vectors and the numpy oracle."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import loss_oracle  # noqa: E402
from wildgs_slam_blackwell_b200 import loss  # noqa: E402
from test_loss_oracle import GOLDEN, KINDS, oracle_for  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
CONFIG = {"Training": {"rgb_boundary_threshold": 0.01, "alpha": 0.5, "ssim_loss": False}}


class View:
    pass


def viewpoint(g):
    vp = View()
    vp.original_image = torch.from_numpy(g["gt"]).to(DEV)
    vp.grad_mask = torch.from_numpy(g["grad_mask"]).to(DEV)
    vp.depth = g["gt_depth"]
    vp.exposure_a = torch.from_numpy(g["a"]).to(DEV).requires_grad_(True)
    vp.exposure_b = torch.from_numpy(g["b"]).to(DEV).requires_grad_(True)
    return vp


@pytest.mark.parametrize("kind", KINDS)
def test_reference_api_matches_reference_golden(kind):
    """Same call the reference's mapper makes (src/mapper.py:896-904, 1117), loss.backward() included."""
    g = dict(np.load(os.path.join(GOLDEN, f"loss_{kind}.npz")))
    vp = viewpoint(g)
    image = torch.from_numpy(g["image"]).to(DEV).requires_grad_(True)
    depth = torch.from_numpy(g["depth"]).to(DEV).requires_grad_(True)
    opacity = torch.from_numpy(g["opacity"]).to(DEV)
    unc = torch.from_numpy(g["unc"]).to(DEV)
    if kind == "tracking":
        l = loss.get_loss_tracking(CONFIG, image, depth, opacity, vp)
    elif kind == "tracking_unc":
        l = loss.get_loss_tracking(CONFIG, image, depth, opacity, vp, uncertainty=unc)
    elif kind == "mapping":
        l = loss.get_loss_mapping(CONFIG, image, depth, vp)
    else:
        l = loss.get_loss_mapping(CONFIG, image, depth, vp, initialization=True)
    l.backward()
    np.testing.assert_allclose(l.item(), g["loss"], rtol=2e-6)
    np.testing.assert_allclose(image.grad.cpu().numpy(), g["g_image"], rtol=1e-5, atol=1e-10)
    if kind.startswith("mapping"):
        np.testing.assert_allclose(depth.grad.cpu().numpy(), g["g_depth"], rtol=1e-5, atol=1e-10)
    else:
        assert depth.grad is None
    if kind != "mapping_init":
        np.testing.assert_allclose(vp.exposure_a.grad.cpu().numpy(), g["g_a"], rtol=2e-4)
        np.testing.assert_allclose(vp.exposure_b.grad.cpu().numpy(), g["g_b"], rtol=2e-4, atol=1e-7)
    else:
        assert vp.exposure_a.grad is None


@pytest.mark.parametrize("H,W", [(480, 640), (1, 1), (3, 5), (1080, 1920)])
def test_raw_call_matches_oracle_and_is_deterministic(H, W):
    rng = np.random.default_rng(H * 7 + W)
    f = lambda *s: rng.uniform(0, 1, s).astype(np.float32)
    a = dict(color=f(3, H, W), gt=f(3, H, W), depth=f(1, H, W) * 5, gt_depth=f(1, H, W) * 5 - 0.5, opacity=f(1, H, W),
             mask=(f(1, H, W) > 0.2).astype(np.float32), unc=0.1 + 3 * f(H, W), ea=np.float32([0.2]), eb=np.float32([0.05]))
    t = {k: torch.from_numpy(v).to(DEV) for k, v in a.items()}
    kw = dict(depth=t["depth"], gt_depth=t["gt_depth"], opacity=t["opacity"], pixel_mask=t["mask"], uncertainty=t["unc"],
              exposure_a=t["ea"], exposure_b=t["eb"], rgb_boundary_threshold=0.3, color_weight=0.9, depth_weight=0.1)
    l1, dc1, dd1, de1 = [x.clone() for x in loss.photometric_loss(t["color"], t["gt"], **kw)]
    l2, dc2, dd2, de2 = loss.photometric_loss(t["color"], t["gt"], **kw)
    assert torch.equal(l1, l2) and torch.equal(dc1, dc2) and torch.equal(dd1, dd2) and torch.equal(de1, de2)
    o = loss_oracle.photometric_loss(a["color"], a["gt"], depth=a["depth"], gt_depth=a["gt_depth"], opacity=a["opacity"],
                                     pixel_mask=a["mask"], uncertainty=a["unc"], exposure_a=a["ea"], exposure_b=a["eb"],
                                     rgb_boundary_threshold=0.3, color_weight=0.9, depth_weight=0.1)
    np.testing.assert_allclose(l1.item(), o["loss"], rtol=5e-6)
    np.testing.assert_allclose(dc1.cpu().numpy(), o["dL_dcolor"], rtol=2e-5, atol=1e-12)
    np.testing.assert_allclose(dd1.cpu().numpy(), o["dL_ddepth"], rtol=2e-5, atol=1e-12)
    np.testing.assert_allclose(de1.cpu().numpy(), [o["dL_da"], o["dL_db"]], rtol=2e-4, atol=1e-7)


def test_argument_errors():
    c = torch.zeros(3, 8, 8, device=DEV)
    with pytest.raises(RuntimeError, match="depth_weight"):
        loss.photometric_loss(c, c, depth_weight=0.5)
    with pytest.raises(RuntimeError, match="go together"):
        loss.photometric_loss(c, c, exposure_a=torch.zeros(1, device=DEV))
    with pytest.raises(RuntimeError, match=r"\(3, H, W\)"):
        loss.photometric_loss(torch.zeros(8, 8, device=DEV), c)
    with pytest.raises(KeyError):      # ssim_loss on needs opt_params.lambda_dssim, as in the reference (slam_utils.py:135)
        loss.get_loss_mapping({"Training": {"rgb_boundary_threshold": 0.01, "ssim_loss": True}}, c, c[:1], View(),
                              initialization=True)
    with pytest.raises(RuntimeError, match="depth_weighted needs"):
        loss.photometric_loss(c, c, depth=c[:1], gt_depth=c[:1], depth_weight=0.5, depth_weighted=True)
