"""GPU: the mapper's loss functions (loss.get_loss_mapping with the SSIM term, loss.get_loss_mapping_uncertainty)
against vectors produced by the reference's own functions (tests/golden/make_mapping_loss_golden.py)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from wildgs_slam_blackwell_b200 import loss as fused  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
CONFIG = {"Training": {"rgb_boundary_threshold": 0.01, "alpha": 0.5, "ssim_loss": True},
          "opt_params": {"lambda_dssim": 0.2}, "full_resolution": False,
          "uncertainty_params": {"ssim_window_size": 7, "ssim_median_filter_size": 5, "opacity_th_for_uncer_loss": 0.9,
                                 "ssim_mult": 0.5, "uncer_depth_mult": 0.2}}


class View:
    pass


def make_net(seed, fd=12):
    torch.manual_seed(seed)          # same construction order as the generator script -> same weights
    return torch.nn.Sequential(torch.nn.Linear(fd, 16), torch.nn.ReLU(), torch.nn.Linear(16, 1), torch.nn.Softplus(),
                               torch.nn.Flatten(-2, -1))


def setup(d):
    t = lambda k: torch.from_numpy(d[k]).to(DEV)
    image, depth = t("image").requires_grad_(True), t("depth").requires_grad_(True)
    vp = View()
    vp.original_image, vp.depth, vp.features = t("gt"), d["gt_depth"], torch.from_numpy(d["features"])
    vp.exposure_a, vp.exposure_b = t("a").requires_grad_(True), t("b").requires_grad_(True)
    return image, depth, vp


def close(name, got, want, rtol=2e-3, atol_scale=2e-4):
    got = np.zeros_like(want) if got is None else got.detach().cpu().numpy()
    np.testing.assert_allclose(got, want, rtol=rtol, atol=atol_scale * max(np.abs(want).max(), 1e-12), err_msg=name)


def test_get_loss_mapping_with_ssim_matches_reference():
    d = np.load(os.path.join(ROOT, "tests", "golden", "mapping_ssim.npz"))
    image, depth, vp = setup(d)
    loss = fused.get_loss_mapping(CONFIG, image, depth, vp)
    np.testing.assert_allclose(loss.item(), d["loss"], rtol=1e-5)
    loss.backward()
    close("image", image.grad, d["g_image"])
    close("depth", depth.grad, d["g_depth"])
    close("a", vp.exposure_a.grad, d["g_a"])
    close("b", vp.exposure_b.grad, d["g_b"])


@pytest.mark.parametrize("name,tf,sf,init,freeze", [("mapping_uncertainty", 0.3, 0.6, False, False),
                                                    ("mapping_uncertainty_init", 0.9, 0.1, True, True)])
def test_get_loss_mapping_uncertainty_matches_reference(name, tf, sf, init, freeze):
    d = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    image, depth, vp = setup(d)
    net = make_net(7).to(DEV)
    opacity = torch.from_numpy(d["opacity"]).to(DEV)
    unc, loss = fused.get_loss_mapping_uncertainty(CONFIG, image, depth, vp, opacity, net, tf, sf, initialization=init,
                                                   freeze_uncertainty_loss=freeze)
    np.testing.assert_allclose(unc.detach().cpu().numpy(), d["uncertainty"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(loss.item(), d["loss"], rtol=2e-5)
    loss.backward()
    close("image", image.grad, d["g_image"])
    close("depth", depth.grad, d["g_depth"])
    close("a", vp.exposure_a.grad, d["g_a"])
    close("b", vp.exposure_b.grad, d["g_b"])
    for i, p in enumerate(net.parameters()):
        close(f"net{i}", p.grad, d[f"g_net{i}"], rtol=5e-3, atol_scale=1e-3)
    # second call: the cached depth threshold / device depth give the same result
    image2, depth2, _ = setup(d)
    _, loss2 = fused.get_loss_mapping_uncertainty(CONFIG, image2, depth2, vp, opacity, net, tf, sf, initialization=init,
                                                  freeze_uncertainty_loss=freeze)
    assert loss2.item() == loss.item()
