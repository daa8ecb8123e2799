"""GPU: the pieces either side of the rasterizer working together the way the mapper drives them
(src/mapper.py map_opt_online): render() in raw-parameter mode -> fused get_loss_mapping -> backward ->
FusedAdam.step(), over a small keyframe window, with a densification step in the middle (Gaussians cloned,
optimizer state extended in place like gaussian_model.py cat_tensors_to_optimizer).  The fit must improve
and stay finite; the window gradient of the first iteration must equal the torch-composed one."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from test_gpu_render import Cam, Model, Pipe, reference_composition  # noqa: E402
from wildgs_slam_blackwell_b200 import loss as fused_loss, renderer, scenes  # noqa: E402
from wildgs_slam_blackwell_b200.adam import FusedAdam  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
CFG = {"Training": {"rgb_boundary_threshold": 0.01, "alpha": 0.9, "ssim_loss": False}}
NAMES = ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation")
LRS = (1.6e-4, 2.5e-3, 1.25e-4, 0.05, 1e-3, 1e-3)


def _optimizer(pc):
    return FusedAdam([{"params": [getattr(pc, n)], "lr": lr, "name": n} for n, lr in zip(NAMES, LRS)], lr=0.0, eps=1e-15)


def _densify(pc, opt, n_new):
    """clone the first n_new Gaussians (slightly displaced); extend parameters and Adam state in place"""
    for group, name in zip(opt.param_groups, NAMES):
        old = group["params"][0]
        ext = old.detach()[:n_new].clone()
        if name == "_xyz":
            ext += 0.01
        st = opt.state.pop(old, None)
        new = torch.nn.Parameter(torch.cat((old.detach(), ext)).contiguous().requires_grad_(True))
        if st is not None:
            st["exp_avg"] = torch.cat((st["exp_avg"], torch.zeros_like(ext)))
            st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"], torch.zeros_like(ext)))
            opt.state[new] = st
        group["params"][0] = new
        setattr(pc, name, new)


def test_mapping_loop_with_densification_improves_the_fit():
    W, H, P, n_views = 160, 112, 12000, 3
    cam0 = scenes.make_camera(W, H, 50)
    g = scenes.make_gaussians(P, W, H, 1, 50, cam0)
    bg = torch.zeros(3, device=DEV)
    cams = [Cam(scenes.make_camera(W, H, 50 + v)) for v in range(n_views)]
    # ground truth: renders of the unperturbed model
    truth = Model(g, 50)
    with torch.no_grad():
        for cam in cams:
            pkg = renderer.render(cam, truth, Pipe(), bg)
            cam.original_image = pkg["render"].clone()
            cam.depth = pkg["depth"][0].cpu().numpy()
            cam.exposure_a = torch.zeros(1, device=DEV, requires_grad=True)
            cam.exposure_b = torch.zeros(1, device=DEV, requires_grad=True)
    rng = np.random.default_rng(0)
    g2 = dict(g)
    g2["sh"] = g["sh"] + rng.normal(0, 0.3, g["sh"].shape).astype(np.float32)
    g2["means3D"] = g["means3D"] + rng.normal(0, 0.01, g["means3D"].shape).astype(np.float32)
    pc = Model(g2, 51)
    opt = _optimizer(pc)

    def window_loss_and_backward(render_fn, loss_fn):
        total = 0.0
        for cam in cams:
            pkg = render_fn(cam)
            l = loss_fn(pkg, cam)
            l.backward()
            total += l.item()
        return total / n_views

    fused_render = lambda cam: renderer.render(cam, pc, Pipe(), bg)
    fused = lambda pkg, cam: fused_loss.get_loss_mapping(CFG, pkg["render"], pkg["depth"], cam)

    # first iteration: fused window gradient == torch-composed window gradient
    window_loss_and_backward(fused_render, fused)
    g_fused = {n: getattr(pc, n).grad.clone() for n in NAMES}
    for n in NAMES:
        getattr(pc, n).grad = None

    def torch_loss(pkg, cam):
        gt, gd = cam.original_image, torch.from_numpy(cam.depth).to(DEV)[None]
        image = torch.exp(cam.exposure_a) * pkg["render"] + cam.exposure_b
        m = (gt.sum(dim=0) > 0.01).view(1, H, W)
        dm = (gd > 0.01).view(1, H, W)
        return 0.9 * torch.abs(image * m - gt * m).mean() + 0.1 * torch.abs(pkg["depth"] * dm - gd * dm).mean()

    window_loss_and_backward(lambda cam: reference_composition(cam, pc, bg), torch_loss)
    for n in NAMES:
        a, b = g_fused[n].cpu().numpy().astype(np.float64), getattr(pc, n).grad.cpu().numpy().astype(np.float64)
        err, scale = np.abs(a - b), np.abs(b).max()
        assert (err > 2e-3 * np.abs(b) + 2e-4 * scale).mean() < 1e-3 and err.max() < 5e-2 * scale, n
        getattr(pc, n).grad = None

    losses = []
    for it in range(16):
        opt.zero_grad(set_to_none=True)
        losses.append(window_loss_and_backward(fused_render, fused))
        opt.step()
        if it == 7:
            _densify(pc, opt, 1500)
            assert pc._xyz.shape[0] == P + 1500
    assert all(np.isfinite(losses)), losses
    assert losses[-1] < 0.8 * losses[0], losses
    for n in NAMES:
        assert torch.isfinite(getattr(pc, n)).all(), n
