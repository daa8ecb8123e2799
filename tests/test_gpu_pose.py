"""GPU: the fused pose step (Adam on the pose/exposure deltas + update_pose + camera matrices, one launch)
against golden vectors produced by the REFERENCE's own update_pose driven by torch.optim.Adam
(tests/golden/make_pose_golden.py), and the whole tracking iteration incl. the pose step as one CUDA graph."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wildgs_slam_blackwell_b200 import pose, renderer  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
G = dict(np.load(os.path.join(ROOT, "tests", "golden", "pose_steps.npz")))


class View:
    pass


def camera(R, T):
    c = View()
    c.R, c.T = torch.from_numpy(R.copy()).to(DEV), torch.from_numpy(T.copy()).to(DEV)
    c.cam_rot_delta = torch.zeros(3, device=DEV, requires_grad=True)
    c.cam_trans_delta = torch.zeros(3, device=DEV, requires_grad=True)
    c.exposure_a = torch.zeros(1, device=DEV, requires_grad=True)
    c.exposure_b = torch.zeros(1, device=DEV, requires_grad=True)
    c.projection_matrix = torch.eye(4, device=DEV) + 0.1 * torch.arange(16.0, device=DEV).view(4, 4)
    return c


def test_optimizer_step_plus_update_pose_matches_reference_golden():
    cam = camera(G["R0"], G["T0"])
    lr = G["lr"]
    opt = pose.FusedPoseOptimizer(cam, lr_rot=lr[0], lr_trans=lr[1], lr_exposure_a=lr[2], lr_exposure_b=lr[3])
    renderer._cached_matrices(cam)
    for it, gv in enumerate(G["grads"]):
        g = torch.from_numpy(gv).to(DEV)
        cam.cam_rot_delta.grad, cam.cam_trans_delta.grad = g[:3].clone(), g[3:6].clone()
        cam.exposure_a.grad, cam.exposure_b.grad = g[6:7].clone(), g[7:8].clone()
        opt.step()
        np.testing.assert_allclose(cam.R.cpu().numpy(), G["R"][it], rtol=2e-6, atol=2e-7, err_msg=f"R it {it}")
        np.testing.assert_allclose(cam.T.cpu().numpy(), G["T"][it], rtol=2e-6, atol=2e-7, err_msg=f"T it {it}")
        e = np.array([cam.exposure_a.item(), cam.exposure_b.item()])
        np.testing.assert_allclose(e, G["exposure"][it], rtol=2e-6, atol=1e-8, err_msg=f"exposure it {it}")
        assert bool(opt.converged.item()) == bool(G["converged"][it])
        assert not cam.cam_rot_delta.any() and not cam.cam_trans_delta.any()
        # the matrices render() will use were refreshed in place by the same launch, and the cache is current
        v, p, c = renderer._cached_matrices(cam)
        v2, p2, c2 = renderer.camera_matrices(cam.R, cam.T, cam.projection_matrix)
        assert torch.equal(v, v2) and torch.allclose(p, p2, rtol=1e-6, atol=1e-6) and torch.allclose(c, c2, rtol=1e-6, atol=1e-7)
    assert int(opt.step_count.item()) == len(G["grads"])


def test_update_pose_alone_matches_reference_golden():
    for tau, R, T, conv in zip(G["taus"], G["Ru"], G["Tu"], G["Cu"]):
        cam = camera(G["R0"], G["T0"])
        with torch.no_grad():
            cam.cam_trans_delta.copy_(torch.from_numpy(tau[:3]))
            cam.cam_rot_delta.copy_(torch.from_numpy(tau[3:]))
        v0 = cam.R._version
        converged = pose.update_pose(cam)
        assert bool(converged) == bool(conv)
        np.testing.assert_allclose(cam.R.cpu().numpy(), R, rtol=2e-6, atol=2e-7)
        np.testing.assert_allclose(cam.T.cpu().numpy(), T, rtol=2e-6, atol=2e-7)
        assert not cam.cam_rot_delta.any() and not cam.cam_trans_delta.any()
        assert cam.R._version > v0 or not hasattr(torch.autograd.graph, "increment_version")


def test_errors():
    cam = camera(G["R0"], G["T0"])
    cam.R = cam.R.cpu()
    with pytest.raises(RuntimeError, match="in place"):
        pose.update_pose(cam)
    cam = camera(G["R0"], G["T0"])
    with pytest.raises(RuntimeError, match="no gradient"):
        pose.FusedPoseOptimizer(cam, 0.003, 0.001).step()


def test_whole_tracking_iteration_as_one_graph():
    """render -> tracking loss -> backward -> Adam + update_pose, captured once, replayed: same trajectory as eager."""
    from test_gpu_render import Cam, Model, Pipe
    from wildgs_slam_blackwell_b200 import graph, loss as fused_loss, scenes
    W, H, P = 160, 112, 6000
    cam_d = scenes.make_camera(W, H, 31)
    g = scenes.make_gaussians(P, W, H, 0, 31, cam_d)
    cfg = {"Training": {"rgb_boundary_threshold": 0.01}}
    bg = torch.zeros(3, device=DEV)
    pc = Model(g, 31)
    with torch.no_grad():
        target = renderer.render(Cam(cam_d), pc, Pipe(), bg)["render"].clone()

    def run(captured, n=6):
        cam = Cam(scenes.make_camera(W, H, 31, tau=np.array([0.02, -0.01, 0.015, 0.01, -0.008, 0.005])))
        cam.R, cam.T = cam.R.contiguous(), cam.T.contiguous()
        cam.original_image, cam.grad_mask = target, torch.ones(1, H, W, dtype=torch.bool, device=DEV)
        cam.exposure_a = torch.zeros(1, device=DEV, requires_grad=True)
        cam.exposure_b = torch.zeros(1, device=DEV, requires_grad=True)
        opt = pose.FusedPoseOptimizer(cam, lr_rot=0.003, lr_trans=0.001)

        def step():
            opt.zero_grad()
            pkg = renderer.render(cam, pc, Pipe(), bg)
            l = fused_loss.get_loss_tracking(cfg, pkg["render"], pkg["depth"], pkg["opacity"], cam)
            l.backward()
            opt.step()
            return l.detach()

        losses = []
        if captured:
            it = graph.CapturedIteration(step, DEV, warmup=2)      # two eager iterations, then the capture
            for _ in range(n):
                losses.append(it.replay().item())
            it.check()
        else:
            for _ in range(n + 2):
                losses.append(step().item())
            losses = losses[2:]
        return losses, cam.R.clone(), cam.T.clone()

    le, Re, Te = run(False)
    lc, Rc, Tc = run(True)
    np.testing.assert_allclose(lc, le, rtol=1e-4)
    # float-atomic summation order differs from run to run and Adam's normalised step magnifies it (a step is
    # ~lr = 1e-3 whatever the gradient's size): 1e-4 is about 1 % of the distance moved in these 8 iterations
    torch.testing.assert_close(Rc, Re, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(Tc, Te, rtol=1e-4, atol=1e-4)
    assert le[-1] < le[0]          # the pose moves towards the target view
