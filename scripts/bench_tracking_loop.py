#!/usr/bin/env python
"""Pose refinement of one non-keyframe (Mapper.refine_pose_non_key_frame, src/mapper.py:885-909): 100 iterations of
render -> get_loss_tracking -> backward -> pose_optimizer.step() -> update_pose, config 2 (300 k Gaussians, 640x480).
  reference   : unmodified reference rasterizer (oracle/_ref) + the reference's torch expressions for wrapper, loss,
                torch.optim.Adam and update_pose (restated below from src/utils/pose_utils.py; test infrastructure)
  composition : the same torch expressions on OUR rasterizer
  fused       : renderer.render + fused loss + FusedPoseOptimizer, eager
  fused_graph : the whole iteration as ONE CUDA graph (captured for the first frame, reused for the following
                frames of the sequence), convergence flag polled every 10 iterations
Prints one JSON line: ms per 100-iteration refinement (median of 5), CUDA events + final synchronise."""
import json, os, statistics, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "scripts"))
from test_gpu_render import Cam, Model, Pipe
from bench_iteration import composition, torch_loss_tracking, CONFIG
from wildgs_slam_blackwell_b200 import graph, loss as fused_loss, pose, renderer, scenes
DEV = "cuda:0"
ITERS = 100


def skew(x):
    s = torch.zeros(3, 3, device=x.device)
    s[0, 1], s[0, 2], s[1, 0], s[1, 2], s[2, 0], s[2, 1] = -x[2], x[1], x[2], -x[0], -x[1], x[0]
    return s


def torch_update_pose(cam, thr=1e-4):
    """pose_utils.py:17-100 as torch expressions (includes the host syncs of its `angle < 1e-5` branches)."""
    tau = torch.cat([cam.cam_trans_delta, cam.cam_rot_delta], axis=0)
    rho, theta = tau[:3], tau[3:]
    W = skew(theta); W2 = W @ W; angle = torch.norm(theta); I = torch.eye(3, device=tau.device)
    if angle < 1e-5:
        R = I + W + 0.5 * W2; V = I + 0.5 * W + (1.0 / 6.0) * W2
    else:
        R = I + (torch.sin(angle) / angle) * W + ((1 - torch.cos(angle)) / (angle ** 2)) * W2
        V = I + W * ((1.0 - torch.cos(angle)) / (angle ** 2)) + W2 * ((angle - torch.sin(angle)) / (angle ** 3))
    T = torch.eye(4, device=tau.device); T[:3, :3] = R; T[:3, 3] = V @ rho
    w2c = torch.eye(4, device=tau.device); w2c[:3, :3] = cam.R; w2c[:3, 3] = cam.T
    new = T @ w2c
    converged = tau.norm() < thr
    cam.R, cam.T = new[:3, :3], new[:3, 3]
    cam.cam_rot_delta.data.fill_(0); cam.cam_trans_delta.data.fill_(0)
    return converged


def main():
    P, W, H, deg, _ = scenes.CONFIGS["config2"]
    cam_d = scenes.make_camera(W, H, 1002)
    g = scenes.make_gaussians(P, W, H, deg, 1002, cam_d)
    bg = torch.zeros(3, device=DEV)
    pc = Model(g, 1)
    with torch.no_grad():
        target = renderer.render(Cam(cam_d), pc, Pipe(), bg)["render"].clone()
    start = scenes.make_camera(W, H, 1002, tau=np.array([0.02, -0.01, 0.015, 0.01, -0.008, 0.005]))
    arms = os.environ.get("ARMS", "composition,fused,fused_graph").split(",")
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if os.path.exists(os.path.join(ref_dir, "diff_gaussian_rasterization", "_C.so")) and "ARMS" not in os.environ:
        arms.append("reference")
    res = {}
    for arm in arms:
        if arm == "reference":
            import importlib
            for k in [k for k in sys.modules if k.startswith("diff_gaussian_rasterization")]:
                del sys.modules[k]
            sys.path.insert(0, ref_dir)
            mod = importlib.import_module("diff_gaussian_rasterization")
        else:
            import wildgs_slam_blackwell_b200.rasterizer as mod
        times, final_loss = [], None
        pc = Model(g, 1)     # fresh leaves per arm (their grad accumulators are bound to the stream of first use)
        it = None
        for rep in range(6):
            if arm == "fused_graph" and it is not None:
                # steady state of a sequence: the iteration captured for the first frame is reused; the new frame's
                # initial pose and exposure are written into the captured buffers
                torch.cuda.synchronize(); t0 = time.perf_counter()
                with torch.no_grad():
                    cam.exposure_a.zero_(); cam.exposure_b.zero_()
                opt.reset(R0, T0)
                for i in range(ITERS):
                    l = it.replay()
                    if i % 10 == 9 and opt.converged.item():
                        break
                it.check()
                torch.cuda.synchronize()
                times.append((time.perf_counter() - t0) * 1e3)
                final_loss = float(l.item())
                continue
            cam = Cam(start)
            cam.R, cam.T = cam.R.contiguous(), cam.T.contiguous()
            R0, T0 = cam.R.clone(), cam.T.clone()
            cam.original_image, cam.grad_mask = target, torch.ones(1, H, W, dtype=torch.bool, device=DEV)
            cam.exposure_a = torch.zeros(1, device=DEV, requires_grad=True)
            cam.exposure_b = torch.zeros(1, device=DEV, requires_grad=True)
            torch.cuda.synchronize(); t0 = time.perf_counter()
            if arm in ("reference", "composition"):
                opt = torch.optim.Adam([{"params": [cam.cam_rot_delta], "lr": 0.003}, {"params": [cam.cam_trans_delta], "lr": 0.001},
                                        {"params": [cam.exposure_a], "lr": 0.01}, {"params": [cam.exposure_b], "lr": 0.01}])
                for _ in range(ITERS):
                    img, depth, opacity = composition(mod, cam, pc, bg)
                    opt.zero_grad()
                    l = torch_loss_tracking(img, opacity, cam)
                    l.backward()
                    with torch.no_grad():
                        opt.step()
                        converged = torch_update_pose(cam, thr=0.0)
                    if converged:
                        break
            else:
                opt = pose.FusedPoseOptimizer(cam, lr_rot=0.003, lr_trans=0.001, converged_threshold=0.0)

                def step():
                    opt.zero_grad()
                    pkg = renderer.render(cam, pc, Pipe(), bg)
                    l = fused_loss.get_loss_tracking(CONFIG, pkg["render"], pkg["depth"], pkg["opacity"], cam)
                    l.backward()
                    opt.step()
                    return l.detach()
                if arm == "fused":
                    for _ in range(ITERS):
                        l = step()
                else:
                    it = graph.CapturedIteration(step, DEV, warmup=2)       # first frame: capture (not in the median)
                    for i in range(ITERS - 2):
                        l = it.replay()
                    it.check()
            torch.cuda.synchronize()
            times.append((time.perf_counter() - t0) * 1e3)
            final_loss = float(l.item())
        res[arm] = {"ms_per_100_iterations": round(statistics.median(times[1:]), 2), "final_loss": round(final_loss, 6)}
    print(json.dumps({"what": "pose refinement, 100 iterations, config2 (wall clock incl. all host work)", **res}))


if __name__ == "__main__":
    main()
