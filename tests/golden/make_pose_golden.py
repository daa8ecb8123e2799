#!/usr/bin/env python
"""Golden vectors for the fused pose step, produced by the REFERENCE's own update_pose
(/root/reference/src/utils/pose_utils.py) driven by torch.optim.Adam exactly as the pose-refinement loop
does (src/mapper.py:861-907), torch CPU float32:   python tests/golden/make_pose_golden.py
Writes tests/golden/pose_steps.npz: initial R, T, exposure; per-iteration gradients; per-iteration R, T,
exposure, converged after pose_optimizer.step() + update_pose()."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from src.utils.pose_utils import SE3_exp, update_pose  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


class Cam:
    def update_RT(self, R, t):
        self.R, self.T = R, t


def main():
    g = torch.Generator().manual_seed(7)
    cam = Cam()
    w2c = SE3_exp(torch.tensor([0.3, -0.2, 0.5, 0.1, -0.25, 0.15]))
    cam.R, cam.T = w2c[:3, :3].clone(), w2c[:3, 3].clone()
    cam.cam_rot_delta = torch.zeros(3, requires_grad=True)
    cam.cam_trans_delta = torch.zeros(3, requires_grad=True)
    cam.exposure_a = torch.tensor([0.0], requires_grad=True)
    cam.exposure_b = torch.tensor([0.0], requires_grad=True)
    lrs = dict(rot=0.003, trans=0.001, a=0.01, b=0.01)
    opt = torch.optim.Adam([{"params": [cam.cam_rot_delta], "lr": lrs["rot"]}, {"params": [cam.cam_trans_delta], "lr": lrs["trans"]},
                            {"params": [cam.exposure_a], "lr": lrs["a"]}, {"params": [cam.exposure_b], "lr": lrs["b"]}])
    out = dict(R0=cam.R.numpy().copy(), T0=cam.T.numpy().copy(), lr=np.array([lrs["rot"], lrs["trans"], lrs["a"], lrs["b"]]))
    n = 14
    grads, Rs, Ts, Es, Cs = [], [], [], [], []
    for it in range(n):
        scale = 10.0 ** (-(it // 3))                  # shrinking gradients; the last ones are exactly zero
        gv = torch.randn(8, generator=g) * scale
        if it >= n - 2:
            gv = torch.zeros(8)
        cam.cam_rot_delta.grad, cam.cam_trans_delta.grad = gv[:3].clone(), gv[3:6].clone()
        cam.exposure_a.grad, cam.exposure_b.grad = gv[6:7].clone(), gv[7:8].clone()
        with torch.no_grad():
            opt.step()
            conv = update_pose(cam)
        grads.append(gv.numpy().copy()); Rs.append(cam.R.numpy().copy()); Ts.append(cam.T.numpy().copy())
        Es.append(np.array([cam.exposure_a.item(), cam.exposure_b.item()], np.float32)); Cs.append(bool(conv))
        assert cam.cam_rot_delta.abs().sum() == 0 and cam.cam_trans_delta.abs().sum() == 0
    out.update(grads=np.stack(grads), R=np.stack(Rs), T=np.stack(Ts), exposure=np.stack(Es), converged=np.array(Cs))
    # update_pose alone (no optimiser), incl. the small-angle branch
    taus = np.array([[0.1, -0.2, 0.05, 0.3, 0.2, -0.4], [1e-7, 0, 0, 2e-6, -1e-6, 3e-6], [0, 0, 0, 0, 0, 0],
                     [0.5, 0.5, -0.5, 2.0, -1.0, 0.7]], np.float32)
    Ru, Tu, Cu = [], [], []
    for tau in taus:
        c = Cam()
        c.R, c.T = torch.from_numpy(out["R0"]).clone(), torch.from_numpy(out["T0"]).clone()
        c.cam_trans_delta = torch.tensor(tau[:3], requires_grad=True)
        c.cam_rot_delta = torch.tensor(tau[3:], requires_grad=True)
        with torch.no_grad():
            Cu.append(bool(update_pose(c)))
        Ru.append(c.R.numpy().copy()); Tu.append(c.T.numpy().copy())
    out.update(taus=taus, Ru=np.stack(Ru), Tu=np.stack(Tu), Cu=np.array(Cu))
    np.savez_compressed(os.path.join(HERE, "pose_steps.npz"), **out)
    print("converged flags:", Cs, Cu)


if __name__ == "__main__":
    main()
