#!/usr/bin/env python
"""Tracking-iteration time (config 2: 300 k Gaussians, 640x480): render() + get_loss_tracking + backward,
the loop body of Mapper.refine_pose_non_key_frame (src/mapper.py:886-904), in three arrangements:
  composition : the reference's wrapper arrangement (torch activations, torch camera properties, torch loss
                expression + autograd) on OUR rasterizer
  fused       : renderer.render (raw-parameter mode + camera kernel) + fused loss kernel
  fused_graph : the fused iteration captured once into a CUDA graph and replayed (graph.CapturedIteration)
  reference   : the composition on the UNMODIFIED reference rasterizer (oracle/_ref), when built
Prints one JSON line (ms per iteration, CUDA events, L2 flushed between iterations)."""
import json
import math
import os
import statistics
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_render import Cam, Model, Pipe  # noqa: E402
from wildgs_slam_blackwell_b200 import loss as fused_loss, renderer, scenes  # noqa: E402

DEV = "cuda:0"
CONFIG = {"Training": {"rgb_boundary_threshold": 0.01}}


def torch_loss_tracking(image, opacity, vp):
    """get_loss_tracking + get_loss_tracking_rgb as torch expressions (slam_utils.py:47-87)."""
    image_ab = torch.exp(vp.exposure_a) * image + vp.exposure_b
    gt = vp.original_image
    m = (gt.sum(dim=0) > CONFIG["Training"]["rgb_boundary_threshold"]).view(1, *gt.shape[1:]) * vp.grad_mask
    return (opacity * torch.abs(image_ab * m - gt * m)).mean()


def composition(mod, cam, pc, bg):
    import math
    sp = torch.zeros_like(pc.get_xyz, requires_grad=True) + 0
    s = mod.GaussianRasterizationSettings(
        image_height=cam.image_height, image_width=cam.image_width, tanfovx=math.tan(cam.FoVx * 0.5),
        tanfovy=math.tan(cam.FoVy * 0.5), bg=bg, scale_modifier=1.0, viewmatrix=cam.world_view_transform,
        projmatrix=cam.full_proj_transform, projmatrix_raw=cam.projection_matrix, sh_degree=pc.active_sh_degree,
        campos=cam.camera_center, prefiltered=False, debug=False)
    img, radii, depth, opacity, n_touched = mod.GaussianRasterizer(s)(
        means3D=pc.get_xyz, means2D=sp, shs=pc.get_features, colors_precomp=None, opacities=pc.get_opacity,
        scales=pc.get_scaling, rotations=pc.get_rotation, cov3D_precomp=None, theta=cam.cam_rot_delta,
        rho=cam.cam_trans_delta)
    return img, depth, opacity


def timeit(fn, flush, warm=10, steps=60):
    for _ in range(warm):
        fn(); flush.add_(1)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        flush.add_(1); a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in ev)


def torch_loss_mapping(image, depth, vp, gt_depth):
    """get_loss_mapping + get_loss_mapping_rgbd as torch expressions (slam_utils.py:107-143, ssim off)."""
    image_ab = torch.exp(vp.exposure_a) * image + vp.exposure_b
    gt = vp.original_image
    m = (gt.sum(dim=0) > CONFIG["Training"]["rgb_boundary_threshold"]).view(1, *gt.shape[1:])
    dm = (gt_depth > 0.01).view(*depth.shape)
    return 0.95 * torch.abs(image_ab * m - gt * m).mean() + 0.05 * torch.abs(depth * dm - gt_depth * dm).mean()


UCFG = {"Training": {"rgb_boundary_threshold": 0.01, "alpha": 0.5, "ssim_loss": True}, "opt_params": {"lambda_dssim": 0.2},
        "full_resolution": False,
        "uncertainty_params": {"ssim_window_size": 7, "ssim_median_filter_size": 5, "opacity_th_for_uncer_loss": 0.9,
                               "ssim_mult": 0.5, "uncer_depth_mult": 0.2}}


def _win(ws, ch, dev):
    gs = torch.tensor([math.exp(-((x - ws // 2) ** 2) / float(2 * 1.5 ** 2)) for x in range(ws)])
    gs = (gs / gs.sum()).unsqueeze(1)
    return gs.mm(gs.t()).float()[None, None].expand(ch, 1, ws, ws).contiguous().to(dev)


def _moments(a, b, ws):
    import torch.nn.functional as F
    w = _win(ws, a.size(-3), a.device)
    c = lambda t: F.conv2d(t, w, padding=ws // 2, groups=a.size(-3))
    mu1, mu2 = c(a), c(b)
    return mu1, mu2, c(a * a) - mu1.pow(2), c(b * b) - mu2.pow(2), c(a * b) - mu1 * mu2


def torch_loss_mapping_uncertainty(image, depth, vp, gt_depth, opacity, net, train_frac=0.3, ssim_frac=0.6):
    """get_loss_mapping_uncertainty + compute_mapping_loss_components as the torch expressions the reference runs
    (slam_utils.py:146-258, mapping_utils.py:125-323, loss_utils.py:79-101) -- the baseline arm's loss."""
    import torch.nn.functional as F
    tr, u = UCFG["Training"], UCFG["uncertainty_params"]
    bias = lambda x, s: x / (1 + (1 - x) * (1 / s - 2))
    rs = lambda t, shape, mode="bilinear": F.interpolate(t.view((1, 1) + t.shape[-2:]), size=shape, mode=mode)[0, 0]
    image_ab = torch.exp(vp.exposure_a) * image + vp.exposure_b
    gt = vp.original_image
    _, h, w = gt.shape
    mask = (gt.sum(dim=0) > tr["rgb_boundary_threshold"]).view(1, h, w)
    mu1, mu2, s1, s2, s12 = _moments(image_ab, gt, 11)
    ssim_loss = 1.0 - (((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1.pow(2) + mu2.pow(2) + 1e-4) * (s1 + s2 + 9e-4))).mean()
    unc = net(vp.features)
    l1_rgb = torch.abs(image_ab * mask - gt * mask)
    thr = min(10 * gt_depth.median(), 50)
    dmask = ((gt_depth > 0.01) & (gt_depth < thr)).view(*depth.shape)
    l1_depth = torch.abs(depth * dmask - gt_depth * dmask)
    proc = torch.clip(unc, min=0.1) + 1e-3
    resized = (rs(proc.detach(), (h, w)) - 0.1) * (1 + bias(train_frac, 0.8)) + 0.1
    op = opacity.detach().view(h, w)
    small_op = rs(op, unc.shape)
    m1, m2, v1, v2, v12 = _moments(gt[None], image_ab[None], u["ssim_window_size"])
    eps = torch.tensor([torch.finfo(torch.float32).eps], device=gt.device)
    v1, v2 = torch.maximum(eps, v1), torch.maximum(eps, v2)
    v12 = torch.sign(v12) * torch.minimum(torch.sqrt(v1 * v2), torch.abs(v12))
    lum = ((2 * m1 * m2 + 1e-4) / (m1.pow(2) + m2.pow(2) + 1e-4)).mean(1).squeeze()
    con = torch.clamp((2 * torch.sqrt(v1) * torch.sqrt(v2) + 9e-4) / (v1 + v2 + 9e-4), max=0.98).mean(1).squeeze()
    stu = torch.clamp((v12 + 4.5e-4) / (torch.sqrt(v1) * torch.sqrt(v2) + 4.5e-4), max=0.98).mean(1).squeeze()
    sl = torch.clip(op * (100 + 900 * bias(ssim_frac, 0.8)) * (1 - lum) * (1 - stu) * (1 - con), max=5.0)
    sm = rs(sl.detach(), unc.shape)
    k = u["ssim_median_filter_size"]
    pad = F.pad(sm[None, None], (k // 2, k // 2, k // 2, k // 2), mode="reflect")
    filt = pad.unfold(2, k, 1).unfold(3, k, 1).contiguous().view(unc.shape + (-1,)).median(dim=-1)[0]
    sdl = rs(torch.clip(l1_depth.squeeze(), max=5.0).detach(), unc.shape, "bicubic")
    sdl[rs(gt_depth.squeeze().detach(), unc.shape, "bicubic") > thr] = 0.0
    ul = filt / proc ** 2 + 0.5 * torch.log(proc) + u["uncer_depth_mult"] * sdl / proc ** 2
    ul[small_op < u["opacity_th_for_uncer_loss"]] = 0
    lam = UCFG["opt_params"]["lambda_dssim"]
    wts = 0.5 / resized.unsqueeze(0) ** 2
    wts = torch.where(wts < 0.1, 0.0, wts)
    rgb_loss = wts * ((1.0 - lam) * l1_rgb + lam * ssim_loss)
    near = gt_depth < depth.detach() + 1.0
    l1_depth[near] = wts[near] * l1_depth[near]
    return tr["alpha"] * rgb_loss.mean() + (1 - tr["alpha"]) * l1_depth.mean() + u["ssim_mult"] * ul.mean()


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "tracking"
    mapping = mode in ("mapping", "mapping_uncertainty")
    wild = mode == "mapping_uncertainty"
    P, W, H, deg, _ = scenes.CONFIGS["config3" if mapping else "config2"]
    cam_d = scenes.make_camera(W, H, 1002)
    g = scenes.make_gaussians(P, W, H, deg, 1002, cam_d)
    rng = np.random.default_rng(3)
    bg = torch.zeros(3, device=DEV)
    flush = torch.zeros((512 << 20) // 4, device=DEV)
    res = {}
    arms = ["composition", "fused"]
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if os.path.exists(os.path.join(ref_dir, "diff_gaussian_rasterization", "_C.so")):
        arms.append("reference")
    if os.environ.get("ARMS"):
        arms = [a for a in arms if a in os.environ["ARMS"].split(",")]
    for arm in arms:
        pc, cam = Model(g, 1), Cam(cam_d)
        cam.original_image = torch.from_numpy(rng.uniform(0, 1, (3, H, W)).astype(np.float32)).to(DEV)
        cam.grad_mask = torch.from_numpy(rng.uniform(0, 1, (1, H, W)) > 0.3).to(DEV)
        cam.exposure_a = torch.zeros(1, device=DEV, requires_grad=True)
        cam.exposure_b = torch.zeros(1, device=DEV, requires_grad=True)
        leaves = list(pc.leaves().values()) + [cam.cam_rot_delta, cam.cam_trans_delta, cam.exposure_a, cam.exposure_b]
        opt = None
        if mapping:   # one keyframe per iteration + Adam over the Gaussian parameters, as map_opt_online does
            cam.depth = rng.uniform(0.5, 6.0, (H, W)).astype(np.float32)
            gt_depth = torch.from_numpy(cam.depth).to(DEV)[None]
            if wild:  # DINO-like patch features and the uncertainty MLP (stand-ins of the real sizes: 34x45x384)
                torch.manual_seed(5)
                cam.features = torch.randn(H // 14, W // 14, 384, device=DEV)
                net = torch.nn.Sequential(torch.nn.Linear(384, 64), torch.nn.ReLU(), torch.nn.Linear(64, 1),
                                          torch.nn.Softplus(), torch.nn.Flatten(-2, -1)).to(DEV)
                leaves = leaves + list(net.parameters())
            groups = [{"params": [t], "lr": lr} for t, lr in zip(pc.leaves().values(), (1.6e-4, 2.5e-3, 1.25e-4, 0.05, 1e-3, 1e-3))]
            if arm == "fused":
                from wildgs_slam_blackwell_b200.adam import FusedAdam
                opt = FusedAdam(groups, lr=0.0, eps=1e-15, capturable=True)   # device-side step count: graph-replayable
            else:
                opt = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
        if arm == "reference":
            import importlib
            for k in [k for k in sys.modules if k.startswith("diff_gaussian_rasterization")]:
                del sys.modules[k]
            sys.path.insert(0, ref_dir)
            mod = importlib.import_module("diff_gaussian_rasterization")
            assert os.path.abspath(mod.__file__).startswith(ref_dir)
        else:
            import wildgs_slam_blackwell_b200.rasterizer as mod

        def step():
            for t in leaves:
                t.grad = None
            if arm == "fused":
                pkg = renderer.render(cam, pc, Pipe(), bg)
                if wild:
                    _, l = fused_loss.get_loss_mapping_uncertainty(UCFG, pkg["render"], pkg["depth"], cam, pkg["opacity"],
                                                                   net, 0.3, 0.6)
                elif mapping:
                    l = fused_loss.get_loss_mapping({"Training": {"rgb_boundary_threshold": 0.01, "alpha": 0.95}},
                                                    pkg["render"], pkg["depth"], cam)
                else:
                    l = fused_loss.get_loss_tracking(CONFIG, pkg["render"], pkg["depth"], pkg["opacity"], cam)
            else:
                img, depth, opacity = composition(mod, cam, pc, bg)
                if wild:
                    l = torch_loss_mapping_uncertainty(img, depth, cam, gt_depth, opacity, net)
                else:
                    l = torch_loss_mapping(img, depth, cam, gt_depth) if mapping else torch_loss_tracking(img, opacity, cam)
            l.backward()
            if opt is not None:
                opt.step()

        res[arm] = round(timeit(step, flush), 4)
        if arm == "fused":
            from wildgs_slam_blackwell_b200 import graph
            it = graph.CapturedIteration(step, DEV)
            res["fused_graph"] = round(timeit(it.replay, flush), 4)
            it.check()
    what = ("mapping iteration with the mapper's real loss (render + get_loss_mapping_uncertainty: SSIM, uncertainty MLP, "
            "weighted L1 + backward + Adam step), config3 Gaussians (1 M), one keyframe" if wild else
            "mapping iteration (render + mapping loss + backward + Adam step), config3 Gaussians (1 M), one keyframe"
            if mapping else "tracking iteration (render + tracking loss + backward), config2")
    print(json.dumps({"what": what + ", ms per iteration (median)",
                      **res}))


if __name__ == "__main__":
    main()
