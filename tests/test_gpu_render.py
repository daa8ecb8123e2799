"""GPU: render() (mirror of the reference's gaussian_renderer.render) with the activations, the dc|rest
concatenation and the camera matrices fused into kernels, against the reference's own composition
(torch activations + torch camera properties) run on top of the parity-pinned drop-in rasterizer.
SURVEY.md §8(f) row 1."""
import math
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import util  # noqa: E402
from wildgs_slam_blackwell_b200 import renderer, scenes  # noqa: E402
from wildgs_slam_blackwell_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


class Pipe:
    compute_cov3D_python = False
    convert_SHs_python = False


class Model:
    """The slice of the reference GaussianModel render() touches (gaussian_model.py:54-63,77-97)."""

    def __init__(self, g, seed):
        rng = np.random.default_rng(seed)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(DEV).requires_grad_(True)
        o = np.clip(g["opacities"], 1e-4, 1 - 1e-4)
        self._xyz = t(g["means3D"])
        self._features_dc = t(g["sh"][:, :1])
        self._features_rest = t(g["sh"][:, 1:])
        self._opacity = t(np.log(o / (1 - o)))
        self._scaling = t(np.log(g["scales"]))
        self._rotation = t(g["rotations"] * rng.uniform(0.5, 2.0, (g["rotations"].shape[0], 1)))  # not unit length
        self.active_sh_degree = g["sh_degree"]
        self.max_sh_degree = g["sh_degree"]

    get_xyz = property(lambda s: s._xyz)
    get_opacity = property(lambda s: torch.sigmoid(s._opacity))
    get_scaling = property(lambda s: torch.exp(s._scaling))
    get_rotation = property(lambda s: torch.nn.functional.normalize(s._rotation))
    get_features = property(lambda s: torch.cat((s._features_dc, s._features_rest), dim=1))

    def leaves(self):
        return dict(xyz=self._xyz, f_dc=self._features_dc, f_rest=self._features_rest, opacity=self._opacity,
                    scaling=self._scaling, rotation=self._rotation)


class Cam:
    """The slice of the reference Camera render() touches; the three properties are the reference's torch
    expressions (camera_utils.py:137-151, graphics_utils.py:33-46)."""

    def __init__(self, cam):
        w2c = np.asarray(cam["viewmatrix"], np.float64).T
        self.R = torch.from_numpy(w2c[:3, :3].astype(np.float32)).to(DEV)
        self.T = torch.from_numpy(w2c[:3, 3].astype(np.float32)).to(DEV)
        self.projection_matrix = torch.from_numpy(np.asarray(cam["projmatrix_raw"], np.float32)).to(DEV)
        self.FoVx, self.FoVy = 2 * math.atan(cam["tanfovx"]), 2 * math.atan(cam["tanfovy"])
        self.image_height, self.image_width = cam["height"], cam["width"]
        self.cam_rot_delta = torch.zeros(3, device=DEV, requires_grad=True)
        self.cam_trans_delta = torch.zeros(3, device=DEV, requires_grad=True)

    @property
    def world_view_transform(self):
        Rt = torch.zeros((4, 4), device=DEV)
        Rt[:3, :3] = self.R
        Rt[:3, 3] = self.T
        Rt[3, 3] = 1.0
        return torch.linalg.inv(torch.linalg.inv(Rt)).transpose(0, 1)

    @property
    def full_proj_transform(self):
        return self.world_view_transform.unsqueeze(0).bmm(self.projection_matrix.unsqueeze(0)).squeeze(0)

    @property
    def camera_center(self):
        return self.world_view_transform.inverse()[3, :3]


def reference_composition(cam, pc, bg):
    """gaussian_renderer/__init__.py:41-141 verbatim in behaviour, on the drop-in rasterizer."""
    screenspace_points = torch.zeros_like(pc.get_xyz, requires_grad=True)
    s = GaussianRasterizationSettings(
        image_height=int(cam.image_height), image_width=int(cam.image_width), tanfovx=math.tan(cam.FoVx * 0.5),
        tanfovy=math.tan(cam.FoVy * 0.5), bg=bg, scale_modifier=1.0, viewmatrix=cam.world_view_transform,
        projmatrix=cam.full_proj_transform, projmatrix_raw=cam.projection_matrix, sh_degree=pc.active_sh_degree,
        campos=cam.camera_center, prefiltered=False, debug=False)
    img, radii, depth, opacity, n_touched = GaussianRasterizer(s)(
        means3D=pc.get_xyz, means2D=screenspace_points, shs=pc.get_features, colors_precomp=None,
        opacities=pc.get_opacity, scales=pc.get_scaling, rotations=pc.get_rotation, cov3D_precomp=None,
        theta=cam.cam_rot_delta, rho=cam.cam_trans_delta)
    return dict(render=img, viewspace_points=screenspace_points, radii=radii, depth=depth, opacity=opacity,
                n_touched=n_touched)


def test_camera_matrices_match_reference_properties():
    cam = Cam(scenes.make_camera(320, 240, 77))
    v, p, c = renderer.camera_matrices(cam.R, cam.T, cam.projection_matrix)
    torch.testing.assert_close(v, cam.world_view_transform, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(p, cam.full_proj_transform, rtol=1e-5, atol=2e-6)
    torch.testing.assert_close(c, cam.camera_center, rtol=1e-5, atol=1e-6)
    # cached until the pose tensors change
    a = renderer._cached_matrices(cam)
    assert renderer._cached_matrices(cam)[0] is a[0]
    cam.T = cam.T + 0.1
    assert renderer._cached_matrices(cam)[0] is not a[0]
    with torch.no_grad():
        cam.R.mul_(1.0)
    b = renderer._cached_matrices(cam)
    assert b[0] is not a[0]


@pytest.mark.parametrize("P,W,H,deg,seed", [(20000, 320, 240, 0, 11), (8000, 200, 136, 3, 12), (5000, 96, 64, 1, 13)])
def test_fused_render_matches_reference_composition(P, W, H, deg, seed):
    cam_d = scenes.make_camera(W, H, seed, bg=(0.2, 0.4, 0.1))
    g = scenes.make_gaussians(P, W, H, deg, seed, scenes.make_camera(W, H, seed))
    dc, dd = scenes.make_cotangents(W, H, seed)
    dc, dd = torch.from_numpy(dc).to(DEV), torch.from_numpy(dd).to(DEV)
    bg = torch.tensor(cam_d["bg"], dtype=torch.float32, device=DEV)
    out = {}
    for name in ("fused", "ref"):
        pc, cam = Model(g, seed), Cam(cam_d)
        pkg = renderer.render(cam, pc, Pipe(), bg) if name == "fused" else reference_composition(cam, pc, bg)
        ((pkg["render"] * dc).sum() + (pkg["depth"] * dd).sum()).backward()
        torch.cuda.synchronize()
        rec = {k: pkg[k].detach().cpu().numpy() for k in ("render", "depth", "opacity", "radii", "n_touched")}
        rec.update({"g_" + k: v.grad.cpu().numpy() for k, v in pc.leaves().items() if v.grad is not None})
        rec["g_means2D"] = pkg["viewspace_points"].grad.cpu().numpy()
        rec["g_tau"] = np.concatenate([cam.cam_trans_delta.grad.cpu().numpy().ravel(), cam.cam_rot_delta.grad.cpu().numpy().ravel()])
        out[name] = rec
    f, r = out["fused"], out["ref"]
    if name == "fused":
        assert set(pkg) == {"render", "viewspace_points", "visibility_filter", "radii", "depth", "opacity", "n_touched"}
    # activations and matrices are evaluated in a different (fused) order: radii may flip on a handful of
    # Gaussians sitting on a ceil() boundary; images within 1e-4 of the image scale otherwise
    assert (f["radii"] != r["radii"]).mean() < 2e-4
    same = f["radii"] == r["radii"]
    # Compositing is discontinuous at its alpha >= 1/255 test: an input perturbation of 1 ulp flips the
    # decision for a few (pixel, Gaussian) pairs, each worth <= 1/255 of the pixel scale.  So: all but
    # 1e-4 of the pixels within 2e-4 of the image scale, and no pixel off by more than two such flips.
    for k in ("render", "depth", "opacity"):
        scale = max(1.0, np.abs(r[k]).max())
        err = np.abs(f[k] - r[k])
        assert (err > 2e-4 * scale + 1e-3 * np.abs(r[k])).mean() < 1e-4, (k, err.max())
        assert err.max() < 2.0 / 255.0 * scale, (k, err.max())
    assert (f["n_touched"][same] != r["n_touched"][same]).mean() < 1e-3
    for k in ("g_xyz", "g_f_dc", "g_opacity", "g_scaling", "g_rotation", "g_means2D", "g_tau") + (("g_f_rest",) if deg else ()):
        # same tolerance as the rasterizer's own gradient parity, except on the Gaussians touched by a flip
        a, b = np.asarray(f[k], np.float64), np.asarray(r[k], np.float64)
        assert a.shape == b.shape, k
        scale = np.abs(b).max()
        err = np.abs(a - b)
        bad = err > 2e-3 * np.abs(b) + 2e-4 * scale
        assert bad.mean() < (5e-4 if a.size > 100 else 1e-9), (k, bad.sum(), a.size, err.max(), scale)
        assert err.max() < 5e-2 * scale, (k, err.max(), scale)


def test_unfused_branches_run():
    W, H, P = 96, 64, 3000
    cam_d = scenes.make_camera(W, H, 5)
    g = scenes.make_gaussians(P, W, H, 0, 5, cam_d)
    pc, cam = Model(g, 5), Cam(cam_d)
    bg = torch.zeros(3, device=DEV)
    full = renderer.render(cam, pc, Pipe(), bg)
    mask = torch.ones(P, dtype=torch.bool, device=DEV)
    masked = renderer.render(cam, pc, Pipe(), bg, mask=mask)
    torch.testing.assert_close(masked["render"], full["render"], rtol=1e-3, atol=2e-4)
    col = torch.rand(P, 3, device=DEV)
    over = renderer.render(cam, pc, Pipe(), bg, override_color=col)
    assert over["render"].shape == (3, H, W)
    empty = Model(g, 5)
    empty._xyz = torch.zeros(0, 3, device=DEV)
    assert renderer.render(cam, empty, Pipe(), bg) is None


def test_captured_iteration_matches_eager_and_follows_updates():
    """The tracking loop body as one CUDA graph: same loss and gradients as the eager iteration, and in-place
    updates of the pose deltas / exposure between replays are picked up."""
    from wildgs_slam_blackwell_b200 import graph, loss as fused_loss
    W, H, P = 160, 112, 6000
    cam_d = scenes.make_camera(W, H, 21)
    g = scenes.make_gaussians(P, W, H, 0, 21, cam_d)
    pc, cam = Model(g, 21), Cam(cam_d)
    rng = np.random.default_rng(0)
    cam.original_image = torch.from_numpy(rng.uniform(0, 1, (3, H, W)).astype(np.float32)).to(DEV)
    cam.grad_mask = torch.ones(1, H, W, dtype=torch.bool, device=DEV)
    cam.exposure_a = torch.zeros(1, device=DEV, requires_grad=True)
    cam.exposure_b = torch.zeros(1, device=DEV, requires_grad=True)
    bg = torch.zeros(3, device=DEV)
    cfg = {"Training": {"rgb_boundary_threshold": 0.01}}
    leaves = [cam.cam_rot_delta, cam.cam_trans_delta, cam.exposure_a, cam.exposure_b, pc._xyz]

    def step():
        for t in leaves:
            t.grad = None
        pkg = renderer.render(cam, pc, Pipe(), bg)
        l = fused_loss.get_loss_tracking(cfg, pkg["render"], pkg["depth"], pkg["opacity"], cam)
        l.backward()
        return l.detach()

    def snapshot(l):
        torch.cuda.synchronize()
        return [l.item()] + [t.grad.clone() for t in leaves]

    eager0 = snapshot(step())
    it = graph.CapturedIteration(step, DEV)
    cap0 = snapshot(it.replay())
    assert abs(cap0[0] - eager0[0]) <= 1e-6 * abs(eager0[0])
    for a, b in zip(cap0[1:], eager0[1:]):
        util.assert_grad_close("captured", a.cpu().numpy(), b.cpu().numpy())
    with torch.no_grad():
        cam.exposure_b.add_(0.05)
        pc._xyz.add_(0.01)
    cap1 = snapshot(it.replay())
    n = it.check()
    assert n > 0
    it = None
    eager1 = snapshot(step())
    assert abs(cap1[0] - eager1[0]) <= 1e-6 * abs(eager1[0]) and abs(cap1[0] - cap0[0]) > 1e-4
    for a, b in zip(cap1[1:], eager1[1:]):
        util.assert_grad_close("captured after update", a.cpu().numpy(), b.cpu().numpy())


def test_async_forward_reports_overflow():
    from wildgs_slam_blackwell_b200 import graph, rasterizer
    W, H, P = 96, 64, 3000
    cam_d = scenes.make_camera(W, H, 5)
    g = scenes.make_gaussians(P, W, H, 0, 5, cam_d)
    s = util.settings_from_cam(cam_d, 0, DEV)
    args = [util.to_dev(g[k], DEV) for k in ("means3D", "sh")] + [torch.Tensor([])] + \
           [util.to_dev(g[k], DEV) for k in ("opacities", "scales", "rotations")] + [torch.Tensor([]), s]
    ref = rasterizer.forward_raw(*args)
    with graph.async_forward(DEV):
        out = rasterizer.forward_raw(*args)
        assert out[-1].num_rendered == -1
    assert graph.check_overflow(DEV) == ref[-1].num_rendered
    assert torch.equal(out[0], ref[0])


def test_sticky_overflow_record_covers_every_no_wait_forward():
    """No-wait mode: a forward that needs more instances than its capacity renders the background only (zero gradients),
    and the STICKY record reports it even when later forwards on the device were fine; the check raises the capacity
    floor so that the rerun fits."""
    from wildgs_slam_blackwell_b200 import graph, rasterizer
    W, H, P = 200, 120, 3000
    cam_d = scenes.make_camera(W, H, 21, bg=(0.2, 0.4, 0.6))
    g = scenes.make_gaussians(P, W, H, 0, 21, cam_d)
    big = dict(g, scales=g["scales"] * 12.0)                 # R >> max(4 P, 65536): the default capacity
    s = util.settings_from_cam(cam_d, 0, DEV)
    mk = lambda gg: [util.to_dev(gg[k], DEV) for k in ("means3D", "sh")] + [torch.Tensor([])] + \
        [util.to_dev(gg[k], DEV) for k in ("opacities", "scales", "rotations")] + [torch.Tensor([]), s]
    rasterizer._capacity_hint.clear()
    rasterizer._capacity_ratio.clear()
    rasterizer._capacity_floor.clear()
    graph.overflow_status(DEV, clear=True)
    ref = rasterizer.forward_raw(*mk(big))                   # blocking forward: grows and retries
    R_big = ref[-1].num_rendered
    assert R_big > max(4 * P, 1 << 16)
    rasterizer._capacity_hint.clear()
    rasterizer._capacity_ratio.clear()
    with graph.async_forward(DEV):
        with graph.async_forward(DEV):                        # nests: the inner exit must not switch the mode off
            pass
        out_big = rasterizer.forward_raw(*mk(big))            # overflows silently ...
        assert out_big[-1].num_rendered == -1
        out_small = rasterizer.forward_raw(*mk(g))            # ... and is followed by a forward that is fine
    torch.cuda.synchronize()
    for c in range(3):
        assert torch.all(out_big[0][c] == float(cam_d["bg"][c]))
    assert not out_big[2].any() and not out_big[3].any()
    n, cap = graph.last_num_rendered(DEV)
    assert n <= cap                                           # the most recent forward alone looks healthy
    with pytest.raises(RuntimeError, match="instance capacity exceeded by 1 forward") as ei:
        graph.check_overflow(DEV)
    assert ei.value.needed == R_big
    assert graph.overflow_status(DEV) == (0, 0)               # cleared by the check
    with graph.async_forward(DEV):
        again = rasterizer.forward_raw(*mk(big))              # the floor was raised: fits now
    assert graph.check_overflow(DEV) == R_big
    assert torch.equal(again[0], ref[0])
    rasterizer._capacity_floor.clear()


def test_backward_of_an_overflowed_no_wait_forward_is_all_zero_and_counts_no_observation():
    """The other half of the overflow contract: the backward of a forward that rendered the background only leaves zero
    gradients (dL_dtau included) and does NOT touch the fused densification statistics -- the view was not observed.
    (An optimiser step already taken with that view missing cannot be undone by rerunning the forward.)"""
    from wildgs_slam_blackwell_b200 import graph, rasterizer
    W, H, P = 200, 120, 3000
    cam_d = scenes.make_camera(W, H, 21, bg=(0.2, 0.4, 0.6))
    g = scenes.make_gaussians(P, W, H, 0, 21, cam_d)
    big = dict(g, scales=g["scales"] * 12.0)
    s = util.settings_from_cam(cam_d, 0, DEV)
    leaves = {k: util.to_dev(big[k], DEV).requires_grad_(True) for k in ("means3D", "sh", "opacities", "scales", "rotations")}
    m2 = torch.zeros(P, 3, device=DEV, requires_grad=True)
    theta = torch.zeros(3, device=DEV, requires_grad=True)
    rho = torch.zeros(3, device=DEV, requires_grad=True)
    acc = torch.full((P, 1), 3.0, device=DEV)
    den = torch.full((P, 1), 5.0, device=DEV)
    mx = torch.full((P,), 7.0, device=DEV)
    rasterizer._capacity_hint.clear()
    rasterizer._capacity_ratio.clear()
    rasterizer._capacity_floor.clear()
    graph.overflow_status(DEV, clear=True)
    rast = rasterizer.GaussianRasterizer(s)
    with graph.async_forward(DEV), rasterizer.densification_stats(acc, den, mx):
        color, radii, depth, opacity, n_touched = rast(leaves["means3D"], m2, leaves["opacities"], shs=leaves["sh"],
                                                       scales=leaves["scales"], rotations=leaves["rotations"],
                                                       theta=theta, rho=rho)
        (color.sum() + depth.sum() + opacity.sum()).backward()
    with pytest.raises(RuntimeError, match="instance capacity exceeded"):
        graph.check_overflow(DEV)
    assert (radii > 0).any()                                 # Gaussians were visible: only the overflow keeps them out
    for k, t in leaves.items():
        assert t.grad is not None and not t.grad.any(), k
    assert not m2.grad.any() and not theta.grad.any() and not rho.grad.any()
    assert torch.all(acc == 3.0) and torch.all(den == 5.0) and torch.all(mx == 7.0)
    # the same view with room: gradients flow and the statistics move
    for t in list(leaves.values()) + [m2, theta, rho]:
        t.grad = None
    with rasterizer.densification_stats(acc, den, mx):
        color, radii, depth, opacity, n_touched = rast(leaves["means3D"], m2, leaves["opacities"], shs=leaves["sh"],
                                                       scales=leaves["scales"], rotations=leaves["rotations"],
                                                       theta=theta, rho=rho)
        (color.sum() + depth.sum() + opacity.sum()).backward()
    assert leaves["means3D"].grad.any() and theta.grad.any() and float(den.max()) == 6.0 and float(mx.max()) > 7.0
    rasterizer._capacity_floor.clear()
    rasterizer._capacity_hint.clear()
    rasterizer._capacity_ratio.clear()


def test_capacity_carries_over_when_the_row_count_changes():
    """A densify / prune changes P.  The instances-per-Gaussian ratio learned by a blocking forward at this resolution
    sizes the first forward at the new P, so a no-wait forward there fits instead of rendering background only."""
    from wildgs_slam_blackwell_b200 import graph, rasterizer
    W, H, P = 200, 120, 3000
    cam_d = scenes.make_camera(W, H, 21, bg=(0.2, 0.4, 0.6))
    g = scenes.make_gaussians(P, W, H, 0, 21, cam_d)
    big = dict(g, scales=g["scales"] * 12.0)                 # R >> max(4 P, 65536)
    s = util.settings_from_cam(cam_d, 0, DEV)
    mk = lambda gg, n: [util.to_dev(gg[k][:n], DEV) for k in ("means3D", "sh")] + [torch.Tensor([])] + \
        [util.to_dev(gg[k][:n], DEV) for k in ("opacities", "scales", "rotations")] + [torch.Tensor([]), s]
    rasterizer._capacity_hint.clear()
    rasterizer._capacity_ratio.clear()
    rasterizer._capacity_floor.clear()
    graph.overflow_status(DEV, clear=True)
    ref = rasterizer.forward_raw(*mk(big, P - 200))          # blocking: learns the ratio at P - 200 rows
    assert ref[-1].num_rendered > max(4 * P, 1 << 16)
    with graph.async_forward(DEV):
        out = rasterizer.forward_raw(*mk(big, P))            # another row count, no hint of its own
    n = graph.check_overflow(DEV)                            # does not raise: the carried capacity was enough
    assert n > ref[-1].num_rendered
    want = rasterizer.forward_raw(*mk(big, P))
    assert torch.equal(out[0], want[0])
    for i in range(100):                                     # the table stays bounded
        rasterizer._capacity_hint[(0, 1, 1, i)] = 1
        rasterizer._capacity_hint.move_to_end((0, 1, 1, i))
    rasterizer.forward_raw(*mk(g, P))
    assert len(rasterizer._capacity_hint) <= rasterizer._HINT_KEEP
    rasterizer._capacity_hint.clear()
    rasterizer._capacity_ratio.clear()
