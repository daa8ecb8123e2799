"""Shared helpers for the tests: scene -> tensors, running the product through its public API and
through the raw C-ABI exports, tolerance helpers."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from wildgs_slam_blackwell_b200 import _lib, rasterizer, scenes  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def golden_cases():
    sys.path.insert(0, GOLDEN_DIR)
    import make_golden
    return make_golden.CASES, make_golden.build_case


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN_DIR, name + ".npz")))


def to_dev(a, dev="cuda:0"):
    return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def settings_from_cam(cam, sh_degree, dev="cuda:0", debug=False, scale_modifier=1.0):
    return rasterizer.GaussianRasterizationSettings(
        image_height=cam["height"], image_width=cam["width"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
        bg=to_dev(cam["bg"], dev), scale_modifier=scale_modifier, viewmatrix=to_dev(cam["viewmatrix"], dev),
        projmatrix=to_dev(cam["projmatrix"], dev), projmatrix_raw=to_dev(cam["projmatrix_raw"], dev),
        sh_degree=sh_degree, campos=to_dev(cam["campos"], dev), prefiltered=False, debug=debug)


def run_ours(g, cam, cot, colors=None, cov=None, dev="cuda:0", export=True):
    """Forward + backward through the public autograd API; optional state export through the C ABI.
    Returns dict of numpy arrays keyed like the golden files."""
    s = settings_from_cam(cam, g["sh_degree"], dev)
    P = g["means3D"].shape[0]
    leaf = lambda a: None if a is None else to_dev(a, dev).requires_grad_(True)
    means3D, opac = leaf(g["means3D"]), leaf(g["opacities"])
    means2D = torch.zeros(P, 3, device=dev, requires_grad=True)
    sh = None if colors is not None else leaf(g["sh"])
    colors_t = leaf(colors)
    scales = None if cov is not None else leaf(g["scales"])
    rots = None if cov is not None else leaf(g["rotations"])
    cov_t = leaf(cov)
    theta = torch.zeros(3, device=dev, requires_grad=True)
    rho = torch.zeros(3, device=dev, requires_grad=True)
    rast = rasterizer.GaussianRasterizer(s)
    color, radii, depth, opacity, n_touched = rast(means3D, means2D, opac, shs=sh, colors_precomp=colors_t,
                                                   scales=scales, rotations=rots, cov3D_precomp=cov_t,
                                                   theta=theta, rho=rho)
    out = dict(color=color, depth=depth, opacity=opacity, radii=radii, n_touched=n_touched)
    rec = {k: v.detach().cpu().numpy() for k, v in out.items()}
    if export:
        rec.update(export_state(g, cam, colors, cov, dev))
    if cot is not None:
        dc, dd = to_dev(cot[0], dev), to_dev(cot[1], dev)
        loss = (color * dc).sum() + (depth * dd).sum()
        loss.backward()
        torch.cuda.synchronize()
        gr = lambda t: None if t is None or t.grad is None else t.grad.detach().cpu().numpy()
        rec["g_dL_dmeans3D"] = gr(means3D)
        rec["g_dL_dmeans2D"] = gr(means2D)
        rec["g_dL_dopacity"] = gr(opac)
        rec["g_dL_dsh"] = gr(sh)
        rec["g_dL_dcolors"] = gr(colors_t)
        rec["g_dL_dscales"] = gr(scales)
        rec["g_dL_drotations"] = gr(rots)
        rec["g_dL_dcov3D"] = gr(cov_t)
        rec["g_dL_dtau_sum"] = np.concatenate([gr(rho).reshape(-1), gr(theta).reshape(-1)])
    return rec


def export_state(g, cam, colors=None, cov=None, dev="cuda:0"):
    """Re-run the forward through forward_raw and pull the reference-format state out of the
    C-ABI exports (wgs_export_binning / wgs_export_geometry)."""
    s = settings_from_cam(cam, g["sh_degree"], dev)
    e = torch.Tensor([])
    sh = e if colors is not None else to_dev(g["sh"], dev)
    col = e if colors is None else to_dev(colors, dev)
    scales = e if cov is not None else to_dev(g["scales"], dev)
    rots = e if cov is not None else to_dev(g["rotations"], dev)
    covt = e if cov is None else to_dev(cov, dev)
    # (outputs are held by these names until the exports below have run)
    color, radii, depth, opacity, n_touched, st = rasterizer.forward_raw(
        to_dev(g["means3D"], dev), sh, col, to_dev(g["opacities"], dev), scales, rots, covt, s)
    P = g["means3D"].shape[0]
    R = st.num_rendered
    W, H = cam["width"], cam["height"]
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    f = lambda *sh_: torch.zeros(*sh_, device=dev)
    i32 = lambda *sh_: torch.zeros(*sh_, dtype=torch.int32, device=dev)
    keys = torch.zeros(max(R, 1), dtype=torch.int64, device=dev)
    plist, ranges, ncon, fT = i32(max(R, 1)), i32(tiles, 2), i32(H * W), f(H * W)
    lib = _lib.load()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.wgs_export_binning(C.byref(st.args), R, keys.data_ptr(), plist.data_ptr(), ranges.data_ptr(),
                                ncon.data_ptr(), fT.data_ptr(), stream)
    assert rc == 0, _lib.last_error()
    depths, m2d, co, rgb = f(P), f(P, 2), f(P, 4), f(P, 3)
    clamped = torch.zeros(P, 3, dtype=torch.uint8, device=dev)
    rc = lib.wgs_export_geometry(C.byref(st.args), depths.data_ptr(), m2d.data_ptr(), co.data_ptr(), rgb.data_ptr(),
                                 clamped.data_ptr(), stream)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    n = lambda t: t.cpu().numpy()
    return dict(num_rendered=np.int64(R), st_point_list_keys=n(keys[:R]), st_point_list=n(plist[:R]),
                st_ranges=n(ranges), st_n_contrib=n(ncon), st_final_T=n(fT), st_depths=n(depths),
                st_means2D=n(m2d), st_conic_opacity=n(co), st_rgb=n(rgb), st_clamped=n(clamped))


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def assert_grad_close(name, ours, ref, rtol=1e-3, atol_scale=1e-4):
    """rtol 1e-3 + atol tied to max|ref| (summation order differs by design, cancellation)."""
    ours, ref = np.asarray(ours, np.float64), np.asarray(ref, np.float64)
    assert ours.shape == ref.shape, (name, ours.shape, ref.shape)
    scale = np.abs(ref).max() if ref.size else 0.0
    err = np.abs(ours - ref)
    tol = rtol * np.abs(ref) + atol_scale * scale + 1e-30
    bad = err > tol
    assert not bad.any(), (f"{name}: {bad.sum()} / {bad.size} entries out of tolerance; max err "
                           f"{err.max():.3e} at scale {scale:.3e}; worst ratio {(err / tol).max():.2f}")
