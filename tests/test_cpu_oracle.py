"""CPU (no GPU): pins the oracle (oracle/cpu_rasterizer.c) against the golden vectors the UNMODIFIED
reference produced on a B200 (tests/golden/*.npz, generator: tests/golden/make_golden.py).

Bars: per-Gaussian forward state, radii, sorted keys / ids, tile ranges: bit-exact.  Compositing uses
host libm expf (CUDA's differs by <= 2 ulp), so n_contrib / n_touched / final_T may differ in a few
pixels where alpha sits on a threshold: <= 0.2 % of pixels, everything else rtol 1e-4 like the GPU
bar.  Gradients: rtol 1e-3 + atol 1e-4*max|ref| like the GPU bar, on >= 99.8 % of entries."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

import make_golden  # noqa: E402
from oracle import cpu_oracle  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module", params=sorted(make_golden.CASES))
def pair(request):
    c = make_golden.CASES[request.param]
    g, cam, cot, colors, cov = make_golden.build_case(c)
    o = cpu_oracle.forward_backward(g, cam, cot, colors, cov, export=True)
    return c, o, dict(np.load(os.path.join(GOLDEN, request.param + ".npz")))


def test_preprocess_and_binning_bit_exact(pair):
    c, o, r = pair
    np.testing.assert_array_equal(o["radii"], r["radii"])
    assert int(o["num_rendered"]) == int(r["num_rendered"])
    for k in ("st_depths", "st_means2D", "st_conic_opacity"):
        np.testing.assert_array_equal(bits(o[k]), bits(r[k]), err_msg=k)
    if not c["precomp"]:
        np.testing.assert_array_equal(bits(o["st_rgb"]), bits(r["st_rgb"]))
        np.testing.assert_array_equal(o["st_clamped"], r["st_clamped"])
        np.testing.assert_array_equal(bits(o["st_cov3D"]), bits(r["st_cov3D"]))
    np.testing.assert_array_equal(o["st_point_list_keys"].view(np.uint64), r["st_point_list_keys"].view(np.uint64))
    np.testing.assert_array_equal(o["st_point_list"].view(np.int32), r["st_point_list"].view(np.int32))
    np.testing.assert_array_equal(o["st_ranges"].view(np.int32), r["st_ranges"].view(np.int32))


def _frac_bad(a, b, rtol, atol):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((np.abs(a - b) > rtol * np.abs(b) + atol).mean())


def test_compositing(pair):
    c, o, r = pair
    N = o["st_n_contrib"].size
    assert (o["st_n_contrib"].view(np.int32) != r["st_n_contrib"].view(np.int32)).sum() <= 0.002 * N
    assert (o["n_touched"] != r["n_touched"]).sum() <= max(2, 0.002 * o["n_touched"].size)
    for k in ("color", "depth", "opacity"):
        assert _frac_bad(o[k], r[k], 1e-4, 1e-6) <= 0.002, k
    assert _frac_bad(o["st_final_T"], r["st_final_T"], 1e-4, 1e-7) <= 0.002


def test_gradients(pair):
    c, o, r = pair
    names = ["dL_dmeans3D", "dL_dmeans2D", "dL_dopacity", "dL_dconic", "dL_ddepths"]
    names += ["dL_dcolors", "dL_dcov3D"] if c["precomp"] else ["dL_dsh", "dL_dscales", "dL_drotations", "dL_dcov3D"]
    for n in names:
        ref = r["g_" + n]
        scale = np.abs(ref).max()
        assert _frac_bad(o["g_" + n], ref, 1e-3, 1e-4 * scale + 1e-30) <= 0.002, n
    ref = r["g_dL_dtau_sum"]
    np.testing.assert_allclose(o["g_dL_dtau_sum"], ref, rtol=2e-3, atol=2e-4 * np.abs(ref).max())


def test_mark_visible_matches_near_cull():
    c = make_golden.CASES["g0_deg0"]
    g, cam, _, _, _ = make_golden.build_case(c)
    vis = cpu_oracle.mark_visible(g["means3D"], cam["viewmatrix"])
    o = cpu_oracle.forward_backward(g, cam)
    assert not (o["radii"][~vis] > 0).any()      # culled by the near plane => never rendered
    assert 0.9 < vis.mean() < 1.0                # 3 % of the scene is drawn behind the near plane


def test_empty_scene_is_background():
    c = make_golden.CASES["g1_deg3_bg_ragged"]
    g, cam, cot, _, _ = make_golden.build_case(c)
    g0 = {k: (v[:0] if isinstance(v, np.ndarray) else v) for k, v in g.items()}
    o = cpu_oracle.forward_backward(g0, cam, cot)
    for ch in range(3):
        np.testing.assert_array_equal(o["color"][ch], np.full_like(o["color"][ch], cam["bg"][ch]))
    assert int(o["num_rendered"]) == 0 and not o["opacity"].any()
    np.testing.assert_array_equal(o["dL_dtau"], np.zeros(6, np.float32))


def test_pose_and_parameter_gradients_are_derivatives_of_the_rendered_loss():
    """Correctness, not agreement: central finite differences of L = <color, dc> + <depth, dd> through the
    oracle's own forward -- along the six pose directions (left perturbation W2C <- SE3_exp(h e_i) W2C, the consumer's
    convention, pose_utils.py:81-98) and along random directions of means3D / opacity / sh / scales -- against the
    analytic dL_dtau and parameter gradients.  The forward is discontinuous where the analytic gradient is not (pixels
    enter or leave a footprint at alpha = 1/255, tile rectangles are rounded), so the scene is smooth (large,
    translucent Gaussians) and the cotangents vary slowly over the image; the remaining gap (2-5 %, always the finite
    difference slightly larger) is those boundary terms.  With white-noise cotangents they reach 20 %."""
    from wildgs_slam_blackwell_b200 import scenes
    W, H, P = 64, 48, 250
    cam = scenes.make_camera(W, H, 77)
    g = scenes.make_gaussians(P, W, H, 1, 77, cam)
    g = dict(g, scales=(g["scales"] * 8.0).astype(np.float32), opacities=np.clip(g["opacities"] * 0.5, 0.05, 0.3).astype(np.float32))
    yy, xx = np.mgrid[0:H, 0:W]
    smooth = (np.sin(xx / 9.0) + np.cos(yy / 7.0)).astype(np.float32)
    dc = (np.stack([smooth, 0.5 * smooth + 0.3, -smooth]) / (3 * H * W)).astype(np.float32)
    dd = (smooth[None] / (H * W)).astype(np.float32)
    rng = np.random.default_rng(5)

    def loss(gg, cc):
        o = cpu_oracle.forward_backward(gg, cc)
        return float((o["color"].astype(np.float64) * dc).sum() + (o["depth"].astype(np.float64) * dd).sum())

    def moved(tau_h):
        w2c = scenes.se3_exp(tau_h) @ cam["w2c"]
        view = w2c.T.astype(np.float32)
        return dict(cam, w2c=w2c, viewmatrix=np.ascontiguousarray(view), campos=np.linalg.inv(w2c)[:3, 3].astype(np.float32),
                    projmatrix=np.ascontiguousarray((view.astype(np.float64) @ cam["projmatrix_raw"].astype(np.float64)).astype(np.float32)))

    o = cpu_oracle.forward_backward(g, cam, (dc, dd))
    tau = o["dL_dtau"].astype(np.float64)
    num, h = np.zeros(6), 5e-4
    for i in range(6):
        e = np.zeros(6)
        e[i] = h
        num[i] = (loss(g, moved(e)) - loss(g, moved(-e))) / (2 * h)
    assert np.linalg.norm(num - tau) < 0.06 * np.linalg.norm(tau), (num, tau)
    for name, key, h in (("means3D", "g_dL_dmeans3D", 1e-3), ("opacities", "g_dL_dopacity", 2e-3), ("sh", "g_dL_dsh", 5e-3),
                         ("scales", "g_dL_dscales", 1e-3)):
        d = rng.normal(0, 1, g[name].shape)
        gp, gm = dict(g), dict(g)
        gp[name], gm[name] = (g[name] + h * d).astype(np.float32), (g[name] - h * d).astype(np.float32)
        fd = (loss(gp, cam) - loss(gm, cam)) / (2 * h)
        ana = float((o[key].astype(np.float64) * d).sum())
        assert abs(fd - ana) < 0.06 * abs(ana) + 1e-7, (name, fd, ana)
