"""GPU: behaviour of the drop-in API and the C ABI beyond the golden scenes — edge cases (empty scene,
capacity overflow + retry, ragged image sizes, markVisible), parity with the CPU oracle on config 1,
parity with the UNMODIFIED reference (oracle/_ref, when it travelled to the box) at config-2 size, and
size-independent properties at BASELINE.json's full size."""
import os
import sys

import numpy as np
import pytest
import torch

import util
from oracle import cpu_oracle
from oracle.ref_cuda import ref_api
from wildgs_slam_blackwell_b200 import rasterizer, scenes

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_config1_matches_cpu_oracle():
    g, cams, cots = scenes.make_scene("config1")
    ours = util.run_ours(g, cams[0], cots[0])
    o = cpu_oracle.forward_backward(g, cams[0], cots[0], export=True)
    np.testing.assert_array_equal(ours["radii"], o["radii"])
    assert int(ours["num_rendered"]) == int(o["num_rendered"])
    np.testing.assert_array_equal(ours["st_point_list"].view(np.uint32), o["st_point_list"])
    np.testing.assert_array_equal(ours["st_ranges"].view(np.uint32), o["st_ranges"])
    np.testing.assert_array_equal(util.bits(ours["st_depths"]), util.bits(o["st_depths"]))
    assert (ours["st_n_contrib"].view(np.uint32) != o["st_n_contrib"]).mean() < 2e-3     # host expf vs CUDA expf
    for k in ("color", "depth", "opacity"):
        bad = np.abs(ours[k] - o[k]) > 1e-4 * np.abs(o[k]) + 1e-6
        assert bad.mean() < 2e-3, k
    util.assert_grad_close("tau", ours["g_dL_dtau_sum"], o["dL_dtau"], rtol=2e-3, atol_scale=2e-4)
    for a, b in (("g_dL_dmeans3D", "g_dL_dmeans3D"), ("g_dL_dopacity", "g_dL_dopacity"), ("g_dL_dscales", "g_dL_dscales"),
                 ("g_dL_drotations", "g_dL_drotations"), ("g_dL_dsh", "g_dL_dsh")):
        ref = o[b]
        bad = np.abs(ours[a] - ref) > 1e-3 * np.abs(ref) + 1e-4 * np.abs(ref).max()
        assert bad.mean() < 2e-3, a


def test_empty_scene_is_background():
    cam = scenes.make_camera(70, 50, 3, bg=(0.2, 0.4, 0.6))
    s = util.settings_from_cam(cam, 0)
    rast = rasterizer.GaussianRasterizer(s)
    z = lambda *sh: torch.zeros(*sh, device=DEV)
    color, radii, depth, opacity, n_touched = rast(z(0, 3), z(0, 3), z(0, 1), shs=z(0, 1, 3), scales=z(0, 3), rotations=z(0, 4))
    for c in range(3):
        assert torch.all(color[c] == cam["bg"][c])
    assert not depth.any() and not opacity.any() and radii.numel() == 0 and n_touched.numel() == 0


def test_capacity_overflow_retries_and_matches_oracle():
    g, cams, cots = scenes.make_scene(3000, 200, 120, 0, seed=21)
    g["scales"] = g["scales"] * 12.0                      # huge splats: R >> 4P and >> the 65536 floor
    o = cpu_oracle.forward_backward(g, cams[0], cots[0], export=True)
    assert int(o["num_rendered"]) > max(4 * 3000, 1 << 16)
    rasterizer._capacity_hint.clear()
    rasterizer._capacity_ratio.clear()
    ours = util.run_ours(g, cams[0], cots[0])
    assert int(ours["num_rendered"]) == int(o["num_rendered"])
    np.testing.assert_array_equal(ours["radii"], o["radii"])
    np.testing.assert_array_equal(ours["st_point_list"].view(np.uint32), o["st_point_list"])
    util.assert_grad_close("tau", ours["g_dL_dtau_sum"], o["dL_dtau"], rtol=2e-3, atol_scale=2e-4)
    assert rasterizer._capacity_hint[(0, 200, 120, 3000)] >= int(o["num_rendered"])   # next call starts large enough


def test_very_long_tile_lists_use_the_large_sort():
    g, cams, cots = scenes.make_scene(30000, 48, 32, 0, seed=22)              # 6 tiles, > 4096 entries each
    o = cpu_oracle.forward_backward(g, cams[0], None, export=True)
    lens = (o["st_ranges"][:, 1] - o["st_ranges"][:, 0])
    assert lens.max() > 4096
    ours = util.run_ours(g, cams[0], None)
    np.testing.assert_array_equal(ours["st_point_list"].view(np.uint32), o["st_point_list"])
    np.testing.assert_array_equal(ours["st_point_list_keys"].view(np.uint64), o["st_point_list_keys"])
    np.testing.assert_array_equal(ours["st_ranges"].view(np.uint32), o["st_ranges"])


@pytest.mark.parametrize("P, W, H, sigma", [(6000, 96, 64, 4e-3), (40000, 48, 32, 2e-2), (6000, 96, 64, 2e-4), (40000, 48, 32, 2e-4)])
def test_clustered_depths_sort_like_the_reference(P, W, H, sigma):
    """Two narrow depth clusters per tile (a foreground sheet in front of a far one, plus exact clones: what densification
    produces and what a fronto-parallel surface looks like) put hundreds of entries into a few buckets of the per-tile
    bucket sort -- the per-bucket warp sort (17..256 entries) and the whole-segment fallback (> 256) must give the
    reference's order (stable on equal depths = ascending id); small tiles (<= 4096 entries) and long ones."""
    g, cams, _ = scenes.make_scene(P, W, H, 0, seed=41)
    g = {k: (v.copy() if hasattr(v, "copy") else v) for k, v in g.items()}
    cam = cams[0]
    # (sigma 4e-3 / 2e-2: ~60 entries per bucket -> the warp sort; 2e-4: everything in one or two buckets -> the fallback)
    # move every Gaussian onto one of two planes of constant view-space depth (+ a tiny spread), keep its pixel position
    V = cam["viewmatrix"].reshape(4, 4)                       # row-vector convention: p_view = [p, 1] @ V
    pv = np.concatenate([g["means3D"], np.ones((P, 1), np.float32)], 1) @ V
    rng = np.random.default_rng(41)
    z_new = np.where(rng.random(P) < 0.5, 2.0, 9.0).astype(np.float32) + rng.normal(0, sigma, P).astype(np.float32)
    pv[:, :3] *= (z_new / np.maximum(pv[:, 2], 1e-3))[:, None]
    g["means3D"] = (pv @ np.linalg.inv(V))[:, :3].astype(np.float32)
    g["means3D"][P // 2:P // 2 + P // 8] = g["means3D"][:P // 8]           # exact clones: equal depth keys
    o = cpu_oracle.forward_backward(g, cam, None, export=True)
    lens = o["st_ranges"][:, 1] - o["st_ranges"][:, 0]
    assert (lens.max() > 4096) == (P == 40000)
    ours = util.run_ours(g, cam, None)
    np.testing.assert_array_equal(ours["st_ranges"].view(np.uint32), o["st_ranges"])
    np.testing.assert_array_equal(ours["st_point_list_keys"].view(np.uint64), o["st_point_list_keys"])
    np.testing.assert_array_equal(ours["st_point_list"].view(np.uint32), o["st_point_list"])


def test_mark_visible_matches_oracle():
    g, cams, _ = scenes.make_scene(5000, 64, 64, 0, seed=5)
    s = util.settings_from_cam(cams[0], 0)
    vis = rasterizer.GaussianRasterizer(s).markVisible(util.to_dev(g["means3D"]))
    assert vis.dtype == torch.bool
    np.testing.assert_array_equal(vis.cpu().numpy(), cpu_oracle.mark_visible(g["means3D"], cams[0]["viewmatrix"]))


def test_no_workspace_growth_over_steps():
    g, cams, cots = scenes.make_scene(20000, 320, 240, 0, seed=7)
    util.run_ours(g, cams[0], cots[0], export=False)
    torch.cuda.synchronize()
    base = torch.cuda.memory_allocated()
    for _ in range(25):
        util.run_ours(g, cams[0], cots[0], export=False)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() <= base + (1 << 20)   # no reference cycle keeps workspaces alive


def test_only_depth_or_only_color_cotangent():
    g, cams, cots = scenes.make_scene(2000, 96, 64, 0, seed=9)
    s = util.settings_from_cam(cams[0], 0)
    leaf = lambda a: util.to_dev(a).requires_grad_(True)
    m3, op, sh, sc, ro = leaf(g["means3D"]), leaf(g["opacities"]), leaf(g["sh"]), leaf(g["scales"]), leaf(g["rotations"])
    m2 = torch.zeros(2000, 3, device=DEV, requires_grad=True)
    th, rh = torch.zeros(3, device=DEV, requires_grad=True), torch.zeros(3, device=DEV, requires_grad=True)
    color, radii, depth, opacity, n_touched = rasterizer.GaussianRasterizer(s)(m3, m2, op, shs=sh, scales=sc, rotations=ro, theta=th, rho=rh)
    ga = torch.autograd.grad(depth.sum(), [m3, th, rh], retain_graph=True)
    gb = torch.autograd.grad(color.sum(), [m3, th, rh], retain_graph=True)
    gc = torch.autograd.grad(color.sum() + depth.sum() + 3.0 * opacity.sum(), [m3, th, rh])   # opacity cotangent is ignored
    for a, b, c in zip(ga, gb, gc):
        torch.testing.assert_close(a + b, c, rtol=2e-3, atol=1e-4 * float(c.abs().max()))
    assert th.grad is None and tuple(gc[1].shape) == (3,) and not radii.requires_grad and not n_touched.requires_grad


# ------------------------------- BASELINE.json full size (config 2) -------------------------------
@pytest.fixture(scope="module")
def config2():
    g, cams, cots = scenes.make_scene("config2")
    ours = util.run_ours(g, cams[0], cots[0])
    return g, cams[0], cots[0], ours


def test_config2_structural_properties(config2):
    g, cam, cot, o = config2
    R = int(o["num_rendered"])
    ranges = o["st_ranges"].view(np.uint32).astype(np.int64)
    lens = ranges[:, 1] - ranges[:, 0]
    nz = lens > 0
    assert lens.sum() == R                                           # ranges partition [0, R)
    assert np.array_equal(ranges[nz][1:, 0], ranges[nz][:-1, 1]) and ranges[nz][0, 0] == 0 and ranges[nz][-1, 1] == R
    keys = o["st_point_list_keys"].view(np.uint64)
    assert np.all(keys[1:] >= keys[:-1])                             # globally sorted (tile | depth)
    tile_of = (keys >> np.uint64(32)).astype(np.int64)
    assert np.array_equal(np.repeat(np.arange(len(lens)), lens), tile_of)
    ids = o["st_point_list"].view(np.uint32).astype(np.int64)
    eq = keys[1:] == keys[:-1]
    assert np.all(ids[1:][eq] > ids[:-1][eq])                        # stable: ties keep ascending id
    depth_bits = (keys & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    assert np.array_equal(depth_bits, util.bits(o["st_depths"])[ids])
    # every instance belongs to its Gaussian's tile rectangle count: sum over Gaussians == R
    vis = o["radii"] > 0
    assert np.array_equal(np.bincount(ids, minlength=len(vis)) > 0, vis)
    gx = (cam["width"] + 15) // 16
    ncon = o["st_n_contrib"].view(np.uint32).reshape(cam["height"], cam["width"])
    tile_len = lens.reshape(-1, gx)
    assert np.all(ncon <= np.repeat(np.repeat(tile_len, 16, 0), 16, 1)[:cam["height"], :cam["width"]])
    T = o["st_final_T"].reshape(cam["height"], cam["width"])
    np.testing.assert_allclose(o["opacity"][0], 1.0 - T, rtol=0, atol=1e-7)
    assert np.all((T >= 1e-4 * 0.01) & (T <= 1.0))


def test_config2_forward_is_deterministic_and_backward_is_linear(config2):
    g, cam, cot, o = config2
    o2 = util.run_ours(g, cam, (2.0 * cot[0], 2.0 * cot[1]), export=False)
    for k in ("color", "depth", "opacity", "radii", "n_touched"):
        np.testing.assert_array_equal(o[k], o2[k])                   # bitwise repeatable forward
    for k in ("g_dL_dmeans3D", "g_dL_dopacity", "g_dL_dsh", "g_dL_dscales", "g_dL_drotations", "g_dL_dtau_sum"):
        util.assert_grad_close(k, o2[k], 2.0 * o[k], rtol=1e-3, atol_scale=1e-4)   # linear in the cotangents


def _assert_matches_unmodified_reference(g, cam, cot, o):
    """`o` (util.run_ours with export) against the UNMODIFIED reference rasterizer built for sm_100 (oracle/_ref) on the
    same inputs: integer state bit-exact, images rtol 1e-4, every gradient incl. pose rtol 1e-3 (north_star's bars)."""
    gt, ct = ref_api.to_cuda(g, DEV), ref_api.to_cuda(cam, DEV)
    ref = ref_api.RefRasterizer()
    out = ref.forward(gt, ct)
    st = ref.export_state()
    gr = ref.backward(torch.from_numpy(cot[0]).to(DEV), torch.from_numpy(cot[1]).to(DEV))
    n = lambda t: t.cpu().numpy()
    np.testing.assert_array_equal(o["radii"], n(out["radii"]))
    assert int(o["num_rendered"]) == int(out["num_rendered"])
    np.testing.assert_array_equal(o["st_point_list_keys"].view(np.uint64), n(st["point_list_keys"]).view(np.uint64))
    np.testing.assert_array_equal(o["st_point_list"], n(st["point_list"]))
    np.testing.assert_array_equal(o["st_ranges"], n(st["ranges"]))
    np.testing.assert_array_equal(o["st_n_contrib"], n(st["n_contrib"]))
    np.testing.assert_array_equal(o["n_touched"], n(out["n_touched"]))
    np.testing.assert_array_equal(util.bits(o["st_final_T"]), util.bits(n(st["final_T"])))
    for k in ("color", "depth", "opacity"):
        np.testing.assert_allclose(o[k], n(out[k]), rtol=1e-4, atol=1e-7, err_msg=k)
    util.assert_grad_close("means3D", o["g_dL_dmeans3D"], n(gr["dL_dmeans3D"]))
    util.assert_grad_close("means2D", o["g_dL_dmeans2D"], n(gr["dL_dmeans2D"]))
    util.assert_grad_close("opacity", o["g_dL_dopacity"], n(gr["dL_dopacity"]))
    util.assert_grad_close("sh", o["g_dL_dsh"], n(gr["dL_dsh"]))
    util.assert_grad_close("scales", o["g_dL_dscales"], n(gr["dL_dscales"]))
    util.assert_grad_close("rotations", o["g_dL_drotations"], n(gr["dL_drotations"]))
    util.assert_grad_close("tau", o["g_dL_dtau_sum"], n(gr["dL_dtau_sum"]))
    return int(out["num_rendered"])


@pytest.mark.skipif(not ref_api.available(), reason="oracle/_ref/libref_rasterizer.so not built")
def test_config2_matches_unmodified_reference(config2):
    g, cam, cot, o = config2
    _assert_matches_unmodified_reference(g, cam, cot, o)


@pytest.mark.skipif(not ref_api.available(), reason="oracle/_ref/libref_rasterizer.so not built")
def test_config3_view_matches_unmodified_reference():
    """BASELINE.json config 3 at FULL size: one keyframe (view 3 of the 10) of the 1 M-Gaussian mapping window.
    ~1840 list entries per tile: the 32 privatised histogram copies, 32-bit offsets and the long-list paths at scale."""
    g, cams, cots = scenes.make_scene("config3")
    assert g["means3D"].shape[0] == 1_000_000 and len(cams) == 10
    o = util.run_ours(g, cams[3], cots[3])
    R = _assert_matches_unmodified_reference(g, cams[3], cots[3], o)
    assert R > 2_000_000


@pytest.mark.skipif(not ref_api.available(), reason="oracle/_ref/libref_rasterizer.so not built")
def test_config4_full_size_sh3_1080p_matches_reference():
    """BASELINE.json config 4 at FULL size: 3 M Gaussians, 1920x1080 (8160 tiles, 45 sort bits), SH degree 3."""
    g, cams, cots = scenes.make_scene("config4")
    assert g["means3D"].shape[0] == 3_000_000 and g["sh"].shape[1] == 16
    o = util.run_ours(g, cams[0], cots[0])
    R = _assert_matches_unmodified_reference(g, cams[0], cots[0], o)
    assert R > 5_000_000


@pytest.mark.skipif(not ref_api.available(), reason="oracle/_ref/libref_rasterizer.so not built")
def test_integration_md_stub_drives_the_library():
    """The reference-side binding printed in INTEGRATION.md section 3 (ctypes stubs standing in for
    _C.rasterize_gaussians / _C.rasterize_gaussians_backward) is executed as it stands and must give the drop-in
    package's results bit for bit (same kernels, same inputs)."""
    import re
    doc = open(os.path.join(util.ROOT, "INTEGRATION.md")).read()
    blocks = {b.split("\n", 1)[0]: b for b in re.findall(r"```python\n(.*?)```", doc, re.S)}
    from wildgs_slam_blackwell_b200 import _lib
    os.environ["WGS_LIB"] = _lib.LIB_PATH
    ns = {}
    exec([b for k, b in blocks.items() if k.startswith("# --- wgs-abi-structs")][0], ns)
    exec([b for k, b in blocks.items() if k.startswith("# --- wgs-abi-calls")][0], ns)
    g, cams, cots = scenes.make_scene(20000, 200, 136, 1, seed=31)
    cam, cot = cams[0], cots[0]
    o = util.run_ours(g, cam, cot, export=False)
    t = lambda a: util.to_dev(a, DEV)
    e = torch.empty(0, device=DEV)
    bg, m3, op, sc, ro, sh = t(cam["bg"]), t(g["means3D"]), t(g["opacities"]), t(g["scales"]), t(g["rotations"]), t(g["sh"])
    view, proj, praw, campos = t(cam["viewmatrix"]), t(cam["projmatrix"]), t(cam["projmatrix_raw"]), t(cam["campos"])
    H, W = cam["height"], cam["width"]
    R, color, radii, geomB, binB, imgB, depth, opacity, n_touched = ns["rasterize_gaussians_C"](
        bg, m3, e, op, sc, ro, 1.0, e, view, proj, praw, cam["tanfovx"], cam["tanfovy"], H, W, sh, 1, campos, False, False)
    assert all(isinstance(b, torch.Tensor) and b.dtype == torch.uint8 for b in (geomB, binB, imgB))   # save_for_backward-able
    grads = ns["rasterize_gaussians_backward_C"](bg, m3, radii, e, sc, ro, 1.0, e, view, proj, praw, cam["tanfovx"],
                                                 cam["tanfovy"], t(cot[0]), t(cot[1]), sh, 1, campos, geomB, R, binB, imgB,
                                                 False)
    torch.cuda.synchronize()
    n = lambda x: x.cpu().numpy()
    for k, v in (("color", color), ("depth", depth), ("opacity", opacity), ("radii", radii), ("n_touched", n_touched)):
        np.testing.assert_array_equal(o[k], n(v), err_msg=k)
    dm2, dcol, dop, dm3, dcov, dsh, dsc, dro, dtau = grads
    assert dcol.numel() == 0 and dcov.numel() == 0 and dtau.shape == (6,)
    util.assert_grad_close("means3D", n(dm3), o["g_dL_dmeans3D"])
    util.assert_grad_close("means2D", n(dm2), o["g_dL_dmeans2D"])
    util.assert_grad_close("sh", n(dsh), o["g_dL_dsh"])
    util.assert_grad_close("opacity", n(dop), o["g_dL_dopacity"])
    util.assert_grad_close("scales", n(dsc), o["g_dL_dscales"])
    util.assert_grad_close("rotations", n(dro), o["g_dL_drotations"])
    util.assert_grad_close("tau", n(dtau), o["g_dL_dtau_sum"])


def test_two_streams_with_different_P_do_not_share_accumulators():
    """Two backwards in flight on two streams of one device, on models of different size: each must give the gradients
    it gives alone (the accumulator workspace / pose ticket are per (device, stream, P))."""
    scn = [scenes.make_scene(P, 160, 112, 0, seed=40 + i) for i, P in enumerate((30000, 17000))]
    alone = [util.run_ours(g, cams[0], cots[0], export=False) for g, cams, cots in scn]
    streams = [torch.cuda.Stream(DEV), torch.cuda.Stream(DEV)]
    leaves, outs = [], []
    for (g, cams, cots), st in zip(scn, streams):
        with torch.cuda.stream(st):
            leaf = lambda a: util.to_dev(a).requires_grad_(True)
            P = g["means3D"].shape[0]
            L = dict(m3=leaf(g["means3D"]), op=leaf(g["opacities"]), sh=leaf(g["sh"]), sc=leaf(g["scales"]), ro=leaf(g["rotations"]),
                     th=torch.zeros(3, device=DEV, requires_grad=True), rh=torch.zeros(3, device=DEV, requires_grad=True))
            m2 = torch.zeros(P, 3, device=DEV, requires_grad=True)
            color, radii, depth, opacity, nt = rasterizer.GaussianRasterizer(util.settings_from_cam(cams[0], 0))(
                L["m3"], m2, L["op"], shs=L["sh"], scales=L["sc"], rotations=L["ro"], theta=L["th"], rho=L["rh"])
            outs.append(((color * util.to_dev(cots[0][0])).sum() + (depth * util.to_dev(cots[0][1])).sum(), st))
            leaves.append(L)
    for _ in range(3):      # interleave the two backwards (and repeat: the workspaces must come back clean)
        for L in leaves:
            for t in L.values():
                t.grad = None
        for loss, st in outs:
            with torch.cuda.stream(st):
                loss.backward(retain_graph=True)
        torch.cuda.synchronize()
        for L, ref in zip(leaves, alone):
            util.assert_grad_close("means3D", L["m3"].grad.cpu().numpy(), ref["g_dL_dmeans3D"])
            util.assert_grad_close("opacity", L["op"].grad.cpu().numpy(), ref["g_dL_dopacity"])
            tau = np.concatenate([L["rh"].grad.cpu().numpy().reshape(-1), L["th"].grad.cpu().numpy().reshape(-1)])
            util.assert_grad_close("tau", tau, ref["g_dL_dtau_sum"])
    keys = [k for k in rasterizer._grad_ws if k[0] == 0]
    assert len({k[1] for k in keys}) >= 2 and len(keys) <= rasterizer._GRAD_WS_KEEP


def test_grad_sink_writes_and_accumulates_in_place():
    """rasterizer.grad_sink (used by the mapping window): gradients land in caller buffers, autograd gets
    None for them, accumulate=True adds; pose gradients still flow through autograd."""
    g, cams, cots = scenes.make_scene(3000, 128, 96, 0, seed=31)
    base = util.run_ours(g, cams[0], cots[0], export=False)
    s = util.settings_from_cam(cams[0], 0)
    leaf = lambda a: util.to_dev(a).requires_grad_(True)
    m3, op, sh, sc, ro = leaf(g["means3D"]), leaf(g["opacities"]), leaf(g["sh"]), leaf(g["scales"]), leaf(g["rotations"])
    m2 = torch.zeros(3000, 3, device=DEV, requires_grad=True)
    th, rh = torch.zeros(3, device=DEV, requires_grad=True), torch.zeros(3, device=DEV, requires_grad=True)
    flat = torch.full((3000 * 14,), 7.0, device=DEV)             # garbage: first pass must overwrite
    bufs, off = {}, 0
    for name, t in (("means3D", m3), ("sh", sh), ("opacity", op), ("scales", sc), ("rotations", ro)):
        bufs[name] = flat[off:off + t.numel()].view(t.shape)
        off += t.numel()
    dc, dd = util.to_dev(cots[0][0]), util.to_dev(cots[0][1])
    rast = rasterizer.GaussianRasterizer(s)
    for k in range(2):
        color, radii, depth, opacity, n_touched = rast(m3, m2, op, shs=sh, scales=sc, rotations=ro, theta=th, rho=rh)
        with rasterizer.grad_sink(bufs, accumulate=k > 0):
            ((color * dc).sum() + (depth * dd).sum()).backward()
        assert m3.grad is None and sh.grad is None and op.grad is None
        mult = float(k + 1)
        util.assert_grad_close("means3D", bufs["means3D"].cpu().numpy(), mult * base["g_dL_dmeans3D"])
        util.assert_grad_close("sh", bufs["sh"].cpu().numpy(), mult * base["g_dL_dsh"])
        util.assert_grad_close("opacity", bufs["opacity"].cpu().numpy(), mult * base["g_dL_dopacity"])
        util.assert_grad_close("scales", bufs["scales"].cpu().numpy(), mult * base["g_dL_dscales"])
        util.assert_grad_close("rotations", bufs["rotations"].cpu().numpy(), mult * base["g_dL_drotations"])
    tau = np.concatenate([rh.grad.cpu().numpy().reshape(-1), th.grad.cpu().numpy().reshape(-1)])
    util.assert_grad_close("tau", tau, 2.0 * base["g_dL_dtau_sum"])       # autograd accumulated the two passes
    with pytest.raises(RuntimeError, match="grad_sink buffer"):
        color, *_ = rast(m3, m2, op, shs=sh, scales=sc, rotations=ro, theta=th, rho=rh)
        with rasterizer.grad_sink({"means3D": torch.zeros(5, 3, device=DEV)}):
            color.sum().backward()


def test_window_trainer_sums_views_with_and_without_sink():
    """WindowTrainer on one GPU, 3 views: bucket == sum over views of the single-view gradients,
    whether autograd accumulates (.grad views) or the backward kernel accumulates (grad_sink)."""
    from wildgs_slam_blackwell_b200.window import WindowTrainer
    g, cams, cots = scenes.make_scene(4000, 160, 112, 0, seed=41, n_views=3)
    singles = [util.run_ours(g, cams[k], cots[k], export=False) for k in range(3)]
    want = {n: sum(s[key] for s in singles) for n, key in (("means3D", "g_dL_dmeans3D"), ("sh", "g_dL_dsh"),
            ("opacities", "g_dL_dopacity"), ("scales", "g_dL_dscales"), ("rotations", "g_dL_drotations"))}
    for sink in (None, {"means3D": "means3D", "sh": "sh", "opacities": "opacity", "scales": "scales",
                        "rotations": "rotations"}):
        params = {k: util.to_dev(g[k]).requires_grad_(True) for k in ("means3D", "sh", "opacities", "scales", "rotations")}
        tr = WindowTrainer(params, sink_names=sink)
        m2 = torch.zeros(4000, 3, device=DEV, requires_grad=True)
        poses = [(torch.zeros(3, device=DEV, requires_grad=True), torch.zeros(3, device=DEV, requires_grad=True)) for _ in range(3)]

        def render_loss(v):
            th, rh = poses[v]
            th.grad = rh.grad = None
            rast = rasterizer.GaussianRasterizer(util.settings_from_cam(cams[v], 0))
            color, radii, depth, opacity, n_touched = rast(params["means3D"], m2, params["opacities"], shs=params["sh"],
                                                           scales=params["scales"], rotations=params["rotations"], theta=th, rho=rh)
            loss = (color * util.to_dev(cots[v][0])).sum() + (depth * util.to_dev(cots[v][1])).sum()
            return loss, (lambda: torch.cat([rh.grad.reshape(-1), th.grad.reshape(-1)]))
        for _ in range(2):                      # second step re-uses the bucket: must not carry over
            out = tr.step(3, render_loss)
        for n in want:
            assert params[n].grad.data_ptr() == tr.bucket.views[n].data_ptr()
            util.assert_grad_close(n, params[n].grad.cpu().numpy(), want[n])
        for v in range(3):
            util.assert_grad_close(f"tau{v}", out["pose_grads"][v].cpu().numpy(), singles[v]["g_dL_dtau_sum"])
        assert out["assignment"] == [[0, 1, 2]]


def test_more_than_8192_tiles_uses_single_block_scan():
    """2048x1100 = 8832 tiles: beyond the 8-CTA cluster scan, exercises the fallback scan kernel."""
    g, cams, cots = scenes.make_scene(20000, 2048, 1100, 0, seed=51)
    o = cpu_oracle.forward_backward(g, cams[0], None, export=True)
    ours = util.run_ours(g, cams[0], None)
    assert int(ours["num_rendered"]) == int(o["num_rendered"])
    np.testing.assert_array_equal(ours["st_ranges"].view(np.uint32), o["st_ranges"])
    np.testing.assert_array_equal(ours["st_point_list"].view(np.uint32), o["st_point_list"])
    np.testing.assert_array_equal(ours["radii"], o["radii"])


def test_tile_row_strips_sum_to_the_whole_image():
    """rasterizer.tile_rows: strips of a partition of the image reproduce the whole-image forward bitwise
    inside their rows, and their n_touched / gradients (incl. pose) sum to the whole-image results."""
    g, cams, cots = scenes.make_scene(6000, 200, 150, 1, seed=61)     # 13 x 10 tiles, ragged edges
    cam, cot = cams[0], cots[0]
    full = util.run_ours(g, cam, cot, export=False)
    s = util.settings_from_cam(cam, 1)
    dc, dd = util.to_dev(cot[0]), util.to_dev(cot[1])
    acc = None
    for r0, r1 in ((0, 3), (3, 4), (4, 10)):
        leaf = lambda a: util.to_dev(a).requires_grad_(True)
        m3, op, sh, sc, ro = leaf(g["means3D"]), leaf(g["opacities"]), leaf(g["sh"]), leaf(g["scales"]), leaf(g["rotations"])
        m2 = torch.zeros(6000, 3, device=DEV, requires_grad=True)
        th, rh = torch.zeros(3, device=DEV, requires_grad=True), torch.zeros(3, device=DEV, requires_grad=True)
        with rasterizer.tile_rows(r0, r1):
            color, radii, depth, opacity, n_touched = rasterizer.GaussianRasterizer(s)(
                m3, m2, op, shs=sh, scales=sc, rotations=ro, theta=th, rho=rh)
        ((color * dc).sum() + (depth * dd).sum()).backward()          # backward outside the context: strip is remembered
        y0, y1 = r0 * 16, min(r1 * 16, 150)
        c = color.detach().cpu().numpy()
        np.testing.assert_array_equal(c[:, y0:y1], full["color"][:, y0:y1])
        assert not c[:, :y0].any() and not c[:, y1:].any()
        np.testing.assert_array_equal(depth.detach().cpu().numpy()[:, y0:y1], full["depth"][:, y0:y1])
        np.testing.assert_array_equal(radii.cpu().numpy(), full["radii"])
        part = dict(n_touched=n_touched.cpu().numpy().astype(np.int64), m3=m3.grad.cpu().numpy(), op=op.grad.cpu().numpy(),
                    sh=sh.grad.cpu().numpy(), sc=sc.grad.cpu().numpy(), ro=ro.grad.cpu().numpy(), m2=m2.grad.cpu().numpy(),
                    tau=np.concatenate([rh.grad.cpu().numpy().reshape(-1), th.grad.cpu().numpy().reshape(-1)]))
        acc = part if acc is None else {k: acc[k] + part[k] for k in acc}
    np.testing.assert_array_equal(acc["n_touched"], full["n_touched"])
    for k, ref in (("m3", "g_dL_dmeans3D"), ("op", "g_dL_dopacity"), ("sh", "g_dL_dsh"), ("sc", "g_dL_dscales"),
                   ("ro", "g_dL_drotations"), ("m2", "g_dL_dmeans2D"), ("tau", "g_dL_dtau_sum")):
        util.assert_grad_close(k, acc[k], full[ref])
    with pytest.raises(RuntimeError, match="tile_rows"):
        with rasterizer.tile_rows(4, 11):
            rasterizer.GaussianRasterizer(s)(m3, m2, op, shs=sh, scales=sc, rotations=ro, theta=th, rho=rh)


@pytest.mark.parametrize("P,deg,active", [(5003, 1, 1), (5003, 2, 2), (2999, 3, 3), (4001, 3, 1), (37, 2, 0)])
def test_staged_sh_rows_match_cpu_oracle(P, deg, active):
    """SH degrees 1-3 (M = 4, 9, 16) go through the shared-memory staging of preprocess fwd/bwd: ragged P (partial
    last warp and block), active degree below the stored one (inactive bands get zero gradients), and the
    accumulating write-back (grad_sink) -- all against the C oracle."""
    g, cams, cots = scenes.make_scene(P, 176, 120, deg, seed=70 + deg)
    g = dict(g, sh_degree=active)
    ours = util.run_ours(g, cams[0], cots[0])
    o = cpu_oracle.forward_backward(g, cams[0], cots[0], export=True)
    np.testing.assert_array_equal(ours["radii"], o["radii"])
    np.testing.assert_array_equal(ours["st_point_list"].view(np.uint32), o["st_point_list"])
    vis = o["radii"] > 0
    np.testing.assert_allclose(ours["st_rgb"][vis], o["st_rgb"][vis], rtol=1e-5, atol=1e-6)      # SH -> RGB
    assert (ours["st_clamped"][vis].astype(bool) != o["st_clamped"][vis].astype(bool)).mean() < 1e-3   # rgb within 1 ulp of 0
    for k in ("g_dL_dsh", "g_dL_dmeans3D"):
        ref = o[k]
        bad = np.abs(ours[k] - ref) > 1e-3 * np.abs(ref) + 1e-4 * np.abs(ref).max()
        assert bad.mean() < 2e-3, k
    M = (deg + 1) ** 2
    assert ours["g_dL_dsh"].shape == (P, M, 3)
    assert not ours["g_dL_dsh"][:, (active + 1) ** 2:].any()         # inactive bands
    assert not ours["g_dL_dsh"][~vis].any()                          # culled rows are written as zeros
    # accumulating write-back: second backward into the same buffers doubles them
    s = util.settings_from_cam(cams[0], active)
    leaf = lambda a: util.to_dev(a).requires_grad_(True)
    m3, op, sh, sc, ro = leaf(g["means3D"]), leaf(g["opacities"]), leaf(g["sh"]), leaf(g["scales"]), leaf(g["rotations"])
    m2 = torch.zeros(P, 3, device=DEV, requires_grad=True)
    bufs = {"means3D": torch.full_like(m3, 3.0), "sh": torch.full_like(sh, 3.0), "opacity": torch.full_like(op, 3.0),
            "scales": torch.full_like(sc, 3.0), "rotations": torch.full_like(ro, 3.0)}
    dc, dd = util.to_dev(cots[0][0]), util.to_dev(cots[0][1])
    rast = rasterizer.GaussianRasterizer(s)
    for k in range(2):
        color, radii, depth, opacity, n_touched = rast(m3, m2, op, shs=sh, scales=sc, rotations=ro)
        with rasterizer.grad_sink(bufs, accumulate=k > 0):
            ((color * dc).sum() + (depth * dd).sum()).backward()
        util.assert_grad_close("sh", bufs["sh"].cpu().numpy(), (k + 1.0) * ours["g_dL_dsh"])


def test_fused_densification_statistics_match_the_mapper_procedure():
    """rasterizer.densification_stats / WindowTrainer(densify_stats=True): the per-Gaussian backward kernel accumulates
    what the mapper computes after each view's backward (mapper.py:1176-1182, gaussian_model.py:745-749):
    sum over views of ||dL_dmeans2D[:, :2]|| and of the visibility counts, running max of the radii."""
    from wildgs_slam_blackwell_b200.window import WindowTrainer
    P = 4000
    g, cams, cots = scenes.make_scene(P, 160, 112, 0, seed=43, n_views=3)
    singles = [util.run_ours(g, cams[k], cots[k], export=False) for k in range(3)]
    accum, denom, maxr = np.zeros((P, 1), np.float32), np.zeros((P, 1), np.float32), np.zeros(P, np.float32)
    for s in singles:                                   # the reference's procedure, view by view
        vis = s["radii"] > 0
        accum[vis] += np.linalg.norm(s["g_dL_dmeans2D"][vis, :2], axis=-1, keepdims=True)
        denom[vis] += 1
        maxr[vis] = np.maximum(maxr[vis], s["radii"][vis])
    params = {k: util.to_dev(g[k]).requires_grad_(True) for k in ("means3D", "sh", "opacities", "scales", "rotations")}
    tr = WindowTrainer(params, sink_names={"means3D": "means3D", "sh": "sh", "opacities": "opacity", "scales": "scales",
                                           "rotations": "rotations"}, densify_stats=True)
    m2 = torch.zeros(P, 3, device=DEV, requires_grad=True)

    def render_loss(v):
        rast = rasterizer.GaussianRasterizer(util.settings_from_cam(cams[v], 0))
        color, radii, depth, opacity, n_touched = rast(params["means3D"], m2, params["opacities"], shs=params["sh"],
                                                       scales=params["scales"], rotations=params["rotations"])
        return (color * util.to_dev(cots[v][0])).sum() + (depth * util.to_dev(cots[v][1])).sum(), None
    for _ in range(2):                                  # per-window sums: the second step must not carry over
        tr.step(3, render_loss)
    util.assert_grad_close("xyz_gradient_accum", tr.grad_accum.cpu().numpy(), accum)
    np.testing.assert_array_equal(tr.denom.cpu().numpy(), denom)
    np.testing.assert_array_equal(tr.reduce_max_radii().cpu().numpy(), maxr)
    with pytest.raises(RuntimeError, match="whole-image"):
        with rasterizer.tile_rows(0, 2), rasterizer.densification_stats(tr.grad_accum, tr.denom, tr.max_radii):
            render_loss(0)[0].backward()


@pytest.mark.parametrize("streams", [1, 2])
def test_window_shares_of_emulated_ranks_sum_to_the_window_gradient(streams):
    """The work units WindowTrainer hands to each of 4 ranks (whole views + tile-row strips of the remainder, on one or
    two streams with atomic accumulation, block-compacted preprocess_bwd), run one rank after the other on this GPU, sum
    to the gradient of the whole 5-view window -- eagerly and with a rank's step captured in one CUDA graph."""
    sys.path.insert(0, os.path.join(util.ROOT, "scripts"))
    from window_workload import LEAVES, WindowWorkload
    scenes.CONFIGS["_w5"] = (40_000, 320, 240, 0, 5)
    scene = scenes.make_scene(40_000, 320, 240, 0, seed=91, n_views=5)
    ref_wl = WindowWorkload("_w5", torch.device(DEV), peer="off", streams=1, use_graph=False, scene=scene)
    ref = {k: torch.zeros_like(p) for k, p in ref_wl.params.items()}
    for v in range(5):
        loss, _ = ref_wl.render_loss(v)
        for k, gk in zip(LEAVES, torch.autograd.grad(loss, [ref_wl.params[k] for k in LEAVES])):
            ref[k] += gk
    for plan, use_graph in (("views", False), ("strips", False), ("balanced", False), ("balanced", True)):
        total = {k: torch.zeros_like(p) for k, p in ref.items()}
        units = 0
        for r in range(4):
            wl = WindowWorkload("_w5", torch.device(DEV), peer="off", plan=plan, streams=streams, use_graph=use_graph,
                                scene=scene, emulate=(r, 4))
            out = wl.step()
            if use_graph:
                out = wl.step()
                wl.check()
            torch.cuda.synchronize()
            units += len(out["assignment"][r])
            for k in total:
                total[k] += wl.params[k].grad
        assert units == (5 if plan == "views" else 8 if plan == "strips" else 8), (plan, units)
        for k in ref:
            util.assert_grad_close(f"{plan}/{k}", total[k].cpu().numpy(), ref[k].cpu().numpy())
