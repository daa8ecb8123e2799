"""CPU: the numpy loss oracle (oracle/loss_oracle.py) against the golden vectors produced by the
REFERENCE's own get_loss_tracking / get_loss_mapping (tests/golden/make_loss_golden.py), the ctypes
struct against include/wgs_loss.h, and the exported symbols."""
import ctypes as C
import os
import re
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import loss_oracle  # noqa: E402
from wildgs_slam_blackwell_b200 import _lib, loss  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
KINDS = ("tracking", "tracking_unc", "mapping", "mapping_init")


def oracle_for(g):
    kind = str(g["kind"])
    kw = dict(rgb_boundary_threshold=float(g["threshold"]))
    if kind.startswith("tracking"):
        kw.update(opacity=g["opacity"], pixel_mask=g["grad_mask"].astype(np.float32), exposure_a=g["a"],
                  exposure_b=g["b"], color_weight=1.0, depth_weight=0.0)
        if kind == "tracking_unc":
            kw["uncertainty"] = g["unc"]
    else:
        alpha = float(g["alpha"])
        kw.update(depth=g["depth"], gt_depth=g["gt_depth"], color_weight=alpha, depth_weight=1.0 - alpha)
        if kind == "mapping":
            kw.update(exposure_a=g["a"], exposure_b=g["b"])
    return kw


@pytest.mark.parametrize("kind", KINDS)
def test_oracle_matches_reference_golden(kind):
    g = dict(np.load(os.path.join(GOLDEN, f"loss_{kind}.npz")))
    o = loss_oracle.photometric_loss(g["image"], g["gt"], **oracle_for(g))
    np.testing.assert_allclose(o["loss"], g["loss"], rtol=2e-6)
    np.testing.assert_allclose(o["dL_dcolor"], g["g_image"], rtol=1e-5, atol=1e-10)
    np.testing.assert_allclose(o["dL_ddepth"], g["g_depth"], rtol=1e-5, atol=1e-10)
    if kind != "mapping_init":
        np.testing.assert_allclose(o["dL_da"], g["g_a"][0], rtol=2e-4)
        np.testing.assert_allclose(o["dL_db"], g["g_b"][0], rtol=2e-4, atol=1e-7)


def test_loss_abi_symbols_and_struct():
    header = open(os.path.join(ROOT, "include", "wgs_loss.h")).read()
    declared = set(re.findall(r"\b(wgs_[a-z_]+)\s*\(", header)) - {"wgs_backward", "wgs_last_error"}
    assert declared == set(loss.EXPORTS), declared ^ set(loss.EXPORTS)
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), name
    body = re.search(r"typedef struct WgsLossArgs \{(.*?)\} WgsLossArgs;", header, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        for part in decl.strip().split(","):
            if part.strip():
                names.append(re.findall(r"[A-Za-z_0-9]+", part)[-1])
    assert names == [f[0] for f in loss.LossArgs._fields_]
    assert loss._lib_loss().wgs_loss_workspace_bytes(640, 480) >= 256 + 300 * 32
    assert loss._lib_loss().wgs_photometric_loss(None, None) == _lib.WGS_ERR_INVALID_ARGUMENT


def test_no_cpu_fallback():
    import torch
    with pytest.raises(RuntimeError, match="CUDA"):
        loss.photometric_loss(torch.zeros(3, 4, 4), torch.zeros(3, 4, 4))


def test_knn_abi_symbols_and_oracle_bruteforce():
    from oracle import knn_oracle
    from wildgs_slam_blackwell_b200 import knn
    header = open(os.path.join(ROOT, "include", "wgs_knn.h")).read()
    declared = set(re.findall(r"\b(wgs_[a-z_0-9]+)\s*\(", header))
    assert declared == set(knn.EXPORTS), declared ^ set(knn.EXPORTS)
    lib = knn._lib_knn()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.wgs_knn_workspace_bytes(100000) > 100000 * 20
    assert lib.wgs_dist_cuda2(-1, None, None, None, None) == _lib.WGS_ERR_INVALID_ARGUMENT
    rng = np.random.default_rng(0)
    pts = rng.uniform(0, 1, (300, 3)).astype(np.float32)
    pts[10] = pts[11]                                               # a duplicate pair
    d = pts[:, None, :] - pts[None, :, :]
    d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1] + d[..., 2] * d[..., 2]).astype(np.float32)
    np.fill_diagonal(d2, np.inf)
    d2.sort(axis=1)
    brute = (d2[:, 0] + d2[:, 1] + d2[:, 2]) / np.float32(3.0)
    np.testing.assert_allclose(knn_oracle.dist_cuda2(pts), brute, rtol=1e-6)


def test_adam_and_peer_abi_symbols():
    from wildgs_slam_blackwell_b200 import adam
    header = open(os.path.join(ROOT, "include", "wgs_adam.h")).read()
    declared = set(re.findall(r"\b(wgs_[a-z_0-9]+)\s*\(", header))
    assert declared == set(adam.EXPORTS)
    lib = adam._lib_adam()
    assert hasattr(lib, "wgs_adam_step")
    body = re.search(r"typedef struct WgsAdamTensor \{(.*?)\} WgsAdamTensor;", header, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = [re.findall(r"[A-Za-z_0-9]+", d)[-1] for d in body.split(";") if d.strip()]
    assert names == [f[0] for f in adam.AdamTensor._fields_]
    assert lib.wgs_adam_step(None, 3, 0.9, 0.999, 1e-8, None) == _lib.WGS_ERR_INVALID_ARGUMENT
    assert lib.wgs_adam_step(None, 0, 0.9, 0.999, 1e-8, None) == _lib.WGS_OK


def test_pose_abi_symbols_and_structs():
    from wildgs_slam_blackwell_b200 import pose
    header = open(os.path.join(ROOT, "include", "wgs_pose.h")).read()
    declared = set(re.findall(r"\b(wgs_[a-z_0-9]+)\s*\(", header)) - {"wgs_camera_matrices"}
    assert declared == set(pose.EXPORTS), declared
    assert hasattr(pose._lib_pose(), "wgs_pose_step")
    for cname, cls in (("WgsPoseAdam", pose.PoseAdam), ("WgsPoseStepArgs", pose.PoseStepArgs)):
        body = re.search(r"typedef struct " + cname + r" \{(.*?)\} " + cname + ";", header, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = []
        for decl in body.split(";"):
            for part in decl.strip().split(","):
                if part.strip():
                    names.append(re.findall(r"[A-Za-z_0-9]+", part)[-1])
        assert names == [f[0] for f in cls._fields_], cname
    assert pose._lib_pose().wgs_pose_step(None, None) == _lib.WGS_ERR_INVALID_ARGUMENT


def test_compact_abi_symbols():
    from wildgs_slam_blackwell_b200 import compaction
    header = open(os.path.join(ROOT, "include", "wgs_compact.h")).read()
    declared = set(re.findall(r"\b(wgs_[a-z_0-9]+)\s*\(", header))
    assert declared == set(compaction.EXPORTS), declared
    lib = compaction._lib_compact()
    for name in declared:
        assert hasattr(lib, name)
    body = re.search(r"typedef struct WgsCompactTensor \{(.*?)\} WgsCompactTensor;", header, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = [re.findall(r"[A-Za-z_0-9]+", d)[-1] for d in body.split(";") if d.strip()]
    assert names == [f[0] for f in compaction.CompactTensor._fields_]
    assert lib.wgs_compact_workspace_bytes(3_000_000) >= 2930 * 4
    assert lib.wgs_compact_plan(-1, None, None, None, None, None) == _lib.WGS_ERR_INVALID_ARGUMENT


@pytest.mark.parametrize("name", ["ssim_a", "ssim_b", "ssim_batch"])
def test_ssim_oracle_matches_reference_golden(name):
    """oracle/ssim_oracle.py (numpy float64) against vectors made by the reference's loss_utils.ssim (+ autograd) and
    mapping_utils.compute_ssim_components (tests/golden/make_ssim_golden.py)."""
    from oracle import ssim_oracle
    d = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    a, b = d["img1"], d["img2"]
    batched = a.ndim == 4
    val = ssim_oracle.ssim(a, b, 11, size_average=not batched)
    np.testing.assert_allclose(val, d["ssim"], rtol=2e-5, atol=2e-6)
    items = list(zip(a, b, d["dmean"], d["grad"])) if batched else [(a, b, d["dmean"], d["grad"])]
    for x, y, g, want in items:
        got = float(g) * ssim_oracle.ssim_grad(x, y, 11)
        np.testing.assert_allclose(got, want, rtol=2e-3, atol=2e-4 * np.abs(want).max())   # the golden is float32 torch
    for ws in (7, 11):
        for x, y, k in (zip(a, b, range(len(a))) if batched else [(a, b, None)]):
            lum, con, stu = ssim_oracle.components(y, x, ws)
            pick = (lambda t: t[k]) if batched else (lambda t: t)
            np.testing.assert_allclose(lum, pick(d[f"lum{ws}"]), rtol=1e-4, atol=1e-5)
            # variances are E[x^2] - mu^2 in float32 in the reference: on smooth images the cancellation noise
            # (~1e-7 against variances of ~1e-4) shows in contrast and structure at the 1e-3 level
            np.testing.assert_allclose(con, pick(d[f"con{ws}"]), rtol=1e-4, atol=5e-3)
            bad = np.abs(stu - pick(d[f"str{ws}"])) > 5e-3      # ill-conditioned where both variances vanish
            assert bad.mean() < 5e-3


def test_ssim_abi_symbols():
    from wildgs_slam_blackwell_b200 import loss_utils
    header = open(os.path.join(ROOT, "include", "wgs_ssim.h")).read()
    code = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(wgs_[a-z_0-9]+)\s*\(", code))
    assert declared == set(loss_utils.EXPORTS), declared
    lib = loss_utils._lib_ssim()
    for name in declared:
        assert hasattr(lib, name)
    body = re.search(r"typedef struct WgsSsimArgs \{(.*?)\} WgsSsimArgs;", code, re.S).group(1)
    names = [n for d in body.split(";") if d.strip() for n in
             [x.strip().lstrip("*").strip() for x in re.sub(r"^\s*(const\s+)?\w+\s*\*?", "", d.strip(), count=1).split(",")]]
    assert names == [f[0] for f in loss_utils.SsimArgs._fields_], names
    assert lib.wgs_ssim_workspace_bytes(640, 480, 3) >= 40 * 30 * 3 * 8
    a = loss_utils.SsimArgs(0, 0, 0, 0, 11, None, None, None, None, None, None, None)
    assert lib.wgs_ssim_forward(C.byref(a), None) == _lib.WGS_ERR_INVALID_ARGUMENT


def _make_net(seed, fd=12):
    import torch
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(fd, 16), torch.nn.ReLU(), torch.nn.Linear(16, 1), torch.nn.Softplus(),
                               torch.nn.Flatten(-2, -1))


def test_mapping_ssim_oracle_matches_reference_golden():
    d = np.load(os.path.join(ROOT, "tests", "golden", "mapping_ssim.npz"))
    o = loss_oracle.mapping_rgbd(d["image"], d["gt"], d["depth"], d["gt_depth"], d["a"], d["b"], alpha=0.5, lam=0.2)
    np.testing.assert_allclose(o["loss"], d["loss"], rtol=1e-5)
    for k in ("g_image", "g_depth"):
        np.testing.assert_allclose(o[k], d[k], rtol=2e-3, atol=2e-4 * np.abs(d[k]).max())
    np.testing.assert_allclose([o["g_a"], o["g_b"]], [d["g_a"][0], d["g_b"][0]], rtol=2e-3)


@pytest.mark.parametrize("name,tf,sf,init,freeze", [("mapping_uncertainty", 0.3, 0.6, False, False),
                                                    ("mapping_uncertainty_init", 0.9, 0.1, True, True)])
def test_mapping_uncertainty_oracle_matches_reference_golden(name, tf, sf, init, freeze):
    import torch
    d = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    cfg = dict(alpha=0.5, lam=0.2, thr=0.01, window=7, median=5, opacity_th=0.9, ssim_mult=0.5, depth_mult=0.2, freeze=freeze)
    a, b = (None, None) if init else (d["a"], d["b"])
    o = loss_oracle.mapping_uncertainty(d["image"], d["gt"], d["depth"], d["gt_depth"], d["opacity"], d["uncertainty"],
                                        cfg, tf, sf, a, b)
    np.testing.assert_allclose(o["loss"], d["loss"], rtol=2e-4)      # the uncertainty term carries the float32 SSIM components
    for k in ("g_image", "g_depth"):
        np.testing.assert_allclose(o[k], d[k], rtol=2e-3, atol=2e-4 * np.abs(d[k]).max())
    if not init:
        np.testing.assert_allclose([o["g_a"], o["g_b"]], [d["g_a"][0], d["g_b"][0]], rtol=2e-3)
    # gradient w.r.t. the uncertainty map, pushed through the same stand-in network as the generator used
    net = _make_net(7)
    unc = net(torch.from_numpy(d["features"]))
    np.testing.assert_allclose(unc.detach().numpy(), d["uncertainty"], rtol=1e-5, atol=1e-6)
    unc.backward(torch.from_numpy(o["g_unc"]).float())
    for i, p in enumerate(net.parameters()):
        want = d[f"g_net{i}"]
        got = np.zeros_like(want) if p.grad is None else p.grad.numpy()
        np.testing.assert_allclose(got, want, rtol=1e-2, atol=2e-3 * max(np.abs(want).max(), 1e-12))


def test_uncertainty_abi_symbols():
    header = open(os.path.join(ROOT, "include", "wgs_uncertainty.h")).read()
    code = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(wgs_[a-z_0-9]+)\s*\(", code))
    assert declared == set(loss.EXPORTS_UNCERTAINTY), declared
    lib = loss._lib_loss()
    for name in declared:
        assert hasattr(lib, name)
    body = re.search(r"typedef struct WgsUncertaintyLossArgs \{(.*?)\} WgsUncertaintyLossArgs;", code, re.S).group(1)
    names = [n for d in body.split(";") if d.strip() for n in
             [x.strip().lstrip("*").strip() for x in re.sub(r"^\s*(const\s+)?\w+\s*\*?", "", d.strip(), count=1).split(",")]]
    assert names == [f[0] for f in loss.UncertaintyLossArgs._fields_], names
    assert lib.wgs_uncertainty_workspace_bytes(45, 34) >= 4 * 45 * 34 * 4
    a = loss.UncertaintyLossArgs()
    assert lib.wgs_uncertainty_loss(C.byref(a), None) == _lib.WGS_ERR_INVALID_ARGUMENT
