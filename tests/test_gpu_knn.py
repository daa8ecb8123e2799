"""GPU: distCUDA2 (grid kNN kernels through the drop-in simple_knn._C and the C ABI) against the CPU oracle
(exact 3-NN) and, where oracle/_ref/libref_simple_knn.so was built, against the UNMODIFIED reference."""
import ctypes as C
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import knn_oracle  # noqa: E402
from simple_knn._C import distCUDA2  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref_simple_knn.so")


def clouds(name, P, seed):
    rng = np.random.default_rng(seed)
    if name == "uniform":
        return rng.uniform(-3, 3, (P, 3)).astype(np.float32)
    if name == "surface":      # back-projected depth map: a thin sheet, strongly non-uniform in the grid
        u, v = rng.uniform(-1, 1, P), rng.uniform(-0.75, 0.75, P)
        z = 2.0 + 0.5 * np.sin(3 * u) * np.cos(2 * v) + rng.normal(0, 0.002, P)
        return np.stack([u * z, v * z, z], 1).astype(np.float32)
    if name == "clusters":     # dense blobs + far outliers
        c = rng.normal(0, 5, (8, 3))
        pts = c[rng.integers(0, 8, P)] + rng.normal(0, 0.05, (P, 3))
        pts[:5] = rng.uniform(-500, 500, (5, 3))
        return pts.astype(np.float32)
    if name == "duplicates":
        base = rng.uniform(0, 1, (P // 4 + 1, 3))
        return base[rng.integers(0, len(base), P)].astype(np.float32)
    if name == "line":         # degenerate extent in two axes
        return np.stack([rng.uniform(0, 1, P), np.zeros(P), np.full(P, 2.0)], 1).astype(np.float32)
    raise KeyError(name)


@pytest.mark.parametrize("name,P", [("uniform", 50_000), ("surface", 100_000), ("clusters", 20_000),
                                    ("duplicates", 5_000), ("line", 3_000), ("uniform", 7), ("uniform", 4)])
def test_matches_exact_knn_oracle(name, P):
    pts = clouds(name, P, 17)
    ours = distCUDA2(torch.from_numpy(pts).to(DEV)).cpu().numpy()
    ref = knn_oracle.dist_cuda2(pts)
    np.testing.assert_allclose(ours, ref, rtol=2e-6, atol=1e-12)


def test_fewer_than_four_points_and_errors():
    for P in (1, 2, 3):
        pts = clouds("uniform", P, 3)
        ours = distCUDA2(torch.from_numpy(pts).to(DEV)).cpu().numpy()
        ref = knn_oracle.dist_cuda2(pts)
        assert np.array_equal(np.isinf(ours), np.isinf(ref)) and np.all(ours[np.isfinite(ours)] > 1e37)
    assert distCUDA2(torch.zeros(0, 3, device=DEV)).shape == (0,)
    with pytest.raises(RuntimeError, match="CUDA"):
        distCUDA2(torch.zeros(5, 3))
    with pytest.raises(RuntimeError, match="num_points, 3"):
        distCUDA2(torch.zeros(5, 2, device=DEV))


@pytest.mark.skipif(not os.path.exists(REF_SO), reason="oracle/_ref/libref_simple_knn.so not built")
@pytest.mark.parametrize("name,P", [("uniform", 50_000), ("surface", 100_000), ("clusters", 20_000), ("duplicates", 5_000)])
def test_matches_unmodified_reference(name, P):
    ref_lib = C.CDLL(REF_SO)
    ref_lib.ref_dist_cuda2.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
    pts = torch.from_numpy(clouds(name, P, 23)).to(DEV)
    out = torch.zeros(P, device=DEV)
    torch.cuda.synchronize()
    assert ref_lib.ref_dist_cuda2(P, pts.data_ptr(), out.data_ptr()) == 0
    ours = distCUDA2(pts)
    # same neighbours, same fp32 expression: equal up to the contraction the two compilations chose
    np.testing.assert_allclose(ours.cpu().numpy(), out.cpu().numpy(), rtol=1e-6, atol=1e-12)
    np.testing.assert_allclose(knn_oracle.dist_cuda2(pts.cpu().numpy()), out.cpu().numpy(), rtol=2e-6, atol=1e-12)
