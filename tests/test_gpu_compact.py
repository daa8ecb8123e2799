"""GPU: fused row compaction (wgs_compact_plan / wgs_compact_rows) against torch boolean indexing -- pure data
movement, so equality is exact -- and prune_points against the reference's procedure
(gaussian_model.py:526-564) executed with torch ops."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wildgs_slam_blackwell_b200 import compaction  # noqa: E402
from wildgs_slam_blackwell_b200.adam import FusedAdam  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("P,frac", [(100_003, 0.7), (0, 0.5), (1, 1.0), (1, 0.0), (5000, 0.0), (5000, 1.0), (1025, 0.5), (2_000_000, 0.9)])
def test_compact_rows_equals_boolean_indexing(P, frac):
    g = torch.Generator(device=DEV).manual_seed(P)
    keep = torch.rand(P, device=DEV, generator=g) < frac
    tensors = [torch.randn(P, 3, device=DEV, generator=g), torch.randn(P, 15, 3, device=DEV, generator=g),
               torch.randn(P, 1, device=DEV, generator=g), torch.randn(P, 4, device=DEV, generator=g),
               torch.randint(0, 1 << 30, (P,), device=DEV, generator=g, dtype=torch.int32),
               torch.randn(P, device=DEV, generator=g)]
    outs = compaction.compact_rows(keep, tensors)
    for o, t in zip(outs, tensors):
        assert o.dtype == t.dtype and torch.equal(o, t[keep])


def test_errors():
    keep = torch.ones(8, dtype=torch.bool, device=DEV)
    with pytest.raises(RuntimeError, match="multiple of 4"):
        compaction.compact_rows(keep, [torch.zeros(8, dtype=torch.uint8, device=DEV)])
    with pytest.raises(RuntimeError, match="P rows"):
        compaction.compact_rows(keep, [torch.zeros(7, 3, device=DEV)])
    with pytest.raises(RuntimeError, match="CUDA"):
        compaction.compact_rows(torch.ones(8, dtype=torch.bool), [])


class Gaussians:
    pass


def _model(P, cls):
    g = torch.Generator().manual_seed(3)
    m = Gaussians()
    shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}
    groups = []
    for name, shp in shapes.items():
        p = torch.nn.Parameter(torch.randn(*shp, generator=g).to(DEV).requires_grad_(True))
        setattr(m, compaction._PARAM_ATTR[name], p)
        groups.append({"params": [p], "lr": 1e-3, "name": name})
    m.optimizer = cls(groups, lr=0.0, eps=1e-15)
    for p in [gr["params"][0] for gr in groups]:
        p.grad = torch.randn(p.shape, generator=g).to(DEV)
    m.optimizer.step()
    m.optimizer.zero_grad(set_to_none=True)
    m.xyz_gradient_accum = torch.rand(P, 1, generator=g).to(DEV)
    m.denom = torch.rand(P, 1, generator=g).to(DEV)
    m.max_radii2D = torch.rand(P, generator=g).to(DEV)
    m.unique_kfIDs = torch.randint(0, 9, (P,), generator=g).int()
    m.n_obs = torch.randint(0, 9, (P,), generator=g).int()
    return m


def _reference_prune(m, mask):
    """gaussian_model.py:526-564 with torch ops"""
    valid = ~mask
    for group in m.optimizer.param_groups:
        old = group["params"][0]
        st = m.optimizer.state.get(old, None)
        if st is not None:
            st["exp_avg"], st["exp_avg_sq"] = st["exp_avg"][valid], st["exp_avg_sq"][valid]
            del m.optimizer.state[old]
            group["params"][0] = torch.nn.Parameter(old[valid].requires_grad_(True))
            m.optimizer.state[group["params"][0]] = st
        else:
            group["params"][0] = torch.nn.Parameter(old[valid].requires_grad_(True))
        setattr(m, compaction._PARAM_ATTR[group["name"]], group["params"][0])
    m.xyz_gradient_accum, m.denom, m.max_radii2D = m.xyz_gradient_accum[valid], m.denom[valid], m.max_radii2D[valid]
    m.unique_kfIDs, m.n_obs = m.unique_kfIDs[valid.cpu()], m.n_obs[valid.cpu()]


def test_prune_points_matches_reference_procedure():
    P = 30_011
    a, b = _model(P, FusedAdam), _model(P, FusedAdam)
    mask = torch.rand(P, generator=torch.Generator().manual_seed(9)).to(DEV) < 0.3
    compaction.prune_points(a, mask)
    _reference_prune(b, mask)
    for ga, gb in zip(a.optimizer.param_groups, b.optimizer.param_groups):
        pa, pb = ga["params"][0], gb["params"][0]
        assert isinstance(pa, torch.nn.Parameter) and pa.requires_grad and torch.equal(pa, pb)
        assert getattr(a, compaction._PARAM_ATTR[ga["name"]]) is pa
        sa, sb = a.optimizer.state[pa], b.optimizer.state[pb]
        assert torch.equal(sa["exp_avg"], sb["exp_avg"]) and torch.equal(sa["exp_avg_sq"], sb["exp_avg_sq"])
        assert sa["step"] == sb["step"] if not torch.is_tensor(sa["step"]) else torch.equal(sa["step"], sb["step"])
        assert sa["exp_avg"].shape == pa.shape
    for n in ("xyz_gradient_accum", "denom", "max_radii2D", "unique_kfIDs", "n_obs"):
        assert torch.equal(getattr(a, n), getattr(b, n)), n
    # and the optimiser keeps working on the pruned model
    for gr in a.optimizer.param_groups:
        gr["params"][0].grad = torch.ones_like(gr["params"][0])
    a.optimizer.step()
