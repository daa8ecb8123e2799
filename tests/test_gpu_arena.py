"""GPU: the capacity-headroom arena (arena.GaussianArena) against the reference's densification procedure written in
torch ops -- torch.cat appends of parameters and Adam moments (gaussian_model.py:566-644 densification_postfix /
cat_tensors_to_optimizer), boolean-index prune (:526-564 prune_points / _prune_optimizer), torch.optim.Adam with one
group per tensor and eps 1e-15 (training_setup) -- and the config-5 workload built on it at a size that runs in seconds."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from wildgs_slam_blackwell_b200.arena import GaussianArena, capacity_class  # noqa: E402
from wildgs_slam_blackwell_b200.window import GradBucket  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
LRS = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "opacity": 5e-2, "scaling": 1e-3, "rotation": 1e-3}


def _leaves(P, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    r = lambda *s: torch.randn(*s, device=DEV, generator=g)
    return {"xyz": r(P, 3), "f_dc": r(P, 1, 3), "opacity": r(P, 1), "scaling": r(P, 3) - 3.0, "rotation": r(P, 4)}


def _grad(p, k, step):           # any deterministic function of the parameters: both sides see the same values
    return torch.sin(p.detach() * (2.0 + k) + 0.37 * step) * (1.0 + 0.1 * k)


class _TorchModel:
    """The reference's procedure: separate tensors, torch.optim.Adam, torch.cat / boolean indexing with the optimizer-state
    surgery of gaussian_model.py."""

    def __init__(self, leaves):
        self.p = {k: torch.nn.Parameter(v.clone()) for k, v in leaves.items()}
        self.opt = torch.optim.Adam([{"params": [self.p[k]], "lr": LRS[k], "name": k} for k in self.p], lr=0.0, eps=1e-15)
        P = leaves["xyz"].shape[0]
        self.stats = {n: torch.zeros(P, device=DEV) for n in GaussianArena.STATS}

    def _surgery(self, fn_param, fn_state):
        for group in self.opt.param_groups:
            k, old = group["name"], group["params"][0]
            st = self.opt.state.pop(old, None)
            new = torch.nn.Parameter(fn_param(k, old.detach()))
            if st is not None and "exp_avg" in st:
                st["exp_avg"], st["exp_avg_sq"] = fn_state(k, st["exp_avg"]), fn_state(k, st["exp_avg_sq"])
                self.opt.state[new] = st
            group["params"][0] = new
            self.p[k] = new

    def append(self, rows):      # densification_postfix / cat_tensors_to_optimizer
        self._surgery(lambda k, old: torch.cat([old, rows[k]], 0), lambda k, s: torch.cat([s, torch.zeros_like(rows[k])], 0))
        P = self.p["xyz"].shape[0]     # gaussian_model.py:635-637: all three re-created as zeros after every densify
        self.stats = {k: torch.zeros(P, device=DEV) for k in self.stats}

    def prune(self, remove):     # prune_points / _prune_optimizer
        keep = ~remove
        self._surgery(lambda k, old: old[keep], lambda k, s: s[keep])
        self.stats = {k: v[keep] for k, v in self.stats.items()}

    def step(self, step):
        for i, (k, p) in enumerate(self.p.items()):
            p.grad = _grad(p, i, step)
        self.opt.step()


def test_arena_matches_the_torch_cat_boolean_index_procedure():
    """10 k -> ~40 k rows over six growth steps with two prunes, one Adam step after each: parameters, Adam moments and
    statistics equal the reference procedure's; pointers only change when a capacity class is entered (3 over the run)."""
    P0 = 10_000
    leaves = _leaves(P0, 1)
    arena = GaussianArena(leaves)
    ref = _TorchModel(leaves)
    bucket = GradBucket(arena.params, capacity=arena.capacity)
    assert arena.capacity == capacity_class(P0) == 16384 and arena.allocations == 1
    g = torch.Generator(device=DEV).manual_seed(5)
    sched = [10_000, 13_000, 17_000, 23_000, 31_000, 40_000]
    ptr_changes, step_no = 0, 0
    for s, rows in enumerate(sched):
        if s in (2, 4):                                   # prune ~7 % (an arbitrary deterministic mask)
            remove = (torch.arange(arena.P, device=DEV) * 2654435761 % 97) < 7
            n_keep = arena.prune(remove)
            ref.prune(remove)
            assert n_keep == ref.p["xyz"].shape[0] == arena.P
        n_new = rows - arena.P
        if n_new > 0:
            idx = torch.randint(0, arena.P, (n_new,), device=DEV, generator=g)
            new = {k: p.detach()[idx] * 0.5 + 0.25 for k, p in arena.params.items()}
            before = arena._param.data_ptr()
            grew = arena.append(new)
            ref.append(new)
            ptr_changes += arena._param.data_ptr() != before
            assert grew == (arena._param.data_ptr() != before)
        assert arena.P == rows == ref.p["xyz"].shape[0]
        if not bucket.matches(arena.params):
            bucket = GradBucket(arena.params, capacity=arena.capacity)
        else:
            bucket.retarget(arena.params)
        if n_new > 0:                                      # densification_postfix restarted the statistics of every row
            assert all(float(getattr(arena, n).abs().max()) == 0.0 for n in GaussianArena.STATS)
        rowval = torch.arange(arena.P, device=DEV, dtype=torch.float32) % 251 + s      # statistics travel with their rows
        arena.xyz_gradient_accum += rowval[:, None] * 0.5
        arena.denom += rowval[:, None]
        torch.maximum(arena.max_radii2D, rowval * 2.0, out=arena.max_radii2D)
        ref.stats["xyz_gradient_accum"] += rowval * 0.5
        ref.stats["denom"] += rowval
        ref.stats["max_radii2D"] = torch.maximum(ref.stats["max_radii2D"], rowval * 2.0)
        for _ in range(2):
            step_no += 1
            for i, k in enumerate(arena.names):
                bucket.views[k].copy_(_grad(arena.params[k], i, step_no))
            arena.adam_step(bucket, LRS)
            ref.step(step_no)
        torch.cuda.synchronize()
        for k in arena.names:
            torch.testing.assert_close(arena.params[k].detach(), ref.p[k].detach(), rtol=2e-6, atol=2e-7, msg=f"{k} at stage {s}")
            m, v = arena.moments(k)
            st = ref.opt.state[ref.p[k]]
            # torch's Adam forms m with lerp (m + (g - m)(1 - b1)), the kernel with b1 m + (1 - b1) g: a few ulp OF THE
            # TERMS apart (|g| ~ 1 here), which is a large relative error where they cancel -> absolute tolerance
            torch.testing.assert_close(m, st["exp_avg"], rtol=1e-5, atol=2e-6)
            torch.testing.assert_close(v, st["exp_avg_sq"], rtol=1e-5, atol=1e-8)
        for n in GaussianArena.STATS:
            assert torch.equal(getattr(arena, n).reshape(-1), ref.stats[n]), f"{n} at stage {s}"
    assert arena.allocations == 3 and ptr_changes == 2 and arena.capacity == 65536      # 16384 -> 32768 -> 65536


def test_config5_workload_small():
    """The densification-stress loop at a size that runs in seconds: 8 k -> 40 k Gaussians over 6 growth steps, 4 keyframes
    160x112, prune every 3 steps, the rank's step recaptured as a CUDA graph per row count, overflow checked."""
    from config5_workload import Config5Workload
    wl = Config5Workload(torch.device(DEV), p0=8_000, p1=40_000, steps=6, n_views=4, W=160, H=112, prune_every=3, streams=2)
    recs = wl.run(reps=2)
    assert [r["rows"] for r in recs] == wl.sched and recs[-1]["rows"] == 40_000
    assert all(np.isfinite(r["loss_sum"]) and r["ms_per_window_step"] > 0 for r in recs)
    first = float(wl.step().sum())                                     # at a fixed row count Adam on the arena reduces the loss
    for _ in range(30):
        last = float(wl.step().sum())
    wl.check()
    assert last < first
    assert wl.arena.allocations == 3 and wl.bucket_allocations == 3    # 16384, 32768, 65536: no allocation per densification
    p = wl.arena.params
    assert all(bool(torch.isfinite(t).all()) for t in p.values())
    assert float(wl.arena.denom.max()) > 0
