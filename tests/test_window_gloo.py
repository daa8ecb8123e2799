"""CPU, world_size 2 over gloo: the mapping-window host logic (view->rank assignment, gradient bucket,
allreduce, per-view pose gradients).  The rasterizer itself has no CPU path, so a stand-in
differentiable "render" produces per-view losses with the same call structure."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wildgs_slam_blackwell_b200.window import GradBucket, WindowTrainer, assign_strips, assign_views  # noqa: E402

N_VIEWS, P = 5, 37


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _params(seed=0):
    g = torch.Generator().manual_seed(seed)
    return {"means3D": torch.randn(P, 3, generator=g).requires_grad_(True),
            "opacity": torch.rand(P, 1, generator=g).requires_grad_(True),
            "sh": torch.randn(P, 1, 3, generator=g).requires_grad_(True)}


def _make_render(params):
    poses = [torch.zeros(6, requires_grad=True) for _ in range(N_VIEWS)]

    def render_loss(v):
        w = torch.linspace(0.5, 1.5, P).view(P, 1) * (v + 1)
        poses[v].grad = None
        img = (params["means3D"] * w).sum(1, keepdim=True) * params["opacity"] + params["sh"].view(P, 3).mean(1, keepdim=True)
        loss = (img ** 2).mean() + (poses[v] * torch.arange(6.0) * (v + 1)).sum() * img.mean().detach()
        return loss, (lambda: poses[v].grad)
    return render_loss


def _make_strip_render(params, rows):
    """Stand-in for a strip render: the loss of view v is a sum over `rows` tile rows; inside
    rasterizer.tile_rows(b, e) only rows [b, e) contribute (what the real rasterizer does per pixel)."""
    from wildgs_slam_blackwell_b200 import rasterizer
    poses = [torch.zeros(6, requires_grad=True) for _ in range(N_VIEWS)]

    def render_loss(v):
        act = rasterizer.tile_rows.current()
        b, e = (0, rows) if act is None else (act.begin, act.end)
        poses[v].grad = None
        loss = 0.0
        for row in range(b, e):
            w = torch.linspace(0.5, 1.5, P).view(P, 1) * (v + 1) * (row + 1)
            img = (params["means3D"] * w).sum(1, keepdim=True) * params["opacity"] + params["sh"].view(P, 3).mean(1, keepdim=True)
            loss = loss + (img ** 2).mean() + (poses[v] * torch.arange(6.0) * (row + 1)).sum()
        return loss, (lambda: poses[v].grad)
    return render_loss


def _strip_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        params = _params()
        tr = WindowTrainer(params)
        out = tr.step(N_VIEWS, _make_strip_render(params, 4), tile_rows=4)     # 5 views on 2 ranks: view 2 is split
        torch.save((rank, {k: p.grad.clone() for k, p in params.items()}, out["losses"].clone(),
                    out["pose_grads"].clone(), out["assignment"]), os.path.join(q, f"strip{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_window_step_with_strips_world2(tmp_path):
    ctx = mp.get_context("spawn")
    port = _free_port()
    procs = [ctx.Process(target=_strip_worker, args=(r, 2, port, str(tmp_path))) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    params = _params()
    render = _make_strip_render(params, 4)
    ref_l, ref_p = [], []
    for v in range(N_VIEWS):
        loss, pg = render(v)
        loss.backward()
        ref_l.append(loss.detach())
        ref_p.append(pg().clone())
    for r in range(2):
        rank, g, losses, pose, plan = torch.load(tmp_path / f"strip{r}.pt")
        for k in params:
            torch.testing.assert_close(g[k], params[k].grad, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(losses, torch.stack(ref_l), rtol=1e-5, atol=1e-6)     # summed over a view's strips
        torch.testing.assert_close(pose, torch.stack(ref_p), rtol=1e-5, atol=1e-6)
        assert plan == [[(0, 0, 4), (1, 0, 4), (2, 0, 2)], [(2, 2, 4), (3, 0, 4), (4, 0, 4)]]


def test_assign_strips_properties():
    for costs, world, rows in [([1] * 10, 8, 30), ([3, 1, 2, 5, 1, 1, 2, 2, 4, 1], 8, 30), ([1] * 3, 2, 4), ([1, 1], 4, 30)]:
        plan = assign_strips(costs, world, rows)
        seen = set()
        for units in plan:
            for v, b, e in units:
                assert 0 <= b < e <= rows
                for row in range(b, e):
                    assert (v, row) not in seen
                    seen.add((v, row))
        assert len(seen) == len(costs) * rows                       # every (view, row) exactly once
        loads = [sum(costs[v] * (e - b) / rows for v, b, e in units) for units in plan]
        assert max(loads) <= sum(costs) / world + max(costs) / rows + 1e-9     # balanced to one tile row
    assert max(len(u) for u in assign_strips([1] * 10, 8, 30)) == 2   # config 3: at most two units per rank


def _single_process_reference():
    params = _params()
    render = _make_render(params)
    losses, pose = [], []
    for v in range(N_VIEWS):
        loss, pg = render(v)
        loss.backward()
        losses.append(loss.detach())
        pose.append(pg().clone())
    return {k: p.grad.clone() for k, p in params.items()}, torch.stack(losses), torch.stack(pose)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        params = _params()
        tr = WindowTrainer(params)
        render = _make_render(params)
        costs = {0: 5.0, 1: 1.0, 2: 3.0, 3: 3.0, 4: 2.0}
        out1 = tr.step(N_VIEWS, render, cost_fn=lambda v: costs[v])
        g1 = {k: p.grad.clone() for k, p in params.items()}
        out2 = tr.step(N_VIEWS, render, cost_fn=lambda v: costs[v])   # second step: cost-balanced assignment
        g2 = {k: p.grad.clone() for k, p in params.items()}
        torch.save((rank, g1, g2, out1["losses"].clone(), out1["pose_grads"].clone(), out1["assignment"],
                    out2["assignment"], out2["pose_grads"].clone()), os.path.join(q, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_window_step_world2_matches_single_process(tmp_path):
    ctx = mp.get_context("spawn")
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, str(tmp_path))) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    res = [torch.load(tmp_path / f"rank{r}.pt") for r in range(2)]
    ref_g, ref_l, ref_p = _single_process_reference()
    for rank, g1, g2, losses, pose, a1, a2, pose2 in res:
        for k in ref_g:
            torch.testing.assert_close(g1[k], ref_g[k], rtol=1e-5, atol=1e-6)   # sum over ALL views on every rank
            torch.testing.assert_close(g2[k], ref_g[k], rtol=1e-5, atol=1e-6)   # independent of the assignment
        torch.testing.assert_close(losses, ref_l, rtol=1e-6, atol=1e-7)
        torch.testing.assert_close(pose, ref_p, rtol=1e-6, atol=1e-7)           # per view, NOT summed
        torch.testing.assert_close(pose2, ref_p, rtol=1e-6, atol=1e-7)
        assert sorted(a1[0] + a1[1]) == list(range(N_VIEWS))
        assert a1 == [[0, 2, 4], [1, 3]]                       # uniform costs: round-robin by LPT
        loads = [sum({0: 5.0, 1: 1.0, 2: 3.0, 3: 3.0, 4: 2.0}[v] for v in part) for part in a2]
        assert max(loads) == 7.0 and sorted(a2[0] + a2[1]) == list(range(N_VIEWS))
    assert res[0][5] == res[1][5] and res[0][6] == res[1][6]   # every rank computes the same assignment


def test_assign_views_properties():
    assert assign_views([], 4) == [[], [], [], []]
    assert assign_views([1, 1, 1], 1) == [[0, 1, 2]]
    a = assign_views([10, 1, 1, 1, 1, 1, 1, 1, 1, 1], 8)      # config 3: 10 keyframes on 8 GPUs
    assert sorted(sum(a, [])) == list(range(10)) and max(len(x) for x in a) == 2 and a[0] == [0]
    a16 = assign_views([1.0] * 16, 8)                          # config 5 divides evenly
    assert all(len(x) == 2 for x in a16)


def test_grad_bucket_layout_and_accumulation():
    params = _params()
    b = GradBucket(params)
    assert b.flat.numel() == P * 3 + P + P * 3 and b.nbytes == 4 * b.flat.numel()
    b.install()
    (params["means3D"].sum() * 2).backward()
    (params["means3D"].sum() * 3).backward()                   # accumulates in place into the bucket
    assert params["means3D"].grad.data_ptr() == b.flat.data_ptr()
    torch.testing.assert_close(b.flat[:P * 3], torch.full((P * 3,), 5.0))
    assert not b.flat[P * 3:].any()
    assert b.matches(params) and not b.matches({"means3D": params["means3D"]})
    with pytest.raises(TypeError):
        GradBucket({"x": torch.zeros(3, dtype=torch.float64)})


def _stats_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from wildgs_slam_blackwell_b200 import rasterizer
        params = _params(seed=rank)                             # ranks start apart ...
        tr = WindowTrainer(params, densify_stats=True)
        tr.broadcast_params(0)                                  # ... and are replicated from rank 0
        same = _params(seed=0)
        assert all(torch.equal(params[k], same[k]) for k in params)
        base = _make_render(params)

        def render_loss(v):     # stand-in for the backward kernel: it fills whatever densification_stats holds
            acc, den, mx = rasterizer.densification_stats.current_for_backward(None).bufs if rasterizer.densification_stats.current_for_backward(None) else (None,) * 3
            loss, pg = base(v)

            class Fill(torch.autograd.Function):
                @staticmethod
                def forward(ctx, x):
                    return x.clone()

                @staticmethod
                def backward(ctx, g):
                    a, d, m = rasterizer.densification_stats.current_for_backward(None).bufs
                    a += float(v + 1); d += 1.0; m.copy_(torch.maximum(m, torch.full_like(m, float(10 * v + rank))))
                    return g
            return Fill.apply(loss), pg
        for _ in range(2):
            tr.step(N_VIEWS, render_loss)
        torch.save((tr.grad_accum.clone(), tr.denom.clone(), tr.reduce_max_radii().clone(), tr.bucket.nbytes),
                   os.path.join(q, f"stats{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_window_densification_statistics_world2(tmp_path):
    """the statistics segment rides in the gradient allreduce (SUM over ranks = over all views), per step; the radius
    maximum is a MAX reduction on demand"""
    ctx = mp.get_context("spawn")
    port = _free_port()
    procs = [ctx.Process(target=_stats_worker, args=(r, 2, port, str(tmp_path))) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    for r in range(2):
        acc, den, mx, nbytes = torch.load(tmp_path / f"stats{r}.pt")
        assert acc.shape == (P, 1) and bool((acc == sum(v + 1 for v in range(N_VIEWS))).all())
        assert bool((den == N_VIEWS).all())
        want = max(10 * max(views) + rk for rk, views in enumerate(assign_views([1.0] * N_VIEWS, 2)))
        assert bool((mx == want).all())                         # identical on both ranks: MAX across ranks
        assert nbytes == 4 * P * 7                              # nbytes counts the gradients only


def test_assign_balanced_properties():
    from wildgs_slam_blackwell_b200.window import assign_balanced
    for costs, world, rows in [([1] * 10, 8, 30), ([1] * 10, 4, 30), ([1.1, 0.9, 1.0, 1.05, 0.95, 1.0, 1.0, 1.02, 0.98, 1.0], 8, 30),
                               ([3, 1, 2, 5, 1, 1, 2, 2, 4, 1], 8, 30), ([1] * 3, 2, 4), ([1, 1], 4, 30), ([1] * 16, 8, 30)]:
        plan = assign_balanced(costs, world, rows)
        seen = set()
        for units in plan:
            for v, b, e in units:
                assert 0 <= b < e <= rows
                for row in range(b, e):
                    assert (v, row) not in seen
                    seen.add((v, row))
        assert len(seen) == len(costs) * rows                       # every (view, row) exactly once
        loads = [sum(costs[v] * (e - b) / rows for v, b, e in units) for units in plan]
        assert max(loads) <= 1.08 * sum(costs) / world + max(costs) / rows + 1e-9
    plan = assign_balanced([1] * 10, 8, 30)                        # config 3: one whole view + a quarter of another per rank
    assert all(len(u) == 2 and u[0][1:] == (0, 30) and 7 <= u[1][2] - u[1][1] <= 8 for u in plan)
    assert sorted(u[1][0] for u in plan) == [8] * 4 + [9] * 4
    assert all(u == [(i, 0, 30), (8 + i, 0, 30)] for i, u in enumerate(assign_balanced([1] * 16, 8, 30)))   # divides: whole views


def test_emulated_rank_runs_that_ranks_units_without_exchange():
    """WindowTrainer(emulate=(r, N)): rank r's share of an N-rank step in one process; the shares of all ranks sum to the
    single-process gradients, for every plan."""
    ref_params = _params()
    ref = WindowTrainer(ref_params)
    ref.step(N_VIEWS, _make_strip_render(ref_params, 4), tile_rows=4)
    for plan in ("views", "strips", "balanced"):
        total = None
        for r in range(2):
            params = _params()
            tr = WindowTrainer(params, emulate=(r, 2))
            out = tr.step(N_VIEWS, _make_strip_render(params, 4), tile_rows=4, plan=plan)
            g = {k: p.grad.clone() for k, p in params.items()}
            total = g if total is None else {k: total[k] + g[k] for k in g}
            assert len(out["assignment"]) == 2
        for k in total:
            torch.testing.assert_close(total[k], ref_params[k].grad, rtol=1e-5, atol=1e-6)


def test_owned_pieces_partition_the_segments_like_the_kernel():
    """peer.owned_pieces: the host-side mirror of the exchange kernel's split (float4 granularity, ceil(n4 / world)
    per rank over the segments laid end to end) -- the pieces of all ranks tile the segments exactly once."""
    from wildgs_slam_blackwell_b200.peer import owned_pieces
    for segments, world in (([(0, 4000)], 8), ([(0, 12), (64, 4), (128, 36)], 3), ([(256, 1000), (4096, 8)], 2),
                            ([(0, 4)], 8), ([(0, 300_000 * 3), (1_048_576 * 3, 300_000 * 4)], 8)):
        covered = []
        for r in range(world):
            pieces = owned_pieces(segments, r, world)
            assert all(o % 4 == 0 and n % 4 == 0 and n > 0 for o, n in pieces)
            covered += pieces
        cells = sorted(x for o, n in covered for x in range(o, o + n, 4))
        expect = sorted(x for o, n in segments for x in range(o, o + n, 4))
        assert cells == expect                                   # every float4 exactly once
        n4 = sum(n // 4 for _, n in segments)
        per = -(-n4 // world)
        sizes = [sum(n for _, n in owned_pieces(segments, r, world)) // 4 for r in range(world)]
        assert sizes == [max(0, min(per, n4 - per * r)) for r in range(world)]


def test_capacity_classes_are_powers_of_two_with_a_floor():
    from wildgs_slam_blackwell_b200.arena import capacity_class
    assert capacity_class(1) == capacity_class(16384) == 16384
    assert capacity_class(16385) == 32768 and capacity_class(100_000) == 131072 and capacity_class(2_000_000) == 2097152
    # BASELINE config 5's schedule enters five classes
    sched = [int(round(100_000 * 20 ** (s / 39))) for s in range(40)]
    assert len({capacity_class(p) for p in sched}) == 5


def test_rounding_pad_behind_the_used_rows_is_zero_after_a_prune():
    """The exchange moves whole 4-float groups, so up to 3 floats behind row P of every segment ride along.  After a prune
    they held a dead row's gradients and were summed over the ranks on every step: retarget() zeroes them, zero_grads() /
    zero_stats() clear the padded length, and the floats in the exchanged segments are exactly the used rows."""
    from wildgs_slam_blackwell_b200.window import GradBucket
    mk = lambda P: {"xyz": torch.zeros(P, 3), "opacity": torch.zeros(P, 1)}
    b = GradBucket(mk(11), capacity=16, stat_floats=22, side_floats=8)
    b._all.fill_(9.0)                                  # gradients, statistics and side block of a step at 11 rows
    b.retarget(mk(6))                                  # prune: 6 rows; 6*3 = 18 -> pad 18..19, 6*1 -> pad 6..7
    for (o, n), used in zip(b._segments[:4], (18, 6, 6, 6)):
        seg = b._all[o:o + n]
        assert n == (used + 3) // 4 * 4 and torch.all(seg[:used] == 9.0) and not seg[used:].any()
    b._all.fill_(9.0)
    b.zero_grads()
    b.zero_stats()
    for o, n in b._segments[:4]:
        assert not b._all[o:o + n].any()
    assert torch.all(b.side == 9.0)                    # (the side block is the trainer's to clear)


def test_multi_gpu_checks_need_a_process_group():
    """selfcheck.multi_gpu_checks is what `bench.py --gpus N` runs where the ranks exist; alone it refuses to pretend."""
    from wildgs_slam_blackwell_b200 import selfcheck
    with pytest.raises(RuntimeError, match="process group"):
        selfcheck.multi_gpu_checks(torch.device("cpu"))
