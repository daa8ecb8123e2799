"""Multi-GPU self-checks that run inside a live process group in well under a second per rank: the checks of
tests/test_gpu_peer.py in a form `bench.py --gpus N` can run where a 1-GPU test box cannot (SURVEY.md §8(e):
"allreduced bucket == sum of the per-rank contributions, independent of the path").

    multi_gpu_checks(dev) -> {"ok": bool, "n_ranks": N, "peer_allreduce": {...}, "sharded_adam": {...}}

Every rank must call it (collectives inside); every rank returns the same verdict.  No CPU path."""
import torch
import torch.distributed as dist


def _all_ok(flag, dev):
    t = torch.tensor([1.0 if flag else 0.0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item() > 0.5)


def _check_peer_allreduce(dev, rank, world):
    """The library's exchange kernel (NVSwitch multimem path and P2P two-shot path) against ncclAllReduce on a ragged
    buffer: relative error <= 1e-6 of the largest sum, and the result bit-identical on every rank."""
    from .peer import PeerBucket
    res = {}
    for name, mc in (("nvls", True), ("p2p", False)):
        worst, same = 0.0, True
        b = PeerBucket(1_000_003, dev, use_multicast=mc)
        for it in range(2):
            x = torch.randn(1_000_003, device=dev, generator=torch.Generator(device=dev).manual_seed(10 * it + rank))
            b.flat.copy_(x)
            ref = x.clone()
            dist.all_reduce(ref)
            b.all_reduce()
            torch.cuda.synchronize(dev)
            first = b.flat.clone()
            dist.broadcast(first, 0)
            same = same and torch.equal(first, b.flat)
            worst = max(worst, (b.flat - ref).abs().max().item() / max(ref.abs().max().item(), 1e-30))
        res[name] = {"max_rel_err": worst, "identical_on_all_ranks": _all_ok(same, dev), "multicast": bool(b.multicast),
                     "ok": _all_ok(worst < 1e-6 and same, dev)}
        del b
    return res


def _check_sharded_adam(dev, rank, world):
    """Three optimiser steps, with an append and a prune between them: reduce-scatter -> Adam on the owned 1/world ->
    allgather (GaussianArena.sharded_adam_step) leaves the same parameters, bit for bit, as allreduce -> Adam over all
    rows on every rank (adam_step)."""
    from .arena import GaussianArena
    from .window import GradBucket
    lrs = {"xyz": 1e-3, "f_dc": 2e-3, "opacity": 5e-2, "scaling": 1e-3, "rotation": 1e-3}

    def leaves(P):
        g = torch.Generator(device=dev).manual_seed(3)
        r = lambda *s: torch.randn(*s, device=dev, generator=g)
        return {"xyz": r(P, 3), "f_dc": r(P, 1, 3), "opacity": r(P, 1), "scaling": r(P, 3), "rotation": r(P, 4)}
    arenas = [GaussianArena(leaves(20_001), group=dist.group.WORLD) for _ in range(2)]          # [replicated, sharded]
    buckets = [GradBucket(a.params, dist.group.WORLD, "auto", side_floats=64, stat_floats=2, capacity=a.capacity)
               for a in arenas]
    step, oks = 0, []

    def fill(a, b):
        for i, k in enumerate(a.names):           # each rank contributes its own gradient
            b.views[k].copy_(torch.cos(a.params[k].detach() * (1.5 + i) + 0.3 * step + rank))
        b.side.fill_(float(rank + 1))

    def both():
        nonlocal step
        step += 1
        fill(arenas[0], buckets[0])
        buckets[0].all_reduce()
        arenas[0].adam_step(buckets[0], lrs)
        fill(arenas[1], buckets[1])
        buckets[1].reduce_scatter()
        arenas[1].sharded_adam_step(buckets[1], lrs)
        torch.cuda.synchronize(dev)
        side = bool((buckets[1].side == sum(range(1, world + 1))).all())
        return side and all(torch.equal(arenas[0].params[k].detach(), arenas[1].params[k].detach()) for k in arenas[0].names)

    oks.append(both())
    new = {k: p.detach()[:5_003] * 0.5 for k, p in arenas[0].params.items()}
    for a, b in zip(arenas, buckets):
        a.append(new)
        b.retarget(a.params)
    oks.append(both())
    remove = torch.arange(arenas[0].P, device=dev) % 5 == 0
    for a, b in zip(arenas, buckets):
        a.prune(remove)
        b.retarget(a.params)
    oks.append(both())
    for a in arenas:
        a._sync_moments()
    torch.cuda.synchronize(dev)
    moments = all(torch.equal(arenas[0].moments(k)[j], arenas[1].moments(k)[j]) for k in arenas[0].names for j in (0, 1))
    return {"steps_equal": [_all_ok(o, dev) for o in oks], "moments_equal_after_gather": _all_ok(moments, dev),
            "rows": int(arenas[0].P), "ok": _all_ok(all(oks) and moments, dev)}


@torch.no_grad()
def multi_gpu_checks(dev):
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() < 2:
        raise RuntimeError("multi_gpu_checks needs an initialised NCCL process group of at least 2 ranks")
    rank, world = dist.get_rank(), dist.get_world_size()
    out = {"n_ranks": world}
    try:
        out["peer_allreduce"] = _check_peer_allreduce(dev, rank, world)
        out["sharded_adam"] = _check_sharded_adam(dev, rank, world)
        out["ok"] = all(v["ok"] for v in out["peer_allreduce"].values()) and out["sharded_adam"]["ok"]
    finally:
        torch.cuda.synchronize(dev)
    torch.cuda.empty_cache()
    return out
