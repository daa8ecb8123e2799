"""GPU, needs >= 2 devices (skipped otherwise; run with `gpurun --gpus 2`): the peer-memory allreduce kernel
(csrc/peer_allreduce.cu, both the NVSwitch-multimem and the P2P path) against NCCL's allreduce of the same
data, bit-identical across ranks, repeated calls, and replay from a CUDA graph."""
import os
import socket
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from wildgs_slam_blackwell_b200.peer import PeerBucket
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    ok = []
    try:
        for n in (1_000_003, 1000, 4):            # ragged (padded to 4), small, minimal
            for mc in (True, False):
                b = PeerBucket(n, dev, use_multicast=mc)
                for it in range(3):
                    x = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(10 * it + rank))
                    b.flat.copy_(x)
                    ref = x.clone()
                    dist.all_reduce(ref)
                    if it == 1:
                        b.rendezvous()            # one-block peer barrier (mode 3): moves nothing, keeps the generations in step
                        assert torch.equal(b.flat, x)
                    b.all_reduce()
                    torch.cuda.synchronize()
                    same = b.flat.clone()
                    dist.broadcast(same, 0)
                    err = (b.flat - ref).abs().max().item() / max(ref.abs().max().item(), 1e-30)
                    ok.append(err < 1e-6 and torch.equal(same, b.flat))
                # CUDA-graph replay (the barrier generation lives on the device)
                g = torch.cuda.CUDAGraph()
                b.flat.fill_(rank + 1.0)
                torch.cuda.synchronize()
                with torch.cuda.graph(g):
                    b.all_reduce()
                for rep in range(2):
                    b.flat.fill_(rank + 1.0 + rep)
                    g.replay()
                    torch.cuda.synchronize()
                    ok.append(bool((b.flat == sum(r + 1.0 + rep for r in range(world))).all()))
                del b, g
        torch.save(ok, os.path.join(out_dir, f"peer{rank}.pt"))
    finally:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one NVLink domain")
def test_peer_allreduce_matches_nccl(tmp_path):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    world = 2
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker, args=(r, world, port, str(tmp_path))) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    for r in range(world):
        ok = torch.load(tmp_path / f"peer{r}.pt")
        assert len(ok) == 30 and all(ok), ok


def _window_worker(rank, world, port, out_dir):
    """Two ranks, the real rasterizer: a 3-view window through WindowTrainer (whole views, contiguous strips, balanced
    plan; one and two streams; eager and as one CUDA graph), every variant compared on every rank with the sum of the
    whole-view gradients that rank computes alone (independent of the assignment)."""
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    from window_workload import WindowWorkload
    from wildgs_slam_blackwell_b200 import scenes
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    res = {}
    try:
        scenes.CONFIGS["_w3"] = (60_000, 320, 240, 0, 3)
        scene = scenes.make_scene(60_000, 320, 240, 0, seed=77, n_views=3)
        for peer in ("p2p", "nvls", "off"):
            for streams in (1, 2):
                wl = WindowWorkload("_w3", dev, peer=peer, streams=streams, use_graph=False, scene=scene)
                par = wl.parity(plans=("views", "strips", "balanced"))
                res[f"{peer}/s{streams}"] = par
                if peer != "off":                      # the captured step (peer exchange inside the graph)
                    wg = WindowWorkload("_w3", dev, peer=peer, streams=streams, use_graph=True, scene=scene)
                    ms, out = wg.time(3, 2)
                    torch.cuda.synchronize()
                    ref = {k: p.grad.clone() for k, p in wl.params.items()}       # eager 'balanced' result of the same plan
                    ok = all(torch.allclose(wg.params[k].grad, ref[k], rtol=2e-3, atol=1e-4 * float(ref[k].abs().max()))
                             for k in ref)
                    res[f"{peer}/s{streams}/graph"] = {"ok": bool(ok), "ms": ms}
                    del wg
                del wl
        torch.save(res, os.path.join(out_dir, f"window{rank}.pt"))
    finally:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one NVLink domain")
def test_window_trainer_two_gpus_real_rasterizer(tmp_path):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    world = 2
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_window_worker, args=(r, world, port, str(tmp_path))) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
        assert p.exitcode == 0
    for r in range(world):
        res = torch.load(tmp_path / f"window{r}.pt")
        assert len(res) == 10
        for name, v in res.items():
            assert v["ok"], (name, v)
            if "plans" in v:
                assert set(v["plans"]) == {"views", "strips", "balanced"} and v["max_rel_err"] <= 1e-3, (name, v)


def _arena_worker(rank, world, port, out_dir):
    """Sharded optimiser step (reduce-scatter -> Adam on the owned 1/world slice -> parameter allgather) against the
    replicated one (allreduce -> Adam on everything) on a capacity-segmented arena, across an append and a prune (the
    ownership slices move: the moments are gathered first); the segmented exchange modes against NCCL."""
    import torch.distributed as dist
    from wildgs_slam_blackwell_b200.arena import GaussianArena
    from wildgs_slam_blackwell_b200.peer import owned_pieces
    from wildgs_slam_blackwell_b200.window import GradBucket
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    ok = {}
    try:
        lrs = {"xyz": 1e-3, "f_dc": 2e-3, "opacity": 5e-2, "scaling": 1e-3, "rotation": 1e-3}
        for mode in ("p2p", "nvls"):
            def leaves(P):
                g = torch.Generator(device=dev).manual_seed(3)
                r = lambda *s: torch.randn(*s, device=dev, generator=g)
                return {"xyz": r(P, 3), "f_dc": r(P, 1, 3), "opacity": r(P, 1), "scaling": r(P, 3), "rotation": r(P, 4)}
            arenas = [GaussianArena(leaves(20_001), group=dist.group.WORLD, peer_mode=mode) for _ in range(2)]   # [replicated, sharded]
            buckets = [GradBucket(a.params, dist.group.WORLD, mode, side_floats=64, stat_floats=2, capacity=a.capacity) for a in arenas]
            step = 0

            def fill(a, b):
                for i, k in enumerate(a.names):       # each rank contributes its own gradient
                    b.views[k].copy_(torch.cos(a.params[k].detach() * (1.5 + i) + 0.3 * step + rank))
                b.side.fill_(float(rank + 1))

            def both_steps():
                nonlocal step
                step += 1
                fill(arenas[0], buckets[0])
                want = {k: buckets[0].views[k].clone() for k in arenas[0].names}
                for t in want.values():
                    dist.all_reduce(t)
                buckets[0].all_reduce()
                torch.cuda.synchronize()
                good = all(torch.allclose(buckets[0].views[k], want[k], rtol=1e-6, atol=1e-6) for k in want)
                good = good and bool((buckets[0].side == sum(range(1, world + 1))).all())
                arenas[0].adam_step(buckets[0], lrs)
                fill(arenas[1], buckets[1])
                buckets[1].reduce_scatter()
                torch.cuda.synchronize()
                for off, n in owned_pieces(arenas[1].segments(), rank, world):        # my slice holds the sum
                    k = next(name for name in arenas[1].names if arenas[1].offsets[name] <= off < arenas[1].offsets[name] + arenas[1].capacity * arenas[1].widths[name])
                    lo = off - arenas[1].offsets[k]
                    n_used = min(n, arenas[1].P * arenas[1].widths[k] - lo)
                    good = good and torch.allclose(buckets[1]._all[off:off + n_used], want[k].reshape(-1)[lo:lo + n_used], rtol=1e-6, atol=1e-6)
                good = good and bool((buckets[1].side == sum(range(1, world + 1))).all())
                arenas[1].sharded_adam_step(buckets[1], lrs)
                torch.cuda.synchronize()
                return good and all(torch.equal(arenas[0].params[k].detach(), arenas[1].params[k].detach()) for k in arenas[0].names)

            res = [both_steps(), both_steps()]
            new = {k: p.detach()[:5_003] * 0.5 for k, p in arenas[0].params.items()}   # append: within the capacity class
            for a, b in zip(arenas, buckets):
                a.append(new)
                b.retarget(a.params)
            res += [both_steps(), both_steps()]
            remove = torch.arange(arenas[0].P, device=dev) % 5 == 0                  # prune
            for a, b in zip(arenas, buckets):
                a.prune(remove)
                b.retarget(a.params)
            res += [both_steps()]
            for a in arenas:
                a._sync_moments()
            torch.cuda.synchronize()
            res.append(all(torch.equal(arenas[0].moments(k)[j], arenas[1].moments(k)[j]) for k in arenas[0].names for j in (0, 1)))
            big = {k: p.detach().repeat(2, *([1] * (p.dim() - 1))) for k, p in arenas[0].params.items()}   # new capacity class
            grew = [a.append(big) for a in arenas]
            buckets = [GradBucket(a.params, dist.group.WORLD, mode, side_floats=64, stat_floats=2, capacity=a.capacity) for a in arenas]
            res += [all(grew), both_steps()]
            ok[mode] = res
            del arenas, buckets
        torch.save(ok, os.path.join(out_dir, f"arena{rank}.pt"))
    finally:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one NVLink domain")
def test_sharded_adam_matches_replicated_two_gpus(tmp_path):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    world = 2
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_arena_worker, args=(r, world, port, str(tmp_path))) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
        assert p.exitcode == 0
    for r in range(world):
        ok = torch.load(tmp_path / f"arena{r}.pt")
        for mode, res in ok.items():
            assert len(res) == 8 and all(res), (mode, res)
