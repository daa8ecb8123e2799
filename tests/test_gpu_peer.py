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
