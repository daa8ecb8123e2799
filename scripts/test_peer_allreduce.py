#!/usr/bin/env python
"""Multi-GPU check + timing of the peer-memory allreduce (run under torchrun on >= 2 GPUs):
result vs NCCL allreduce of the same data, bit-identical across ranks, repeated calls, both paths."""
import json, os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wildgs_slam_blackwell_b200.peer import PeerBucket

rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
res = {}
for n in (4_200_000, 14_000_000, 1000):
    for mc in (True, False):
        b = PeerBucket(n, dev, use_multicast=mc)
        tag = f"n{n}_{'nvls' if b.multicast else 'p2p'}"
        ok = True
        for it in range(3):
            g = torch.Generator(device=dev).manual_seed(100 * it + rank)
            x = torch.randn(n, device=dev, generator=g)
            b.flat.copy_(x)
            ref = x.clone(); dist.all_reduce(ref)
            b.all_reduce()
            torch.cuda.synchronize()
            err = (b.flat - ref).abs().max().item() / max(ref.abs().max().item(), 1e-30)
            same = b.flat.clone(); dist.broadcast(same, 0)
            ok = ok and err < 1e-6 and torch.equal(same, b.flat)
        # timing: max over ranks of CUDA-event time, 20 calls after a barrier each
        def timed(fn):
            ts = []
            for _ in range(25):
                dist.barrier(); torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            t = torch.tensor(sorted(ts)[len(ts) // 2], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return round(t.item() * 1e3, 1)
        y = torch.randn(n, device=dev)
        res[tag] = {"ok": ok, "peer_us": timed(b.all_reduce), "nccl_us": timed(lambda: dist.all_reduce(y))}
        del b
if rank == 0:
    print(json.dumps({"world": world, **res}))
dist.barrier(); dist.destroy_process_group()
