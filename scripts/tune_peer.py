#!/usr/bin/env python
"""Launch-shape sweep of the peer exchange kernel (torchrun, N ranks): allreduce time of 16.8 / 56 / 128 MB buckets per
(blocks, unroll), 10 back-to-back calls per measurement (one event pair), median of 7, max over ranks."""
import json, os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wildgs_slam_blackwell_b200 import peer
from wildgs_slam_blackwell_b200.peer import PeerBucket

rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
res = {}
SWEEP_BLOCKS = [int(x) for x in os.environ.get("SWEEP_BLOCKS", "72,144,296,592").split(",")]
SWEEP_UNROLL = [int(x) for x in os.environ.get("SWEEP_UNROLL", "2,4,8").split(",")]
for n in (4_200_000, 14_000_000, 32_000_000):
    b = PeerBucket(n, dev, use_multicast=True)
    b.flat.normal_()
    for blocks in SWEEP_BLOCKS:
        for unroll in SWEEP_UNROLL:
            dist.barrier(); torch.cuda.synchronize()
            peer.configure(blocks, unroll)
            ts = []
            for rep in range(8):
                dist.barrier(); torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(10):
                    b.all_reduce()
                e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1) / 10)
            t = torch.tensor(sorted(ts[1:])[3], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res[f"{n * 4 / 1e6:.0f}MB_{'nvls' if b.multicast else 'p2p'}_b{blocks}_u{unroll}"] = round(t.item() * 1e3, 1)
    y = torch.randn(n, device=dev)
    ts = []
    for rep in range(6):
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            dist.all_reduce(y)
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / 10)
    t = torch.tensor(sorted(ts[1:])[2], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res[f"{n * 4 / 1e6:.0f}MB_nccl"] = round(t.item() * 1e3, 1)
    del b, y
if rank == 0:
    print(json.dumps({"world": world, "unit": "us per allreduce", **res}))
dist.barrier(); dist.destroy_process_group()
