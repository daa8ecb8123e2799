#!/usr/bin/env python
"""Mapping-window throughput (BASELINE.json config 3: 1 M Gaussians, 10 keyframes at 640x480; config5 = the final shape
of the densification workload, 2 M Gaussians / 16 keyframes).

    python scripts/bench_window.py [--steps K] [--warmup W] [--workload config3] [--plan balanced] [--streams 2]
    python -m torch.distributed.run --nproc-per-node N ... scripts/bench_window.py --peer auto
    python scripts/bench_window.py --emulate 8          # per-rank compute time of an 8-GPU step, on ONE GPU

One step = every keyframe of the window rendered + back-propagated (views / strips sharded over ranks by WindowTrainer,
the backward kernels accumulating into one flat bucket) + ONE exchange of the Gaussian-gradient bucket.  STRONG scaling:
the window is fixed.  Prints one JSON line (rank 0).  (scripts/window_workload.py holds the workload; bench.py runs the
same object for the `window` entry of its line.)"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from window_workload import WindowWorkload  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="config3")
    ap.add_argument("--no-sink", action="store_true", help="let autograd allocate/add per-view gradients")
    ap.add_argument("--peer", default="auto", choices=["off", "auto", "nvls", "p2p"],
                    help="bucket in NVLink peer memory, summed by wgs_peer_allreduce instead of NCCL")
    ap.add_argument("--plan", default="auto", choices=["views", "strips", "balanced", "auto"])
    ap.add_argument("--streams", type=int, default=2, help="CUDA streams a rank spreads its work units over")
    ap.add_argument("--eager", action="store_true", help="do not capture the rank's step in a CUDA graph")
    ap.add_argument("--emulate", type=int, default=0,
                    help="N: run, one after the other, what each rank of an N-rank job would compute (no exchange)")
    ap.add_argument("--emulate-rank", type=int, default=-1, help="with --emulate: only this rank (for a profiler run)")
    ap.add_argument("--parity", action="store_true", help="also run the multi-GPU gradient parity check")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    if a.emulate:
        from wildgs_slam_blackwell_b200 import scenes
        scene = scenes.make_scene(a.workload)
        per_rank = []
        for r in (range(a.emulate) if a.emulate_rank < 0 else [a.emulate_rank]):
            wl = WindowWorkload(a.workload, dev, peer="off", plan=a.plan, streams=a.streams, use_graph=not a.eager,
                                sink=not a.no_sink, emulate=(r, a.emulate), scene=scene)
            ms, out = wl.time(a.steps, a.warmup)
            per_rank.append(round(ms, 4))
            del wl
        print(json.dumps({"metric": "per-rank compute time of a mapping-window step (no exchange), emulated on one GPU",
                          "emulated_world": a.emulate, "ms_per_rank": per_rank, "max_ms": max(per_rank),
                          "plan": a.plan, "streams": a.streams, "cuda_graph": not a.eager, "workload": a.workload,
                          "assignment": out["assignment"]}), flush=True)
        return
    wl = WindowWorkload(a.workload, dev, peer=a.peer, plan=a.plan, streams=a.streams, use_graph=not a.eager,
                        sink=not a.no_sink)
    ms, out = wl.time(a.steps, a.warmup)
    par = wl.parity() if a.parity else None
    if rank == 0:
        line = {"metric": "mapping-window iterations/sec", "value": round(1e3 / ms, 3), "unit": "window steps/s",
                "frames_per_s": round(wl.n_views * 1e3 / ms, 2), "ms_per_step": round(ms, 4), "n_gpus": world,
                "steps": a.steps, "warmup": a.warmup, "scaling": "strong", "higher_is_better": True, "dtype": "f32",
                "data": "synthetic", "config": wl.describe(out)}
        if par:
            line["multi_gpu_parity"] = par
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
