#!/usr/bin/env python
"""BASELINE.json config 5 (densification stress): 100 k -> 2 M Gaussians over 40 growth steps, 16-keyframe window,
uncertainty-weighted loss, prune every 10 steps, Adam on the capacity-headroom arena.

    python scripts/bench_config5.py [--sharded] [--reps 3] [--small]
    python -m torch.distributed.run --nproc-per-node 8 ... scripts/bench_config5.py --peer auto [--sharded]

Prints one JSON line (rank 0): steps/s per growth stage, the capacity classes entered (= allocations / peer rendezvous over
the whole run), the densify cost per step."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from config5_workload import Config5Workload  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--peer", default="auto", choices=["off", "auto", "nvls", "p2p"])
    ap.add_argument("--sharded", action="store_true", help="reduce-scatter + Adam on 1/world + allgather of the parameters")
    ap.add_argument("--streams", type=int, default=2)
    ap.add_argument("--reserve", action="store_true", help="allocate the final capacity class once, before the run")
    ap.add_argument("--eager", action="store_true")
    ap.add_argument("--small", action="store_true", help="10 k -> 200 k over 12 steps (smoke)")
    ap.add_argument("--profile-last", action="store_true", help="cudaProfilerStart before the timed steps of the last stage")
    ap.add_argument("--stop-at", type=int, default=0, help="cut the schedule after this many growth steps")
    ap.add_argument("--fixed-rows", type=int, default=0, help="one stage at this row count instead of the schedule")
    ap.add_argument("--emulate", type=int, default=0, help="with --fixed-rows: per-rank compute of an N-rank step (rank 0), no exchange")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    kw = dict(p0=10_000, p1=200_000, steps=12, prune_every=4) if a.small else {}
    if a.fixed_rows:
        kw = dict(sched=[a.fixed_rows])
        if a.emulate:
            kw["emulate"] = (0, a.emulate)
    if a.stop_at:
        from config5_workload import schedule
        kw["sched"] = schedule()[:a.stop_at]
    wl = Config5Workload(dev, peer=a.peer, streams=a.streams, sharded=a.sharded, use_graph=not a.eager, reserve=a.reserve, **kw)
    recs = wl.run(reps=a.reps, profile_last=a.profile_last)
    if rank == 0:
        stages = [recs[i] for i in sorted({0, len(recs) // 4, len(recs) // 2, 3 * len(recs) // 4, len(recs) - 1})]
        print(json.dumps({
            "metric": "config 5: mapping-window steps/s while the model grows (16 keyframes 640x480, uncertainty-weighted L1, Adam)",
            "n_gpus": world, "unit": "window steps/s", "scaling": "strong", "dtype": "f32", "data": "synthetic",
            "value_at_final_size": recs[-1]["steps_per_s"], "rows_final": recs[-1]["rows"],
            "stages": [{k: r[k] for k in ("step", "rows", "capacity", "ms_per_window_step", "steps_per_s", "densify_ms")} for r in stages],
            "all_steps_ms": [r["ms_per_window_step"] for r in recs], "rows": [r["rows"] for r in recs],
            "densify_ms": [r["densify_ms"] for r in recs], "loss_sums": [r["loss_sum"] for r in recs],
            # everything the schedule costs, one window step per growth stage: the densify / prune (incl. any capacity-class
            # change: allocation + peer rendezvous), the first step at the new row count (eager pass + graph capture, host
            # wall clock), and the replayed step -- max over ranks each
            "totals_ms": {"window_steps": round(sum(r["ms_per_window_step"] for r in recs), 2),
                          "densify": round(sum(r["densify_ms"] for r in recs), 2),
                          "first_steps": round(sum(r["first_step_ms"] for r in recs), 2),
                          "all_in": round(sum(r["ms_per_window_step"] + r["densify_ms"] + r["first_step_ms"] for r in recs), 2)},
            "first_step_ms": [r["first_step_ms"] for r in recs], "reserved_capacity": bool(a.reserve),
            "capacity_classes_entered": wl.arena.allocations, "bucket_allocations": wl.bucket_allocations,
            "growth_steps": len(recs),
            "config": {"optimizer": "sharded Adam (reduce-scatter, 1/world slice per rank, parameter allgather)" if wl.sharded
                       else "replicated Adam on the arena after the allreduce",
                       "exchange": "none (one rank)" if world == 1 else ("NCCL" if a.peer == "off" else "peer kernel: " + a.peer),
                       "streams_per_rank": a.streams, "cuda_graph_per_row_count": not a.eager,
                       "prune": "5 % least opaque every 10 steps", "loss": "fused uncertainty-weighted L1 (0.5 / u^2, u clipped at 0.1)",
                       "overflow_checked": True}}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
