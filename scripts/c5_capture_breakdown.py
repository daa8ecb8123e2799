#!/usr/bin/env python
"""Where the ~90 ms of config 5's first step at a new row count go (eager pass, gc, warm-up, capture), and what capturing
into the previous graph's memory pool saves.  One GPU."""
import gc
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import torch  # noqa: E402

from config5_workload import Config5Workload  # noqa: E402
from wildgs_slam_blackwell_b200 import graph  # noqa: E402


def wall(fn):
    torch.cuda.synchronize()
    t = time.perf_counter()
    r = fn()
    torch.cuda.synchronize()
    return 1e3 * (time.perf_counter() - t), r


def main():
    dev = torch.device("cuda", 0)
    rows = [400_000, 420_000, 440_000, 460_000, 480_000, 500_000, 520_000]
    wl = Config5Workload(dev, sched=rows, reserve=True, use_graph=True)
    out = []
    prev = None
    for i, r in enumerate(rows):
        wl.grow_to(r, prune=False)
        rec = {"rows": r}
        with graph.async_forward(dev):
            rec["eager_ms"], _ = wall(wl._step)
        rec["gc_ms"], _ = wall(gc.collect)
        share = i >= 3 and prev is not None          # the last two captures reuse the previous graph's pool
        pool = prev.pool() if share else None
        rec["capture_ms"], it = wall(lambda: graph.CapturedIteration(wl._step, dev, warmup=1, pool=pool))
        rec["shared_pool"] = bool(share)
        rec["replay_ms"], _ = wall(it.replay)
        rec["drop_previous_ms"], _ = wall(lambda: None) if prev is None else wall(lambda: prev.__dict__.clear())
        prev = it
        out.append(rec)
    graph.check_overflow(dev)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
