#!/usr/bin/env python
"""Stage times (library events, wgs_stage_timing) of ONE work unit of the mapping window: a whole view of config 3 and
tile-row strips of it, overwrite / accumulate / atomic gradient sinks.   python scripts/strip_stages.py [--steps 20]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import torch  # noqa: E402

from wildgs_slam_blackwell_b200 import _lib, rasterizer  # noqa: E402
from window_workload import WindowWorkload, SINK_NAMES  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--workload", default="config3")
    ap.add_argument("--only", default="", help="r0,r1,mode: one configuration only (profiler runs), mode = 0 | 1 | atomic")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    wl = WindowWorkload(a.workload, dev, peer="off", use_graph=False, streams=1)
    b = wl.trainer.bucket
    sink = {g: b.views[k] for k, g in SINK_NAMES.items()}
    flush = torch.zeros((512 << 20) // 4, device=dev)
    rows = wl.tile_rows
    out = []
    cases = [(r, acc) for r in [(0, rows), (0, rows // 4), (rows // 4, rows // 2), (0, rows // 8)] for acc in (False, True, "atomic")]
    if a.only:
        r0, r1, mode = a.only.split(",")
        cases = [((int(r0), int(r1)), {"0": False, "1": True}.get(mode, "atomic"))]
    for (r0, r1), acc in cases:
        if True:
            tot = {}
            _lib.stage_timing(0, True)
            for _ in range(a.steps + 2):
                flush.add_(1)
                with (rasterizer.tile_rows(r0, r1) if (r0, r1) != (0, rows) else torch.no_grad.__new__(torch.no_grad) if False else _Null()), \
                        rasterizer.grad_sink(sink, accumulate=acc):
                    loss, _ = wl.render_loss(0)
                    loss.backward()
                torch.cuda.synchronize()
                for k, v in _lib.stage_times(0).items():
                    tot[k] = tot.get(k, 0.0) + v
            _lib.stage_timing(0, False)
            rec = {"rows": [r0, r1], "accumulate": acc, "ms": {k: round(v / (a.steps + 2), 4) for k, v in tot.items()}}
            rec["sum_ms"] = round(sum(rec["ms"].values()), 4)
            out.append(rec)
            print(json.dumps(rec), flush=True)


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *e):
        return False


if __name__ == "__main__":
    main()
