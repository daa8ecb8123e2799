"""prune_points at 1 M Gaussians: fused compaction vs the reference procedure (boolean indexing per tensor,
gaussian_model.py:526-564).  Prints one JSON line; bytes = rows read + rows written over the 21 device tensors."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_compact import _model, _reference_prune  # noqa: E402
from wildgs_slam_blackwell_b200 import compaction  # noqa: E402
from wildgs_slam_blackwell_b200.adam import FusedAdam  # noqa: E402

P = int(os.environ.get("P", 1_000_000))
mask = torch.rand(P, generator=torch.Generator().manual_seed(1)).cuda() < 0.1


def run(fn, reps=10):
    ts = []
    for _ in range(reps + 2):
        m = _model(P, FusedAdam)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn(m, mask)
        torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    return sorted(ts[2:])[len(ts[2:]) // 2]




def device_only():
    """the two native calls alone (CUDA events), outputs preallocated"""
    m = _model(P, FusedAdam)
    ts = [m._xyz, m._features_dc, m._features_rest, m._opacity, m._scaling, m._rotation]
    ts = [t.detach() for t in ts] * 3 + [m.xyz_gradient_accum, m.denom, m.max_radii2D]
    e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    best = 1e9
    for _ in range(5):
        torch.cuda.synchronize()
        e[0].record()
        compaction.compact_rows(~mask, ts)
        e[1].record()
        torch.cuda.synchronize()
        best = min(best, e[0].elapsed_time(e[1]))
    return best


dev = device_only()
ours, ref = run(compaction.prune_points), run(_reference_prune)
row_bytes = 3 * (3 + 3 + 45 + 1 + 3 + 4) * 4 + 3 * 4
kept = int((~mask).sum())
print(json.dumps({"what": "prune_points", "P": P, "kept": kept, "fused_ms": round(ours, 3), "compact_rows_call_ms": round(dev, 3), "reference_procedure_ms": round(ref, 3),
                  "algorithmic_GB": round((P + kept) * row_bytes / 1e9, 3)}))
