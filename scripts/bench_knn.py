#!/usr/bin/env python
"""distCUDA2 timing: ours vs the unmodified reference simple-knn (oracle/_ref/libref_simple_knn.so) on a
back-projected-depth-map-like cloud.  Prints one JSON line (ms per call, median of 10, CUDA events for ours,
wall clock around the reference's blocking call)."""
import ctypes as C, json, os, statistics, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_knn import clouds, REF_SO
from simple_knn._C import distCUDA2
res = {}
ref = C.CDLL(REF_SO) if os.path.exists(REF_SO) else None
if ref is not None:
    ref.ref_dist_cuda2.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
for P in (30_000, 100_000, 300_000, 1_000_000):
    pts = torch.from_numpy(clouds("surface", P, 1)).cuda()
    out = torch.zeros(P, device="cuda")
    for _ in range(3):
        distCUDA2(pts)
    ts = []
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); distCUDA2(pts); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    r = {"ours_ms": round(statistics.median(ts), 4)}
    if ref is not None:
        tr = []
        for _ in range(4):
            torch.cuda.synchronize(); t0 = time.perf_counter(); ref.ref_dist_cuda2(P, pts.data_ptr(), out.data_ptr()); tr.append((time.perf_counter() - t0) * 1e3)
        r["reference_ms"] = round(statistics.median(tr[1:]), 3)
    res[f"P{P}"] = r
print(json.dumps({"what": "distCUDA2, surface-like cloud, ms per call", **res}))
