import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch, bench
from wildgs_slam_blackwell_b200 import rasterizer
mod = bench.load_impl("ours"); dev = torch.device("cuda", 0)
wl = bench.Workload(mod, "config2", 0, dev)
rec = {}
class P:
    def __getattr__(self, n): return getattr(torch, n)
    def empty(self, *a, **k):
        t0 = time.perf_counter(); r = torch.empty(*a, **k); dt = time.perf_counter() - t0
        rec.setdefault(str(a[0]), []).append(dt); return r
rasterizer.torch = P()
for _ in range(10): wl.step_resident(1)
torch.cuda.synchronize(); rec.clear()
s0 = torch.cuda.memory_stats()
for _ in range(50): wl.step_resident(1); torch.cuda.synchronize()
s1 = torch.cuda.memory_stats()
for k, v in rec.items(): v.sort(); print("%-22s n=%d p50 %.1f us max %.1f us" % (k, len(v), v[len(v)//2]*1e6, v[-1]*1e6))
for k in ("segment.all.allocated", "segment.all.freed", "num_alloc_retries"): print(k, s0.get(k), s1.get(k))
print(torch.cuda.memory_summary(abbreviated=True)[:1500])
