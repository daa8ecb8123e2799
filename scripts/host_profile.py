"""Where does the host time of one fwd+bwd go?  (diagnostic; run on the GPU box)"""
import ctypes as C, os, sys, time, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from wildgs_slam_blackwell_b200 import _lib, rasterizer

mod = bench.load_impl("ours")
dev = torch.device("cuda", 0)
wl = bench.Workload(mod, sys.argv[1] if len(sys.argv) > 1 else "config2", 0, dev)
lib = _lib.load()
T = {}
def wrap(name):
    f = getattr(lib, name)
    def g(*a):
        t0 = time.perf_counter(); r = f(*a); T.setdefault(name, []).append(time.perf_counter() - t0); return r
    return g
class L:  # proxy with timed entry points
    def __getattr__(self, n):
        return wrap(n) if n in ("wgs_forward", "wgs_backward", "wgs_workspace_sizes") else getattr(lib, n)
proxy = L()
_lib._lib = proxy
def timeit(fn, n=50, sync=True):
    for _ in range(5): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(n):
        t0 = time.perf_counter(); fn()
        if sync: torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    if not sync: torch.cuda.synchronize()
    ts.sort(); return ts[len(ts)//2]*1e3, ts[int(len(ts)*.9)]*1e3
e = torch.Tensor([])
m3, sh, op, sc, ro = wl.leaves
def fwd_only():
    with torch.no_grad():
        return rasterizer.forward_raw(m3, sh, e, op, sc, ro, e, wl.settings)
print("forward_raw  (sync each)   p50 %.3f p90 %.3f ms" % timeit(fwd_only))
print("  inside: wgs_forward p50 %.3f ms, workspace_sizes %.3f ms" % (sorted(T["wgs_forward"])[len(T["wgs_forward"])//2]*1e3, sorted(T["wgs_workspace_sizes"])[len(T["wgs_workspace_sizes"])//2]*1e3))
T.clear()
print("step_resident (sync each)  p50 %.3f p90 %.3f ms" % timeit(lambda: wl.step_resident(1)))
for k, v in T.items(): v.sort(); print("  inside:", k, "p50 %.3f p90 %.3f ms" % (v[len(v)//2]*1e3, v[int(len(v)*.9)]*1e3))
T.clear()
print("step_resident (no sync)    p50 %.3f p90 %.3f ms" % timeit(lambda: wl.step_resident(1), sync=False))
for k, v in T.items(): v.sort(); print("  inside:", k, "p50 %.3f p90 %.3f ms" % (v[len(v)//2]*1e3, v[int(len(v)*.9)]*1e3))
p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm", "--format=csv,noheader", "-lms", "100"], stdout=subprocess.DEVNULL)
time.sleep(0.5); T.clear()
print("step_resident (sync, nvidia-smi -lms 100 running)  p50 %.3f p90 %.3f ms" % timeit(lambda: wl.step_resident(1)))
for k, v in T.items(): v.sort(); print("  inside:", k, "p50 %.3f p90 %.3f ms" % (v[len(v)//2]*1e3, v[int(len(v)*.9)]*1e3))
p.terminate()
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(50): wl.step_resident(1); torch.cuda.synchronize()
pr.disable(); pstats.Stats(pr).sort_stats("tottime").print_stats(14)
