import os, sys, time, torch
print("ALLOC_CONF", os.environ.get("PYTORCH_CUDA_ALLOC_CONF"), os.environ.get("PYTORCH_ALLOC_CONF"))
dev = torch.device("cuda", 0)
def t(n, sz, dt=torch.uint8):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): x = torch.empty(sz, dtype=dt, device=dev)
    return (time.perf_counter() - t0) / n * 1e6
for sz in (1024, 1 << 20, 10 << 20, 40 << 20):
    print(sz, "us per torch.empty: %.1f" % t(2000, sz))
s0 = torch.cuda.memory_stats()
xs = []
t0 = time.perf_counter()
for i in range(200):
    xs = [torch.empty(sz, dtype=torch.uint8, device=dev) for sz in (17 << 20, 9 << 20, 2 << 20, 30000, 3600000, 1200000)]
print("6 allocs/iter held then dropped: %.1f us/iter" % ((time.perf_counter() - t0) / 200 * 1e6))
s1 = torch.cuda.memory_stats()
for k in ("segment.all.allocated", "segment.all.freed", "num_alloc_retries", "allocation.all.allocated"):
    print(k, s0.get(k), s1.get(k))
