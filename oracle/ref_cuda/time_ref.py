#!/usr/bin/env python
"""TEST INFRASTRUCTURE: time the reference rasterizer (oracle/_ref, sm_100 build) fwd / bwd
on a named synthetic config with CUDA events.  Usage: time_ref.py [config2] [iters]"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.ref_cuda import ref_api  # noqa: E402
from wildgs_slam_blackwell_b200 import scenes  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "config2"
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    g, cams, cots = scenes.make_scene(name)
    dev = "cuda:0"
    gt, ct = ref_api.to_cuda(g, dev), ref_api.to_cuda(cams[0], dev)
    dc, dd = (torch.from_numpy(a).to(dev) for a in cots[0])
    ref = ref_api.RefRasterizer()
    tf, tb = [], []
    for i in range(iters + 3):
        t0 = time.perf_counter()
        out = ref.forward(gt, ct)          # synchronises on both sides
        t1 = time.perf_counter()
        ref.backward(dc, dd)
        t2 = time.perf_counter()
        if i >= 3:
            tf.append(t1 - t0)
            tb.append(t2 - t1)
    tf.sort(), tb.sort()
    mf, mb = tf[len(tf) // 2] * 1e3, tb[len(tb) // 2] * 1e3
    print(f"{name}: R={out['num_rendered']} visible={(out['radii'] > 0).sum().item()} "
          f"ref fwd {mf:.3f} ms  bwd {mb:.3f} ms (incl. output allocs/zero-fills)  "
          f"=> {1e3 / (mf + mb):.1f} frames/s")
    nc = ref.export_state()["n_contrib"].long()
    print("sum n_contrib =", nc.sum().item(), "max =", nc.max().item())


if __name__ == "__main__":
    main()
