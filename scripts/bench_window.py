#!/usr/bin/env python
"""Mapping-window throughput (BASELINE.json config 3: 1 M Gaussians, 10 keyframes at 640x480).

    python scripts/bench_window.py [--steps K] [--warmup W] [--workload config3]
    python -m torch.distributed.run --nproc-per-node N ... scripts/bench_window.py

One step = every keyframe of the window rendered + back-propagated (views sharded over ranks by
WindowTrainer, the backward kernels accumulating into one flat bucket) + ONE NCCL allreduce of the
Gaussian-gradient bucket.  STRONG scaling: the window is fixed, so 10 keyframes on 8 GPUs cap at
10/2 = 5x (SURVEY.md §7 "hard parts").  Prints one JSON line (rank 0)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import diff_gaussian_rasterization as dgr  # noqa: E402
from wildgs_slam_blackwell_b200 import scenes  # noqa: E402
from wildgs_slam_blackwell_b200.window import WindowTrainer  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="config3")
    ap.add_argument("--no-sink", action="store_true", help="let autograd allocate/add per-view gradients")
    ap.add_argument("--peer", default="off", choices=["off", "auto", "nvls", "p2p"],
                    help="bucket in NVLink peer memory, summed by wgs_peer_allreduce instead of NCCL")
    ap.add_argument("--strips", action="store_true",
                    help="sub-frame units (tile-row strips) when the window does not divide by the world size")
    ap.add_argument("--graph", action="store_true",
                    help="capture the rank's whole window step (all its units + the peer allreduce) in one CUDA graph")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    P, W, H, deg, n_views = scenes.CONFIGS[a.workload]
    g, cams, cots = scenes.make_scene(a.workload)
    t = lambda x: torch.from_numpy(x).to(dev)
    params = {k: t(g[k]).requires_grad_(True) for k in ("means3D", "sh", "opacities", "scales", "rotations")}
    sink_names = None if a.no_sink else {"means3D": "means3D", "sh": "sh", "opacities": "opacity",
                                         "scales": "scales", "rotations": "rotations"}
    trainer = WindowTrainer(params, sink_names=sink_names, peer=a.peer != "off", peer_mode=a.peer)
    means2D = torch.zeros(P, 3, device=dev, requires_grad=True)
    views = []
    for cam, cot in zip(cams, cots):
        s = dgr.GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=torch.zeros(3, device=dev),
            scale_modifier=1.0, viewmatrix=t(cam["viewmatrix"]), projmatrix=t(cam["projmatrix"]),
            projmatrix_raw=t(cam["projmatrix_raw"]), sh_degree=deg, campos=t(cam["campos"]), prefiltered=False,
            debug=False)
        views.append(dict(rast=dgr.GaussianRasterizer(s), dc=t(cot[0]), dd=t(cot[1]),
                          theta=torch.zeros(3, device=dev, requires_grad=True),
                          rho=torch.zeros(3, device=dev, requires_grad=True)))
    last_R = {}

    def render_loss(v):
        vw = views[v]
        vw["theta"].grad = vw["rho"].grad = None
        color, radii, depth, opacity, n_touched = vw["rast"](
            params["means3D"], means2D, params["opacities"], shs=params["sh"], scales=params["scales"],
            rotations=params["rotations"], theta=vw["theta"], rho=vw["rho"])
        loss = (color * vw["dc"]).sum() + (depth * vw["dd"]).sum()
        last_R[v] = color.grad_fn.num_rendered if hasattr(color.grad_fn, "num_rendered") else 1.0
        return loss, (lambda: torch.cat([vw["rho"].grad.reshape(-1), vw["theta"].grad.reshape(-1)]))

    def step():
        return trainer.step(n_views, render_loss, cost_fn=None if (a.strips or a.graph) else (lambda v: last_R.get(v, 1.0)),
                            tile_rows=(H + 15) // 16 if a.strips else None)

    flush = torch.zeros((512 << 20) // 4, device=dev)
    if a.graph:
        if world > 1 and a.peer == "off":
            raise SystemExit("--graph needs --peer (an NCCL allreduce is not captured here)")
        from wildgs_slam_blackwell_b200 import graph
        eager_out = step()
        it = graph.CapturedIteration(lambda: step()["losses"], dev, warmup=2)
        eager_step, step = step, (lambda: (it.replay(), eager_out)[1])
    for _ in range(a.warmup):
        step()
        flush.add_(1)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    evs = []
    for _ in range(a.steps):
        flush.add_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = step()
        e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    total = torch.tensor([sum(x.elapsed_time(y) for x, y in evs)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total, op=dist.ReduceOp.MAX)
    ms = float(total.item()) / a.steps
    if rank == 0:
        print(json.dumps({
            "metric": "mapping-window iterations/sec", "value": round(1e3 / ms, 3), "unit": "window steps/s",
            "frames_per_s": round(n_views * 1e3 / ms, 2), "ms_per_step": round(ms, 4), "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "scaling": "strong", "higher_is_better": True, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"{a.workload}: {P} Gaussians, {n_views} keyframes {W}x{H}, SH degree {deg}",
                       "assignment": out["assignment"], "bucket_bytes": out["bucket_bytes"],
                       "grad_sink": not a.no_sink, "strips": a.strips, "cuda_graph": a.graph, "allreduce": "nccl" if a.peer == "off" or world == 1 else "peer:" + a.peer, "l2": "flushed between steps"}}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
