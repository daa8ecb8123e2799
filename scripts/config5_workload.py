"""BASELINE.json config 5 -- the densification-stress workload -- as it is stated: the Gaussian count grows 100 k -> 2 M
geometrically over 40 steps inside the mapping loop, every step is a 16-keyframe window under the uncertainty-weighted
loss, with a prune every few steps; at 1 and at N GPUs.

What the reference does per mapping iteration (src/mapper.py:1083-1198): render a keyframe, `get_loss_mapping_uncertainty`
(weights 0.5 / u^2 with u clipped at 0.1, src/utils/slam_utils.py:224-232), backward, densification statistics
(:1176-1182), every `gaussian_update_every` iterations `densify_and_prune`
(thirdparty/gaussian_splatting/scene/gaussian_model.py:728-743 -> clone :646-672, split :674-726, prune :526-564), Adam.
Here, per growth step:

  1. prune (every `prune_every` steps: the 5 % least opaque Gaussians) and densify up to the schedule's row count:
     sources = the Gaussians with the largest accumulated view-space gradient, half of them CLONED (same parameters),
     half SPLIT (position jittered by the Gaussian's own scale, scale / 1.6) -- appends into the capacity-headroom arena
     (arena.GaussianArena): rows written behind row P, no tensor re-created, the peer-memory gradient bucket re-targeted
     in place; a new capacity class (power of two) is entered a handful of times over the whole run.
  2. the window step: the views / strips of this rank, raw-parameter rasterizer (activations fused), the fused
     uncertainty-weighted L1 loss kernel feeding the rasterizer backward, densification statistics fused into the
     backward and summed in the bucket; one exchange; Adam on the arena (replicated, or sharded after a reduce-scatter).
     The rank's step is captured as ONE CUDA graph per row count (the capture is redone when P changes) and the sticky
     overflow record is checked.

Identical on every rank without communication: parameters are replicated bit for bit, the choices above are pure
functions of them (topk / kthvalue on the device) and the jitter comes from a seeded device generator.
"""
import math
import time
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from wildgs_slam_blackwell_b200 import graph, rasterizer, scenes  # noqa: E402
from wildgs_slam_blackwell_b200.arena import GaussianArena  # noqa: E402
from wildgs_slam_blackwell_b200.loss import photometric_loss  # noqa: E402
from wildgs_slam_blackwell_b200.window import WindowTrainer  # noqa: E402

SINK_NAMES = {"xyz": "means3D", "f_dc": "sh", "opacity": "opacity", "scaling": "scales", "rotation": "rotations"}
LRS = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "opacity": 5e-2, "scaling": 1e-3, "rotation": 1e-3}   # configs/wildgs_slam.yaml opt_params


def schedule(p0=100_000, p1=2_000_000, steps=40):
    return [int(round(p0 * (p1 / p0) ** (s / (steps - 1)))) for s in range(steps)]


class Config5Workload:
    def __init__(self, dev, p0=100_000, p1=2_000_000, steps=40, n_views=16, W=640, H=480, peer="auto", plan="balanced",
                 streams=2, sharded=False, prune_every=10, prune_frac=0.05, use_graph=True, seed=1005, emulate=None,
                 sched=None, reserve=False):
        self.dev, self.W, self.H, self.n_views = dev, W, H, n_views
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.sched = list(sched) if sched is not None else schedule(p0, p1, steps)
        p0 = self.sched[0]
        self.prune_every, self.prune_frac, self.use_graph = prune_every, prune_frac, use_graph
        self.peer = peer if self.world > 1 else "off"
        self.sharded = sharded and self.world > 1
        if self.sharded and self.peer == "off":
            raise SystemExit("the sharded optimiser step needs the peer exchange")
        self.plan = plan
        cams = [scenes.make_camera(W, H, seed + k) for k in range(n_views)]
        g = scenes.make_gaussians(p0, W, H, 0, seed, cams[0])
        t = lambda x: torch.from_numpy(np.ascontiguousarray(x, np.float32)).to(dev)
        o = np.clip(g["opacities"], 1e-4, 1 - 1e-4)
        leaves = {"xyz": t(g["means3D"]), "f_dc": t(g["sh"][:, :1]), "opacity": t(np.log(o / (1 - o))),
                  "scaling": t(np.log(g["scales"])), "rotation": t(g["rotations"])}
        group = dist.group.WORLD if self.peer != "off" else None
        # reserve: the final capacity class is allocated (and, in peer memory, rendezvoused) ONCE, before the run -- 2 M rows
        # x 14 floats x 4 buffers is 450 MB of 180 GB -- so no growth step pays an allocation or a symmetric-memory exchange
        self.arena = GaussianArena(leaves, group=group, peer_mode=self.peer if self.peer != "off" else "auto",
                                   capturable=True, capacity=max(self.sched) if reserve else None)
        self.trainer = WindowTrainer(self.arena.params, sink_names=SINK_NAMES, peer=self.peer != "off",
                                     peer_mode=self.peer if self.peer != "off" else "auto", densify_stats=True,
                                     streams=streams, capacity=self.arena.capacity, emulate=emulate,
                                     exchange="reduce_scatter" if self.sharded else "allreduce")
        self.bucket_allocations = 1
        rng = np.random.default_rng(seed + 500)
        self.views = []
        for cam in cams:
            s = rasterizer.GaussianRasterizationSettings(
                image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=torch.zeros(3, device=dev),
                scale_modifier=1.0, viewmatrix=t(cam["viewmatrix"]), projmatrix=t(cam["projmatrix"]),
                projmatrix_raw=t(cam["projmatrix_raw"]), sh_degree=0, campos=t(cam["campos"]), prefiltered=False, debug=False)
            u = np.maximum(rng.lognormal(0.0, 0.5, (H, W)), 0.1)          # u ~ clip(LogNormal(0, 0.5), 0.1)
            self.views.append(dict(settings=s, target=t(rng.uniform(0, 1, (3, H, W))), unc=t(u),
                                   theta=torch.zeros(3, device=dev, requires_grad=True),
                                   rho=torch.zeros(3, device=dev, requires_grad=True)))
        self.gen = torch.Generator(device=dev)
        self.gen.manual_seed(seed)
        self.tile_rows = (H + 15) // 16
        self._it = None
        self._captured_before = False
        self._means2D = None

    # ---- one window step (+ optimiser) ----
    def render_loss(self, v):
        vw, p = self.views[v], self.arena.params
        vw["theta"].grad = vw["rho"].grad = None
        color, radii, depth, opacity, n_touched = rasterizer.rasterize_gaussian_model(
            p["xyz"], self._means2D, p["f_dc"], None, p["opacity"], p["scaling"], p["rotation"], vw["theta"], vw["rho"],
            vw["settings"])
        # uncertainty-weighted L1 (weights 0.5 / u^2, slam_utils.py:224-232): one fused kernel that also produces the
        # cotangents the rasterizer backward reads
        loss, dcol, ddep, _ = photometric_loss(color, vw["target"], uncertainty=vw["unc"], rgb_boundary_threshold=-1e30)

        class _Feed(torch.autograd.Function):        # loss value for the trainer's side block, cotangents for backward
            @staticmethod
            def forward(ctx, c):
                return loss.reshape(())

            @staticmethod
            def backward(ctx, gout):
                return dcol * gout
        return _Feed.apply(color), (lambda: torch.cat([vw["rho"].grad.reshape(-1), vw["theta"].grad.reshape(-1)]))

    def _step(self):
        out = self.trainer.step(self.n_views, self.render_loss, tile_rows=self.tile_rows, plan=self.plan)
        if self.sharded:
            self.arena.sharded_adam_step(self.trainer.bucket, LRS)
        else:
            self.arena.adam_step(self.trainer.bucket, LRS)
        self.arena.add_window_statistics(self.trainer)
        return out["losses"]

    def step(self):
        if not self.use_graph:
            with graph.async_forward(self.dev):
                return self._step()
        if self._it is None:
            # The capture's own warm-up pass is the eager (blocking) step that sizes the instance capacity at this row
            # count; the arena's leaves are new views after every grow_to and are first stepped there, on the capture's
            # side stream, so no full garbage collection is needed to drop accumulators bound to another stream
            # (profiles/r03d_capture_breakdown.json: eager pass 12 ms + gc 22-34 ms of the ~90 ms a first step cost).
            self._it = graph.CapturedIteration(self._step, self.dev, warmup=1, collect=0 if self._captured_before else 2)
            self._captured_before = True
        return self._it.replay()

    def check(self):
        return self._it.check() if self._it is not None else graph.check_overflow(self.dev)

    # ---- densify / prune between window steps ----
    @torch.no_grad()
    def grow_to(self, target_rows, prune):
        a = self.arena
        self._it = None                                # the row count is a launch parameter: recapture
        if prune:
            k = max(int(a.P * self.prune_frac), 1)
            thr = torch.kthvalue(a.params["opacity"].detach().reshape(-1), k).values
            a.prune(a.params["opacity"].detach().reshape(-1) <= thr)
        n_new = max(target_rows - a.P, 0)
        if n_new:
            score = (a.xyz_gradient_accum / a.denom.clamp_min(1.0)).reshape(-1)
            idx = torch.topk(score, min(n_new, a.P)).indices
            if idx.numel() < n_new:
                idx = idx.repeat((n_new + idx.numel() - 1) // idx.numel())[:n_new]
            split = torch.arange(n_new, device=self.dev) % 2 == 1
            rows = {k: p.detach()[idx].clone() for k, p in a.params.items()}
            jitter = torch.randn(n_new, 3, device=self.dev, generator=self.gen) * torch.exp(rows["scaling"])
            rows["xyz"] = torch.where(split[:, None], rows["xyz"] + jitter, rows["xyz"])
            rows["scaling"] = torch.where(split[:, None], rows["scaling"] - math.log(1.6), rows["scaling"])
            a.append(rows)
        before = self.trainer.bucket
        self.trainer.rebind(a.params, capacity=a.capacity)
        self.bucket_allocations += self.trainer.bucket is not before
        if n_new:
            self.trainer.reset_statistics()            # with arena.append: a new densification interval (:635-637)
        self._means2D = torch.zeros(a.P, 3, device=self.dev, requires_grad=True)

    # ---- the whole schedule ----
    def run(self, reps=3, log=None, profile_last=False):
        """Returns one record per growth step: rows, ms per window step (max over ranks, median of `reps` timed replays
        after one untimed), ms of the densify/prune that led to it."""
        flush = torch.zeros((512 << 20) // 4, device=self.dev)
        records = []
        for s, rows in enumerate(self.sched):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            self.grow_to(rows, prune=(s > 0 and s % self.prune_every == 0))
            e1.record()
            torch.cuda.synchronize()
            w0 = time.perf_counter()
            self.step()                                 # not in ms_per_window_step: eager pass + graph capture at this row count
            torch.cuda.synchronize()
            first_ms = torch.tensor([1e3 * (time.perf_counter() - w0), e0.elapsed_time(e1)], dtype=torch.float64, device=self.dev)
            if self.world > 1:
                dist.all_reduce(first_ms, op=dist.ReduceOp.MAX)
            if profile_last and s == len(self.sched) - 1:
                torch.cuda.synchronize()
                torch.cuda.profiler.start()             # ncu --profile-from-start off: only the last stage's steps
            times = []
            for _ in range(reps):
                flush.add_(1)
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record()
                losses = self.step()
                t1.record()
                times.append((t0, t1))
            torch.cuda.synchronize()
            self.check()
            ms = torch.tensor([sorted(a.elapsed_time(b) for a, b in times)[len(times) // 2]], dtype=torch.float64, device=self.dev)
            if self.world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            rec = {"step": s, "rows": self.arena.P, "capacity": self.arena.capacity, "ms_per_window_step": round(float(ms), 4),
                   "steps_per_s": round(1e3 / float(ms), 2), "densify_ms": round(float(first_ms[1]), 3),
                   "first_step_ms": round(float(first_ms[0]), 3),
                   "loss_sum": round(float(losses.sum()), 6)}
            records.append(rec)
            if log and self.rank == 0:
                log(rec)
        return records
