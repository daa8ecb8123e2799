"""The mapping-window workload (BASELINE.json configs 3 and 5's final shape) as an object: bench.py times it for the
`window` entry of its JSON line, scripts/bench_window.py is its command line.

One step = every keyframe of the window rendered + back-propagated against the replicated Gaussian set (views / tile-row
strips sharded over the ranks by WindowTrainer, the backward kernels accumulating straight into the flat gradient
bucket) + ONE exchange of the bucket (the library's NVLink peer kernel, or NCCL).  STRONG scaling: the window is fixed.
Per-view semantics follow the reference's mapping loop (src/mapper.py:1083-1093,1170: one keyframe per iteration),
summed over `window_size` keyframes (configs/wildgs_slam.yaml:34).  The loss is get_loss_mapping's L1 form
(slam_utils.py:116-143: alpha mean|rgb - gt| + (1 - alpha) mean|depth - gt_depth| on the valid pixels, alpha = 0.5 as in
configs/wildgs_slam.yaml:21) through the fused loss kernel; a tile-row strip evaluates it on its rows only (pixel mask),
so that the strips of a view add up to the view's loss.  loss="fixed": fixed cotangents instead (sum of color * dc +
depth * dd in torch ops), the form bench.py's headline frame uses."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import numpy as np  # noqa: E402

import diff_gaussian_rasterization as dgr  # noqa: E402
from wildgs_slam_blackwell_b200 import graph, rasterizer, scenes  # noqa: E402
from wildgs_slam_blackwell_b200.loss import _PhotometricLoss  # noqa: E402
from wildgs_slam_blackwell_b200.window import WindowTrainer  # noqa: E402

LEAVES = ("means3D", "sh", "opacities", "scales", "rotations")
SINK_NAMES = {"means3D": "means3D", "sh": "sh", "opacities": "opacity", "scales": "scales", "rotations": "rotations"}


class WindowWorkload:
    def __init__(self, name, dev, peer="auto", plan="balanced", streams=2, use_graph=True, sink=True, emulate=None,
                 scene=None, loss="mapping"):
        """peer: 'off' (NCCL) | 'auto' | 'nvls' | 'p2p'; plan: WindowTrainer.plan(); emulate=(rank, world): one process
        runs that rank's share without the exchange; scene: (gaussians, cameras, cotangents) to use instead of `name`."""
        self.name, self.dev = name, dev
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.P, self.W, self.H, self.deg, self.n_views = scenes.CONFIGS[name]
        g, cams, cots = scene if scene is not None else scenes.make_scene(name)
        self.P, self.n_views = g["means3D"].shape[0], len(cams)
        t = lambda x: torch.from_numpy(x).to(dev)
        self.params = {k: t(g[k]).requires_grad_(True) for k in LEAVES}
        self.peer = peer if self.world > 1 else "off"
        self.plan, self.use_graph, self.emulate, self.loss = plan, use_graph, emulate, loss
        if use_graph and self.world > 1 and self.peer == "off":
            raise SystemExit("a captured window step needs the peer exchange (an NCCL allreduce is not captured here)")
        self.streams, self.sink = streams, sink
        self.trainer = WindowTrainer(self.params, sink_names=SINK_NAMES if sink else None, peer=self.peer != "off",
                                     peer_mode=self.peer if self.peer != "off" else "auto", streams=streams, emulate=emulate)
        self.means2D = torch.zeros(self.P, 3, device=dev, requires_grad=True)
        self.views = []
        rng = np.random.default_rng(20260)
        for cam, cot in zip(cams, cots):
            s = dgr.GaussianRasterizationSettings(
                image_height=self.H, image_width=self.W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=t(cam["viewmatrix"]),
                projmatrix=t(cam["projmatrix"]), projmatrix_raw=t(cam["projmatrix_raw"]), sh_degree=self.deg,
                campos=t(cam["campos"]), prefiltered=False, debug=False)
            self.views.append(dict(rast=dgr.GaussianRasterizer(s), dc=t(cot[0]), dd=t(cot[1]),
                                   gt_color=t(rng.uniform(0.0, 1.0, (3, self.H, self.W)).astype(np.float32)),
                                   gt_depth=t(rng.uniform(0.5, 6.0, (1, self.H, self.W)).astype(np.float32)), strips={},
                                   theta=torch.zeros(3, device=dev, requires_grad=True),
                                   rho=torch.zeros(3, device=dev, requires_grad=True)))
        self.tile_rows = (self.H + 15) // 16
        self._it = None
        self._eager_out = None
        self.opt = self.arena = None

    def enable_optimizer(self, lrs=None, sharded=False):
        """From here on a step ends like the reference's mapping iteration (src/mapper.py:1212-1219): optimizer.step() on
        the window's summed gradients -- replicated fused Adam (adam.FusedAdam, capturable: ONE launch over the five
        leaves, inside the captured step), every rank updating all rows from the exchanged bucket.  Learning rates: the
        reference's (configs/wildgs_slam.yaml opt_params) scaled by 1e-2 so that the synthetic scene stays where the
        instance capacities were sized over a timed run; the work per step does not depend on them."""
        from wildgs_slam_blackwell_b200.adam import FusedAdam
        lrs = lrs or {"means3D": 1.6e-6, "sh": 2.5e-5, "opacities": 5e-4, "scales": 1e-5, "rotations": 1e-5}
        self._it, self.lrs = None, lrs
        if sharded:
            # SHARDED form (arena.GaussianArena.sharded_adam_step): the leaves move into an arena in peer memory, the
            # window's exchange becomes a reduce-scatter, each rank runs Adam on the 1/world of the rows it owns and the
            # updated rows return by an allgather -- the same bytes over NVLink as the one allreduce before, Adam's 28 B
            # per element of HBM traffic divided by the world size.
            if self.world < 2 or self.peer == "off" or self.emulate is not None or not self.sink:
                raise SystemExit("the sharded optimiser step needs >= 2 ranks, the peer exchange and the gradient sink")
            from wildgs_slam_blackwell_b200.arena import GaussianArena
            self.arena = GaussianArena({k: self.params[k].detach() for k in LEAVES}, group=dist.group.WORLD,
                                       peer_mode=self.peer, capturable=True)
            self.params = self.arena.params
            self.trainer = WindowTrainer(self.params, sink_names=SINK_NAMES, peer=True, peer_mode=self.peer,
                                         streams=self.streams, capacity=self.arena.capacity, exchange="reduce_scatter")
            self.opt = None
            return
        self.opt = FusedAdam([{"params": [self.params[k]], "lr": lrs[k], "name": k} for k in LEAVES], lr=0.0, eps=1e-15,
                             capturable=True)
        self._it = None

    # ---- one window step ----
    def render_loss(self, v):
        vw, p = self.views[v], self.params
        vw["theta"].grad = vw["rho"].grad = None
        color, radii, depth, opacity, n_touched = vw["rast"](
            p["means3D"], self.means2D, p["opacities"], shs=p["sh"], scales=p["scales"], rotations=p["rotations"],
            theta=vw["theta"], rho=vw["rho"])
        if self.loss == "fixed":
            loss = (color * vw["dc"]).sum() + (depth * vw["dd"]).sum()
        else:
            kw = dict(gt_color=vw["gt_color"], gt_depth=vw["gt_depth"], rgb_boundary_threshold=0.01, color_weight=0.5,
                      depth_weight=0.5)
            strip = rasterizer.tile_rows.current()
            if strip is not None:
                kw.update(self._strip_targets(vw, strip))
            loss = _PhotometricLoss.apply(color, depth, None, None, kw)
        return loss, (lambda: torch.cat([vw["rho"].grad.reshape(-1), vw["theta"].grad.reshape(-1)]))

    def _strip_targets(self, vw, strip):
        """pixel mask of the strip's rows (colour term) and the target depth with the other rows invalid (depth term)"""
        r0, r1 = strip.begin, strip.end
        if (r0, r1) not in vw["strips"]:
            m = torch.zeros(1, self.H, self.W, device=self.dev)
            m[:, 16 * r0:16 * r1] = 1.0
            vw["strips"][(r0, r1)] = dict(pixel_mask=m, gt_depth=vw["gt_depth"] * m)
        return vw["strips"][(r0, r1)]

    def eager_step(self, plan=None):
        return self.trainer.step(self.n_views, self.render_loss, tile_rows=self.tile_rows,
                                 plan=self.plan if plan is None else plan)

    def full_step(self):
        """eager_step + (after enable_optimizer) the optimiser step on the exchanged gradients"""
        out = self.eager_step()
        if self.arena is not None:
            self.arena.sharded_adam_step(self.trainer.bucket, self.lrs)
        if self.opt is not None:
            for k in LEAVES:                               # (sink mode leaves .grad unset: the bucket holds the gradient)
                if self.params[k].grad is None:
                    self.params[k].grad = self.trainer.bucket.views[k]
            self.opt.step()
        return out

    def step(self):
        if not self.use_graph:
            return self.full_step()
        if self._it is None:
            self._eager_out = self.full_step()             # also sizes the instance capacity per view
            self._it = graph.CapturedIteration(lambda: self.full_step()["losses"], self.dev, warmup=2)
        self._it.replay()
        return self._eager_out

    def check(self):
        """raises if any forward of any replay overflowed its captured instance capacity"""
        if self._it is not None:
            return self._it.check()
        return graph.check_overflow(self.dev)

    # ---- timing: W warm-up steps, K timed steps, L2 flushed between steps, max over ranks ----
    def choose_plan(self, flush, candidates=("balanced", "strips"), steps=3):
        """plan="auto": time a few steps of each candidate plan (max over ranks, so every rank picks the same one) and
        keep the faster -- whole views + strips of the remainder against contiguous strips of everything; which one wins
        depends on how the window divides by the world size."""
        self.plan_trials = {}
        if self.world == 1 and self.emulate is None:
            self.plan = candidates[0]            # one rank: every plan is "all views, whole"
            return
        for c in candidates:
            self.plan, self._it = c, None
            self.plan_trials[c] = round(self.time(steps, 2, flush)[0], 4)
        self.plan, self._it = min(self.plan_trials, key=self.plan_trials.get), None

    def time(self, steps, warmup, flush=None):
        if flush is None:
            flush = torch.zeros((512 << 20) // 4, device=self.dev)
        if self.plan == "auto":
            self.choose_plan(flush)
        for _ in range(max(warmup, 1)):
            out = self.step()
            flush.add_(1)
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        evs = []
        for _ in range(steps):
            flush.add_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = self.step()
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        total = torch.tensor([sum(x.elapsed_time(y) for x, y in evs)], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(total, op=dist.ReduceOp.MAX)
        self.check()
        return float(total.item()) / steps, out

    # ---- allreduced bucket == sum over views of single-GPU per-view gradients, whatever the assignment ----
    def parity(self, plans=("views", "balanced", "strips")):
        """Every rank renders ALL views whole through plain autograd (no sink, no strips, no exchange) and sums the
        per-view gradients; then one eager window step per plan, and the bucket every rank ends up with is compared
        with that sum.  Tolerance of the parity suite: |err| <= 1e-3 |ref| + 1e-4 max|ref| per tensor."""
        if self.emulate is not None:
            return None
        ref = {k: torch.zeros_like(p) for k, p in self.params.items()}
        ref_loss, ref_pose = [], []
        leaves = [self.params[k] for k in LEAVES]
        for p in leaves:
            p.grad = None
        for v in range(self.n_views):
            loss, pose = self.render_loss(v)
            grads = torch.autograd.grad(loss, leaves + [self.views[v]["rho"], self.views[v]["theta"]])
            for k, gk in zip(LEAVES, grads):
                ref[k] += gk
            ref_loss.append(loss.detach())
            ref_pose.append(torch.cat([grads[-2].reshape(-1), grads[-1].reshape(-1)]))
        ref_loss, ref_pose = torch.stack(ref_loss), torch.stack(ref_pose)
        res = {}
        for plan in plans:
            out = self.eager_step(plan)
            torch.cuda.synchronize()
            worst, rel = 0.0, 0.0
            for k in LEAVES:
                b, r = self.params[k].grad.double(), ref[k].double()
                scale = float(r.abs().max())
                err = (b - r).abs()
                worst = max(worst, float((err / (1e-3 * r.abs() + 1e-4 * scale + 1e-300)).max()))
                big = r.abs() >= 1e-2 * scale
                rel = max(rel, float((err[big] / r.abs()[big]).max()) if bool(big.any()) else 0.0)
            pose_err = float(((out["pose_grads"] - ref_pose).abs() / (1e-3 * ref_pose.abs() + 1e-4 * ref_pose.abs().max())).max())
            loss_err = float(((out["losses"] - ref_loss).abs() / (ref_loss.abs() + 1e-12)).max())
            stats = torch.tensor([worst, rel, pose_err, loss_err], dtype=torch.float64, device=self.dev)
            if self.world > 1:
                dist.all_reduce(stats, op=dist.ReduceOp.MAX)
            worst, rel, pose_err, loss_err = (float(x) for x in stats.tolist())
            res[plan] = {"max_rel_err": rel, "worst_tol_ratio": worst, "pose_tol_ratio": pose_err, "loss_rel_err": loss_err,
                         "ok": bool(worst <= 1.0 and pose_err <= 1.0 and loss_err <= 1e-4)}
        return {"plans": res, "max_rel_err": max(r["max_rel_err"] for r in res.values()),
                "ok": all(r["ok"] for r in res.values()), "n_ranks": self.world,
                "reference": "every rank: sum over all views of whole-view gradients through plain autograd on that rank alone",
                "tolerance": "|err| <= 1e-3 |ref| + 1e-4 max|ref| per tensor (worst_tol_ratio <= 1); max_rel_err over "
                             "entries with |ref| >= 1e-2 max|ref|"}

    def describe(self, out):
        b = self.trainer.bucket
        return {"workload": f"{self.name}: {self.P} Gaussians, {self.n_views} keyframes {self.W}x{self.H}, SH degree {self.deg}",
                "plan": self.plan, "plan_trials_ms": getattr(self, "plan_trials", None), "assignment": out["assignment"], "bucket_bytes": out["bucket_bytes"],
                "streams_per_rank": self.trainer.streams, "cuda_graph": self.use_graph,
                "exchange": ("none (one rank)" if self.world == 1 else "NCCL allreduce" if self.peer == "off" else
                             "peer-memory allreduce kernel (" + ("NVLS multimem" if b.peer.multicast else "P2P two-shot") + ")"),
                "l2": "flushed between steps (512 MiB write)"}
