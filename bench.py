#!/usr/bin/env python
"""bench.py — rasterizer fwd+bwd frames/s on the BASELINE.json workload.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload config2]

One "step" = one forward + backward of the differentiable Gaussian rasterizer (pose gradients
on) for one view of the named synthetic scene (wildgs_slam_blackwell_b200/scenes.py).  At N > 1
(torchrun, one rank per GPU) every rank renders its own view of the replicated Gaussian set
and the Gaussian-gradient bucket is summed with one NCCL allreduce per step (weak scaling:
N frames per step).

Own arm: `value` = frames/s through the public autograd API with every input resident in HBM
and fixed cotangents; `e2e` = the same with all step inputs copied from pinned host memory,
a product-like L1 loss, and the loss + pose gradient read back, inside the timed region.
`roofline` = dominant kernel (stage events recorded inside the C library on the launching
stream) against MEASURED_PEAKS.json; `cpu_baseline` = oracle/ C restatement on host cores.
Reference arm (--impl reference): the UNMODIFIED reference rasterizer built for sm_100 in
oracle/_ref (its own __init__.py + _C.so), same scene, same two measurements.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from wildgs_slam_blackwell_b200 import scenes  # noqa: E402

METRIC = "rasterizer fwd+bwd frames/sec (640x480, 300k Gaussians)"
L2_FLUSH_BYTES = 512 << 20


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="config2", choices=sorted(scenes.CONFIGS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region (B200_PROFILING.md recipe)
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1]))
            except ValueError:
                continue
            for n, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------------------
def load_impl(impl):
    """The package a user imports: ours from the repo root, the reference from oracle/_ref."""
    if impl == "ours":
        import diff_gaussian_rasterization as mod
        assert os.path.dirname(os.path.dirname(os.path.abspath(mod.__file__))) == ROOT
    else:
        ref_dir = os.path.join(ROOT, "oracle", "_ref")
        if not os.path.exists(os.path.join(ref_dir, "diff_gaussian_rasterization", "_C.so")):
            raise FileNotFoundError("oracle/_ref/diff_gaussian_rasterization/_C.so missing "
                                    "(python oracle/ref_cuda/build_ref.py builds it)")
        sys.path.insert(0, ref_dir)
        import diff_gaussian_rasterization as mod
        assert os.path.abspath(mod.__file__).startswith(ref_dir)
    return mod


class _Set:
    """One set of device-resident step inputs (+ the autograd leaves / rasterizer bound to them).  The
    inputs are views into ONE flat device buffer so that a step's H2D traffic is a single copy."""

    def __init__(self, wl, mod, dev):
        self.flat = torch.empty_like(wl.host_flat, device=dev)
        self.flat.copy_(wl.host_flat)
        self.d = {k: self.flat[o:o + n].view(shape) for k, (o, n, shape) in wl.layout.items()}
        self.leaves = [self.d[k].requires_grad_(True) for k in ("means3D", "sh", "opacities", "scales", "rotations")]
        self.settings = mod.GaussianRasterizationSettings(
            image_height=wl.H, image_width=wl.W, tanfovx=wl.cam["tanfovx"], tanfovy=wl.cam["tanfovy"],
            bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=self.d["viewmatrix"],
            projmatrix=self.d["projmatrix"], projmatrix_raw=torch.from_numpy(wl.cam["projmatrix_raw"]).to(dev),
            sh_degree=wl.deg, campos=self.d["campos"], prefiltered=False, debug=False)
        self.rast = mod.GaussianRasterizer(self.settings)
        self.copy_done = torch.cuda.Event()
        self.compute_done = torch.cuda.Event()


class Workload:
    """One view of the scene on one device, driven through the public API of `mod`."""

    def __init__(self, mod, name, rank, dev):
        P, W, H, deg, _ = scenes.CONFIGS[name]
        seed = 1000 + int(name[-1])
        cam0 = scenes.make_camera(W, H, seed)
        self.g = scenes.make_gaussians(P, W, H, deg, seed, cam0)          # replicated Gaussian set
        self.cam = scenes.make_camera(W, H, seed + rank)                   # this rank's view
        dc, dd = scenes.make_cotangents(W, H, seed + rank)
        rng = np.random.default_rng(seed + 900 + rank)
        self.P, self.W, self.H, self.deg, self.dev, self.mod = P, W, H, deg, dev, mod
        # pinned host copies of everything a step consumes (e2e copies them every step)
        # ... packed into one pinned buffer (each array 256-B aligned): one H2D copy per step
        arrays = dict(means3D=self.g["means3D"], sh=self.g["sh"], opacities=self.g["opacities"],
                      scales=self.g["scales"], rotations=self.g["rotations"], viewmatrix=self.cam["viewmatrix"],
                      projmatrix=self.cam["projmatrix"], campos=self.cam["campos"],
                      target_color=rng.uniform(0, 1, (3, H, W)), target_depth=rng.uniform(0.5, 6, (1, H, W)))
        self.layout, off = {}, 0
        for k, a in arrays.items():
            a = np.asarray(a)
            self.layout[k] = (off, a.size, tuple(a.shape))
            off += (a.size + 63) // 64 * 64
        self.host_flat = torch.zeros(off, dtype=torch.float32).pin_memory()
        for k, a in arrays.items():
            o, n, _ = self.layout[k]
            self.host_flat[o:o + n] = torch.from_numpy(np.ascontiguousarray(a, np.float32).reshape(-1))
        # two input sets: the e2e loop copies step i+1's inputs (copy stream) while step i computes
        self.sets = [_Set(self, mod, dev), _Set(self, mod, dev)]
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.e2e_i = 0
        main = self.sets[0]
        self.d, self.leaves, self.settings, self.rast = main.d, main.leaves, main.settings, main.rast
        self.means2D = torch.zeros(P, 3, device=dev, requires_grad=True)
        self.theta = torch.zeros(3, device=dev, requires_grad=True)
        self.rho = torch.zeros(3, device=dev, requires_grad=True)
        self.dc, self.dd = torch.from_numpy(dc).to(dev), torch.from_numpy(dd).to(dev)
        self.out_host = torch.zeros(7, dtype=torch.float32).pin_memory()
        self.h2d_bytes = self.host_flat.numel() * 4
        self.d2h_bytes = 7 * 4
        self.bucket_floats = sum(t.numel() for t in self.leaves)
        # multi-GPU: the backward kernel writes the Gaussian gradients straight into one flat bucket
        # (rasterizer.grad_sink) that NCCL then sums in place -- no packing copy
        # ... which lives in NVLink peer memory and is summed by the library's own kernel (peer.PeerBucket,
        # csrc/peer_allreduce.cu) unless WGS_ALLREDUCE=nccl
        self.peer = None
        mode = os.environ.get("WGS_ALLREDUCE", "auto")
        if hasattr(mod, "grad_sink") and dist.is_initialized() and dist.get_world_size() > 1 and mode != "nccl":
            from wildgs_slam_blackwell_b200.peer import PeerBucket
            w = dist.get_world_size()
            self.peer = PeerBucket(self.bucket_floats, dev, use_multicast=(mode == "nvls" or (mode == "auto" and w >= 4)))
            self.bucket = self.peer.flat
        else:
            self.bucket = torch.zeros(self.bucket_floats, dtype=torch.float32, device=dev)
        self.sink = None
        self.fused_loss = None
        self.use_graph, self.graphs = hasattr(mod, "grad_sink"), {}
        if hasattr(mod, "grad_sink"):
            from wildgs_slam_blackwell_b200.loss import photometric_loss
            self.fused_loss = photometric_loss
            views, off = {}, 0
            for name, t in zip(("means3D", "sh", "opacity", "scales", "rotations"), self.leaves):
                views[name] = self.bucket[off:off + t.numel()].view(t.shape)
                off += t.numel()
            self.sink = mod.grad_sink(views)

    def _backward(self, world, fn):
        if world > 1 and self.sink is not None:
            with self.sink:
                fn()
            if self.peer is not None:
                self.peer.all_reduce()
            else:
                dist.all_reduce(self.bucket)
        else:
            fn()
            if world > 1:
                st = self.sets[0]
                dist.all_reduce(torch.cat([t.grad.reshape(-1) for t in st.leaves]))

    def _zero_grads(self, leaves):
        for t in leaves + [self.means2D, self.theta, self.rho]:
            t.grad = None

    def _forward(self, st):
        m3, sh, op, sc, ro = st.leaves
        return st.rast(m3, self.means2D, op, shs=sh, colors_precomp=None, scales=sc, rotations=ro,
                       cov3D_precomp=None, theta=self.theta, rho=self.rho)

    def step_resident(self, world):
        """fwd+bwd, inputs and cotangents resident in HBM (+ gradient allreduce when sharded)."""
        st = self.sets[0]
        self._zero_grads(st.leaves)
        color, radii, depth, opacity, n_touched = self._forward(st)
        self._backward(world, lambda: torch.autograd.backward([color, depth], [self.dc, self.dd]))

    def prefetch(self, i):
        """Enqueue the H2D copy of step i's inputs (all of them, from pinned memory) on the copy stream."""
        st = self.sets[i % 2]
        with torch.cuda.stream(self.copy_stream), torch.no_grad():
            self.copy_stream.wait_event(st.compute_done)      # the step that last used this set has finished
            st.flat.copy_(self.host_flat, non_blocking=True)
            st.copy_done.record(self.copy_stream)

    def _compute_e2e(self, st, world):
        self._zero_grads(st.leaves)
        color, radii, depth, opacity, n_touched = self._forward(st)
        if self.fused_loss is not None:
            # own arm: the fused loss + cotangent kernel -- same value and gradients
            loss, dcol, ddep, _ = self.fused_loss(color, st.d["target_color"], depth=depth,
                                                  gt_depth=st.d["target_depth"], rgb_boundary_threshold=-1e30,
                                                  depth_min=-1e30, color_weight=1.0, depth_weight=1.0)
            self._backward(world, lambda: torch.autograd.backward([color, depth], [dcol, ddep]))
        else:
            loss = (color - st.d["target_color"]).abs().mean() + (depth - st.d["target_depth"]).abs().mean()
            self._backward(world, loss.backward)
        res = torch.cat([loss.detach().reshape(1), self.rho.grad.reshape(-1), self.theta.grad.reshape(-1)])
        self.out_host.copy_(res, non_blocking=True)

    def step_e2e(self, world):
        """Host buffers in, loss + pose gradient out.  Every step's inputs cross PCIe inside the timed
        region; the copy for step i+1 is issued at the start of step i so that it overlaps compute."""
        i = self.e2e_i
        self.e2e_i += 1
        st = self.sets[i % 2]
        if i == 0:
            self.prefetch(0)
        cur = torch.cuda.current_stream()
        cur.wait_event(st.copy_done)
        self.prefetch(i + 1)
        if self.use_graph and (world == 1 or self.peer is not None):
            # own arm: the step's device work (forward, fused loss, backward, result gather, D2H) is captured
            # once per input set into a CUDA graph (graph.CapturedIteration) and replayed with one launch
            it = self.graphs.get(i % 2)
            if it is None:
                from wildgs_slam_blackwell_b200 import graph
                it = self.graphs[i % 2] = graph.CapturedIteration(lambda: self._compute_e2e(st, world), self.dev)
            it.replay()
        else:
            self._compute_e2e(st, world)
        st.compute_done.record(cur)


def timed_loop(fn, steps, warmup, flush, world):
    """W warm-up steps, then K steps each bracketed by CUDA events on the current stream with an L2
    flush between steps; barrier + synchronize on both sides.  Returns total ms (max over ranks)."""
    for _ in range(warmup):
        fn()
        flush.add_(1)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in evs:
        flush.add_(1)          # evict L2 (512 MiB > 126 MB) outside the timed bracket
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    per = [a.elapsed_time(b) for a, b in evs]
    total = torch.tensor([sum(per)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(total, op=dist.ReduceOp.MAX)
    return float(total.item()), per


# ------------------------------------------------------------------------------------------
def algorithmic_bytes(P, P_vis, R, N, tiles, D):
    """SURVEY.md §8(d) / DESIGN.md: algorithmic HBM bytes per launch of each stage."""
    c = (D + 1) ** 2
    return {
        "preprocess_fwd": P * (44 + 12 * c) + P_vis * 75 + (P - P_vis) * 8,
        "tile_scan": 8 * R + 8 * tiles,          # identifyTileRanges' share
        "scatter": 20 * P + 12 * R,              # duplicateWithKeys
        "tile_sort": 152 * R,                    # histogram + 6 radix passes of 24 B in the reference formulation
        "render_fwd": 44 * R + 28 * N,
        "render_bwd": 84 * R + 24 * N,
        "preprocess_bwd": P_vis * (155 + 24 * c),
    }


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def committed_traffic(kernel):
    """dram bytes/launch of `kernel` from the committed `ncu --set full` summary, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(kernel)
    except Exception:
        return None


def roofline_pass(wl, steps, flush):
    """Own arm only: K more steps with stage events on (inside the C library, on the launching
    stream); returns (roofline dict, per-stage table)."""
    from wildgs_slam_blackwell_b200 import _lib, rasterizer
    idx = torch.cuda.current_device()
    with torch.no_grad():
        m3, sh, op, sc, ro = wl.leaves
        e = torch.Tensor([])
        _, radii, _, _, _, st = rasterizer.forward_raw(m3, sh, e, op, sc, ro, e, wl.settings)
    P_vis, R = int((radii > 0).sum().item()), int(st.num_rendered)
    gx, gy = (wl.W + 15) // 16, (wl.H + 15) // 16
    alg = algorithmic_bytes(wl.P, P_vis, R, wl.W * wl.H, gx * gy, wl.deg)
    _lib.stage_timing(idx, True)
    acc = {}
    for _ in range(steps):
        flush.add_(1)
        wl.step_resident(1)
        for k, v in _lib.stage_times(idx).items():
            acc[k] = acc.get(k, 0.0) + v
    _lib.stage_timing(idx, False)
    peak, peak_src = measured_peaks()
    stages = {}
    for k, ms in acc.items():
        ms /= steps
        gbs = alg[k] / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        stages[k] = {"ms": round(ms, 5), "alg_bytes": int(alg[k]), "gbs": round(gbs, 1), "frac": round(gbs / peak, 4)}
    top = max(stages, key=lambda k: stages[k]["ms"])
    roof = {"kernel": top + "_kernel", "bound": "hbm", "achieved": stages[top]["gbs"], "peak": peak, "unit": "GB/s",
            "frac": stages[top]["frac"], "traffic": committed_traffic(top), "peak_source": peak_src,
            "ms_per_launch": stages[top]["ms"], "alg_bytes_per_launch": stages[top]["alg_bytes"],
            "note": "render stages are FP32-issue bound, not HBM bound; HBM fraction reported per contract, "
                    "issue_active_pct (ncu smsp__issue_active, committed profile) is the utilisation that bounds them"}
    issue = committed_traffic("_issue_active_pct")
    if isinstance(issue, dict) and top in issue:
        roof["issue_active_pct"] = issue[top]
    return roof, stages, dict(P_vis=P_vis, R=R)


def cpu_baseline(name, budget_s=20.0):
    """oracle/ C restatement (TEST INFRASTRUCTURE) timed on the host cores: fwd+bwd frames of the
    same workload, as many as fit the budget (>= 1)."""
    try:
        from oracle import cpu_oracle
        cpu_oracle.load()
    except Exception as ex:  # not built / not present: report, never substitute
        return {"value": None, "unit": "frames/s", "cores": 0, "kind": "port", "sample": f"unavailable: {ex}"}
    g, cams, cots = scenes.make_scene(name)
    threads = os.cpu_count() or 1
    n, t_total = 0, 0.0
    while n < 1 or (t_total < budget_s and n < 20):
        t0 = time.perf_counter()
        cpu_oracle.forward_backward(g, cams[0], cots[0], threads=threads)
        t_total += time.perf_counter() - t0
        n += 1
    return {"value": round(n / t_total, 4), "unit": "frames/s", "cores": threads, "kind": "port",
            "sample": f"{n} fwd+bwd frame(s) of {name} ({t_total:.1f} s of CPU work, OpenMP over tiles/Gaussians)"}


# ------------------------------------------------------------------------------------------
def main():
    a = parse()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference" and world > 1:
        # the reference has no multi-GPU mode: rank 0 alone measures it
        if rank != 0:
            return 0
        world, rank, local = 1, 0, 0
    if not torch.cuda.is_available():
        print(json.dumps({"impl": a.impl, "unavailable": "no CUDA device (the rasterizer has no CPU path)"}))
        return 1 if a.impl == "ours" else 0
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    try:
        mod = load_impl(a.impl)
    except Exception as ex:
        if a.impl == "reference":
            print(json.dumps({"impl": "reference", "unavailable": str(ex).splitlines()[0][:200]}))
            return 0
        raise
    wl = Workload(mod, a.workload, rank, dev)
    flush = torch.zeros(L2_FLUSH_BYTES // 4, dtype=torch.float32, device=dev)

    sampler = ClockSampler(local) if rank == 0 else None
    if a.impl == "ours":
        # no-wait forwards (graph.async_forward -> wgs_set_async): the host does not block on num_rendered between the
        # scan and the scatter kernel, so the device never idles; the instance capacity was sized by the eager step
        # above and is checked after the loop
        from wildgs_slam_blackwell_b200 import graph
        wl.step_resident(world)          # eager: sizes the instance capacity the no-wait forwards reuse
        torch.cuda.synchronize(dev)
        with graph.async_forward(dev):
            total_ms, per = timed_loop(lambda: wl.step_resident(world), a.steps, a.warmup, flush, world)
        graph.check_overflow(dev)
    else:
        total_ms, per = timed_loop(lambda: wl.step_resident(world), a.steps, a.warmup, flush, world)
    clocks = sampler.stop() if sampler else None
    frames = a.steps * world
    value = frames / (total_ms * 1e-3)

    e2e = None
    if not a.no_e2e:
        if a.impl == "ours":
            # no-wait forwards (wgs_set_async): the host does not block on num_rendered, so it can run ahead
            # of the device; the instance capacity (1.5x the count of the resident loop) is checked afterwards
            from wildgs_slam_blackwell_b200 import graph
            with graph.async_forward(dev):
                e2e_ms, _ = timed_loop(lambda: wl.step_e2e(world), a.steps, a.warmup, flush, world)
            graph.check_overflow(dev)
        else:
            e2e_ms, _ = timed_loop(lambda: wl.step_e2e(world), a.steps, a.warmup, flush, world)
        e2e = {"value": round(frames / (e2e_ms * 1e-3), 3), "unit": "frames/s", "h2d_bytes_per_step": wl.h2d_bytes,
               "d2h_bytes_per_step": wl.d2h_bytes, "ms_per_step": round(e2e_ms / a.steps, 4),
               "pipelining": "inputs of step i+1 copied (one packed H2D copy) on a second stream during step i "
                              "(double-buffered)" + ("; the step's device work replayed from a CUDA graph (graph.CapturedIteration), "
                                                     "instance-capacity overflow checked after the loop"
                                                     if a.impl == "ours" else ""),
               "loss": "mean|color - target| + mean|depth - target|: " +
                       ("fused loss+cotangent kernel (wgs_photometric_loss)" if wl.fused_loss is not None
                        else "torch expression + autograd (what the reference's callers run)")}

    roof = stages = counts = None
    if a.impl == "ours" and rank == 0:
        roof, stages, counts = roofline_pass(wl, min(a.steps, 50), flush)
    cpu = None
    if rank == 0 and not a.no_cpu_baseline and (a.impl == "ours" and world == 1 or a.impl == "reference"):
        cpu = cpu_baseline(a.workload)

    if rank == 0:
        P, W, H, deg, _ = scenes.CONFIGS[a.workload]
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": "frames/s", "n_gpus": a.gpus if a.impl == "ours" else 1,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(total_ms / a.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "impl": a.impl,
            "config": {"workload": f"{a.workload}: {P} Gaussians, {W}x{H}, SH degree {deg}, fwd+bwd with rho/theta "
                                   "pose gradients, one view per GPU per step",
                       "frames_per_step": world, "l2": "flushed between steps (512 MiB write)",
                       "forward": "no-wait (wgs_set_async), capacity overflow checked after the loop" if a.impl == "ours" else "stock",
                       "sharding": "views across ranks, replicated Gaussians, " +
                                   (("peer-memory allreduce kernel (" + ("NVLS multimem" if wl.peer.multicast else "P2P two-shot") + ")")
                                    if wl.peer is not None else "NCCL allreduce") +
                                   f" of the {wl.bucket_floats * 4} B gradient bucket per step" if world > 1 else "single GPU",
                       "p50_ms": round(statistics.median(per), 4), "p10_ms": round(sorted(per)[len(per) // 10], 4),
                       "p90_ms": round(sorted(per)[(len(per) * 9) // 10], 4)},
            "clocks": clocks,
            # our kernels per step: preprocess_fwd, tile_scan, scatter, tile_sort_small, tile_sort_large,
            # render_fwd, render_bwd, preprocess_bwd
            "gpu_launches": 8 * a.steps if a.impl == "ours" else 0,   # (+1 per e2e step: photometric_loss_kernel)
        }
        if e2e:
            line["e2e"] = e2e
        if roof:
            line["roofline"] = roof
            line["stages"] = stages
            line["counts"] = counts
        if cpu:
            line["cpu_baseline"] = cpu
        if a.impl == "reference":
            line["reference"] = "unmodified reference sources built for sm_100 (oracle/_ref), its own Python API"
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
