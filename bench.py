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

METRIC = "rasterizer fwd+bwd frames/sec (640x480, 300k Gaussians); mapping iters/sec @1/2/4/8"   # BASELINE.json
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
    ap.add_argument("--no-window", action="store_true", help="skip the mapping-window half of the metric (config 3)")
    ap.add_argument("--window-steps", type=int, default=10)
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
        self.mem = None
        if wl.peer is not None:
            # multi-GPU: the buffer lives in NVLink peer memory, so that each rank uploads 1/world of the replicated
            # Gaussian set and the library's allgather kernel completes it (Workload.prefetch / _compute_e2e)
            from wildgs_slam_blackwell_b200.peer import PeerMemory
            self.mem = PeerMemory(wl.host_flat.numel(), dev, use_multicast=wl.peer.multicast)
            self.flat = self.mem.flat
        else:
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
        self.n_gauss_floats = self.layout["viewmatrix"][0]     # the replicated Gaussian set: the head of the buffer
        self.bucket_floats = sum(self.layout[k][1] for k in ("means3D", "sh", "opacities", "scales", "rotations"))
        # multi-GPU: the backward kernel writes the Gaussian gradients straight into one flat bucket
        # (rasterizer.grad_sink) ... which lives in NVLink peer memory and is summed by the library's own kernel
        # (peer.PeerBucket, csrc/peer_allreduce.cu) unless WGS_ALLREDUCE=nccl
        self.peer = None
        mode = os.environ.get("WGS_ALLREDUCE", "auto")
        if hasattr(mod, "grad_sink") and dist.is_initialized() and dist.get_world_size() > 1 and mode != "nccl":
            from wildgs_slam_blackwell_b200.peer import PeerBucket, owned_pieces
            w = dist.get_world_size()
            self.peer = PeerBucket(self.bucket_floats, dev, use_multicast=(mode == "nvls" or (mode == "auto" and w >= 4)))
            self.bucket = self.peer.flat
            # e2e: this rank's share of the replicated Gaussian set (the split the allgather kernel uses)
            self.upload_pieces = owned_pieces([(0, self.n_gauss_floats)], dist.get_rank(), w)
            self.upload_sync = PeerBucket(4, dev, use_multicast=self.peer.multicast)   # flags of the upload allgather
        else:
            self.bucket = torch.zeros(self.bucket_floats, dtype=torch.float32, device=dev)
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
        if self.peer is not None:
            self.h2d_bytes = 4 * (sum(n for _, n in self.upload_pieces) + self.host_flat.numel() - self.n_gauss_floats)
        self.d2h_bytes = 7 * 4
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
            if self.peer is None:
                st.flat.copy_(self.host_flat, non_blocking=True)
            else:
                # this rank's 1/world of the replicated Gaussian set + its own camera and targets; the other ranks'
                # shares arrive over NVLink (allgather kernel behind the copies)
                for o, n in self.upload_pieces:
                    st.flat[o:o + n].copy_(self.host_flat[o:o + n], non_blocking=True)
                g = self.n_gauss_floats
                st.flat[g:].copy_(self.host_flat[g:], non_blocking=True)
                # (on the copy stream, with its own flag area: it overlaps the compute of the step before)
                # a one-block rendezvous first: the wait for the slowest rank's copies is not spent by the whole
                # allgather grid on SM slots the running step's kernels need)
                self.upload_sync.rendezvous()
                self.upload_sync.all_gather(target=st.mem, segments=[(0, g)])
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
def algorithmic_bytes(P, P_vis, R, N, tiles, D, copies):
    """SURVEY.md §8(d) / DESIGN.md §4: algorithmic HBM bytes per launch of each stage.  tile_scan / tile_sort are
    charged with what THIS design moves (the reference's 6-pass radix formulation, 152 R bytes, is a different
    algorithm; dividing it by our time says nothing): the scan reads the privatised histogram and writes cursors,
    ranges and the tile order; the sort reads and writes each 8-byte bucket entry once."""
    c = (D + 1) ** 2
    return {
        "preprocess_fwd": P * (44 + 12 * c) + P_vis * 75 + (P - P_vis) * 8,
        "tile_scan": (2 * copies + 4) * 4 * tiles,
        "scatter": 20 * P + 12 * R,              # duplicateWithKeys
        "tile_sort": 16 * R + 8 * tiles,
        "render_fwd": 44 * R + 28 * N,
        "render_bwd": 84 * R + 24 * N,
        "preprocess_bwd": P_vis * (155 + 24 * c),
    }


def algorithmic_flops(n_eval, n_blend):
    """FP32 work of the compositing stages (add/mul/compare 1, FMA 2, expf = 1 MUFU + 2)."""
    return {"render_fwd": 15 * n_eval + 13 * n_blend, "render_bwd": 15 * n_eval + 80 * n_blend}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def committed_profile():
    """(dict, id) of the committed `ncu --set full` summary (profiles/traffic.json): dram bytes per launch and
    issue-slot utilisation per kernel.  Values read from it are labelled `from_profile` in the bench line -- they
    are properties of that capture, not of this run."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            d = json.load(f)
        return d, str(d.get("_source", "profiles/traffic.json")).split(":")[0]
    except Exception:
        return {}, None


def count_pairs(st):
    """N_eval, N_blend of the forward `st` describes (debug counting pass, wgs_count_pairs; outside any timed region)."""
    import ctypes as C
    from wildgs_slam_blackwell_b200 import _lib
    counts = torch.zeros(2, dtype=torch.int64, device="cuda")
    rc = _lib.load().wgs_count_pairs(C.byref(st.args), counts.data_ptr(),
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream))
    if rc != 0:
        raise RuntimeError(_lib.last_error())
    n_eval, n_blend = (int(x) for x in counts.tolist())
    return n_eval, n_blend


def roofline_pass(wl, steps, flush, clocks):
    """Own arm only: K more steps with stage events on (inside the C library, on the launching stream).  Per stage
    T_roof = max(bytes / HBM peak, flops / FP32 peak); the frame's `achieved` = sum of T_roof / sum of measured stage
    times.  Everything here is measured in this run except the entries labelled from_profile."""
    from wildgs_slam_blackwell_b200 import _lib, rasterizer
    idx = torch.cuda.current_device()
    with torch.no_grad():
        m3, sh, op, sc, ro = wl.leaves
        e = torch.Tensor([])
        _, radii, _, _, _, st = rasterizer.forward_raw(m3, sh, e, op, sc, ro, e, wl.settings)
        n_eval, n_blend = count_pairs(st)
    P_vis, R = int((radii > 0).sum().item()), int(st.num_rendered)
    gx, gy = (wl.W + 15) // 16, (wl.H + 15) // 16
    copies = 32
    while copies > 1 and copies * gx * gy > 65536:
        copies >>= 1
    alg = algorithmic_bytes(wl.P, P_vis, R, wl.W * wl.H, gx * gy, wl.deg, copies)
    flops = algorithmic_flops(n_eval, n_blend)
    _lib.stage_timing(idx, True)
    acc = {}
    for _ in range(steps):
        flush.add_(1)
        wl.step_resident(1)
        for k, v in _lib.stage_times(idx).items():
            acc[k] = acc.get(k, 0.0) + v
    _lib.stage_timing(idx, False)
    peak, peak_src = measured_peaks()
    props = torch.cuda.get_device_properties(idx)
    sm_mhz = (clocks or {}).get("sm_mhz") or (clocks or {}).get("sm_max_mhz") or 1965.0
    fp32_peak = props.multi_processor_count * 128 * 2 * sm_mhz * 1e6 / 1e12       # TFLOP/s at the clock sampled under load
    prof, prof_id = committed_profile()
    stages, t_roof_sum, t_meas_sum = {}, 0.0, 0.0
    for k, ms in acc.items():
        ms /= steps
        t_hbm = alg[k] / (peak * 1e9) * 1e3                                       # ms
        t_fp = flops.get(k, 0) / (fp32_peak * 1e12) * 1e3
        bound = "fp32" if t_fp > t_hbm else ("smem" if k in ("tile_sort", "tile_scan") else "hbm")
        t_roof = max(t_hbm, t_fp)
        gbs = alg[k] / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        stages[k] = {"ms": round(ms, 5), "bound": bound, "alg_bytes": int(alg[k]), "gbs": round(gbs, 1),
                     "hbm_frac": round(gbs / peak, 4), "t_roof_ms": round(t_roof, 5), "frac": round(t_roof / ms, 4) if ms > 0 else 0.0}
        if k in flops:
            stages[k].update(alg_flops=int(flops[k]), tflops=round(flops[k] / (ms * 1e-3) / 1e12, 3))
        t_roof_sum += t_roof
        t_meas_sum += ms
    top = max(stages, key=lambda k: stages[k]["ms"])
    tk = stages[top]
    roof = {"kernel": top + "_kernel", "bound": tk["bound"], "ms_per_launch": tk["ms"],
            "n_eval": n_eval, "n_blend": n_blend,
            "frame_t_roof_ms": round(t_roof_sum, 5), "frame_t_stages_ms": round(t_meas_sum, 5),
            "frame_achieved": round(t_roof_sum / t_meas_sum, 4) if t_meas_sum > 0 else None,
            "fp32_peak_tflops": round(fp32_peak, 2), "fp32_peak_source": f"{props.multi_processor_count} SMs x 128 x 2 x {sm_mhz:.0f} MHz (clock sampled during the timed loop)",
            "hbm_peak_gbs": peak, "hbm_peak_source": peak_src,
            "alg_bytes_per_launch": tk["alg_bytes"],
            "note": "per stage T_roof = max(bytes / HBM peak, flops / FP32 peak), frame_achieved = sum T_roof / sum of stage times "
                    "(SURVEY 8(d)); the stage event pairs add a few % over ms_per_step"}
    if tk["bound"] == "fp32":
        roof.update(achieved=tk["tflops"], peak=round(fp32_peak, 2), unit="TFLOP/s", frac=tk["frac"], alg_flops_per_launch=tk["alg_flops"])
    else:
        roof.update(achieved=tk["gbs"], peak=peak, unit="GB/s", frac=tk["hbm_frac"])
    # dram bytes / issue utilisation of the same kernel from the committed ncu capture -- NOT measured in this run
    roof["traffic"] = prof.get(top)
    roof["from_profile"] = {"id": prof_id, "traffic_bytes": prof.get(top),
                            "issue_active_pct": (prof.get("_issue_active_pct") or {}).get(top)}
    return roof, stages, dict(P_vis=P_vis, R=R, n_eval=n_eval, n_blend=n_blend)


def cpu_baseline(name, budget_s=15.0):
    """oracle/ C restatement (TEST INFRASTRUCTURE) timed on the host cores: fwd+bwd frames of the same workload on all
    cores, as many as fit the budget (>= 1), one frame of it on ONE thread, and BASELINE.md B2's config 1 (10 k
    Gaussians, 320x240) on all cores and on one."""
    try:
        from oracle import cpu_oracle
        cpu_oracle.load()
    except Exception as ex:  # not built / not present: report, never substitute
        return {"value": None, "unit": "frames/s", "cores": 0, "kind": "port", "sample": f"unavailable: {ex}"}
    threads = os.cpu_count() or 1

    def run(scene, nthreads, budget, cap):
        g, cams, cots = scene
        n, t_total = 0, 0.0
        while n < 1 or (t_total < budget and n < cap):
            t0 = time.perf_counter()
            cpu_oracle.forward_backward(g, cams[0], cots[0], threads=nthreads)
            t_total += time.perf_counter() - t0
            n += 1
        return n, t_total

    scene = scenes.make_scene(name)
    n, t_total = run(scene, threads, budget_s, 20)
    n1, t1 = run(scene, 1, 0.0, 1)
    out = {"value": round(n / t_total, 4), "unit": "frames/s", "cores": threads, "kind": "port",
           "sample": f"{n} fwd+bwd frame(s) of {name} ({t_total:.1f} s of CPU work, OpenMP over tiles/Gaussians)",
           "single_thread": {"value": round(n1 / t1, 4), "cores": 1, "sample": f"{n1} frame of {name} ({t1:.1f} s)"}}
    if name != "config1":
        c1 = scenes.make_scene("config1")
        na, ta = run(c1, threads, 2.0, 200)
        nb, tb = run(c1, 1, 2.0, 50)
        out["config1"] = {"all_cores": {"value": round(na / ta, 3), "cores": threads, "ms_per_frame": round(1e3 * ta / na, 3)},
                          "single_thread": {"value": round(nb / tb, 3), "cores": 1, "ms_per_frame": round(1e3 * tb / nb, 3)},
                          "sample": "config1: 10000 Gaussians, 320x240 (BASELINE.md B2)"}
    return out


def window_entry(a, dev, world, rank, flush):
    """The mapping-window half of BASELINE.json's metric: config 3 (1 M Gaussians, 10 keyframes 640x480), STRONG scaling
    over the ranks of this run through WindowTrainer (whole views + tile-row strips of the remainder, two streams per
    rank, the rank's step as one CUDA graph, NVLink peer exchange), with the multi-GPU gradient parity check."""
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    from window_workload import WindowWorkload
    wl = WindowWorkload("config3", dev, peer=os.environ.get("WGS_ALLREDUCE", "auto").replace("nccl", "off"),
                        plan="auto", streams=2, use_graph=os.environ.get("WGS_ALLREDUCE", "auto") != "nccl")
    ms, out = wl.time(a.window_steps, max(a.warmup, 3), flush)
    par = wl.parity() if world > 1 else None
    desc = wl.describe(out)
    # the same step ending like a mapping iteration (src/mapper.py:1212-1219): optimizer.step() inside the captured step
    wl.enable_optimizer()
    ms_opt, _ = wl.time(a.window_steps, max(a.warmup, 3), flush)
    with_opt = {"optimizer": "replicated fused Adam over the 5 leaves (1 launch, inside the captured step) on the exchanged bucket",
                "ms_per_step": round(ms_opt, 4), "value": round(1e3 / ms_opt, 3)}
    if world > 1 and wl.peer != "off":
        # and with the SHARDED optimiser step: reduce-scatter -> Adam on the owned 1/world of the rows -> allgather
        plan_used = wl.plan
        wl.enable_optimizer(sharded=True)
        wl.plan = plan_used
        ms_sh, _ = wl.time(a.window_steps, max(a.warmup, 3), flush)
        with_opt["sharded"] = {"optimizer": "reduce-scatter of the bucket, Adam on the rank's 1/world slice of the arena, "
                                            "allgather of the updated parameters (peer kernels)",
                               "ms_per_step": round(ms_sh, 4), "value": round(1e3 / ms_sh, 3)}
    # (the long description first, the numbers last: the driver's record keeps the END of the line)
    entry = {"config": desc, "metric": "mapping-window iterations/sec (config 3)", "scaling": "strong",
             "steps": a.window_steps, "overflow_checked": True, "with_optimizer": with_opt, "unit": "window steps/s",
             "frames_per_s": round(wl.n_views * 1e3 / ms, 2), "ms_per_step": round(ms, 4), "value": round(1e3 / ms, 3)}
    del wl
    torch.cuda.empty_cache()
    return entry, par


def reference_window(mod, dev, steps, warmup, flush):
    """Reference arm of the window metric: the unmodified reference rasterizer on one GPU, the per-view loop of
    src/mapper.py:1083-1093 over the 10 keyframes of config 3 (gradients accumulate in .grad through autograd)."""
    P, W, H, deg, n_views = scenes.CONFIGS["config3"]
    g, cams, cots = scenes.make_scene("config3")
    t = lambda x: torch.from_numpy(x).to(dev)
    prm = {k: t(g[k]).requires_grad_(True) for k in ("means3D", "sh", "opacities", "scales", "rotations")}
    m2 = torch.zeros(P, 3, device=dev, requires_grad=True)
    views = []
    for cam, cot in zip(cams, cots):
        st = mod.GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=torch.zeros(3, device=dev),
            scale_modifier=1.0, viewmatrix=t(cam["viewmatrix"]), projmatrix=t(cam["projmatrix"]),
            projmatrix_raw=t(cam["projmatrix_raw"]), sh_degree=deg, campos=t(cam["campos"]), prefiltered=False, debug=False)
        views.append((mod.GaussianRasterizer(st), t(cot[0]), t(cot[1]), torch.zeros(3, device=dev, requires_grad=True),
                      torch.zeros(3, device=dev, requires_grad=True)))

    def step():
        for p_ in list(prm.values()) + [m2]:
            p_.grad = None
        for rast, dc, dd, th, rh in views:
            th.grad = rh.grad = None
            color, radii, depth, opacity, n_touched = rast(prm["means3D"], m2, prm["opacities"], shs=prm["sh"],
                                                           scales=prm["scales"], rotations=prm["rotations"], theta=th, rho=rh)
            torch.autograd.backward([color, depth], [dc, dd])
    total_ms, _ = timed_loop(step, steps, warmup, flush, 1)
    ms = total_ms / steps
    return {"metric": "mapping-window iterations/sec (config 3)", "value": round(1e3 / ms, 3), "unit": "window steps/s",
            "ms_per_step": round(ms, 4), "frames_per_s": round(n_views * 1e3 / ms, 2), "scaling": "strong", "steps": steps,
            "config": {"workload": f"config3: {P} Gaussians, {n_views} keyframes {W}x{H}, SH degree {deg}",
                       "plan": "one GPU, views one after the other (the reference has no multi-GPU mode)"}}


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
    blocking = None
    if a.impl == "ours":
        # the same loop under the reference's own contract: every forward blocks the host until num_rendered is known
        b_ms, _ = timed_loop(lambda: wl.step_resident(world), a.steps, a.warmup, flush, world)
        blocking = {"value": round(frames / (b_ms * 1e-3), 3), "unit": "frames/s", "ms_per_step": round(b_ms / a.steps, 4),
                    "forward": "blocking (host waits for num_rendered, like the reference's D2H copy of R)"}

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
               "pipelining": ("inputs of step i+1 copied (one packed H2D copy) on a second stream during step i "
                              "(double-buffered)" if wl.peer is None else
                              "inputs of step i+1 copied on a second stream during step i (double-buffered): each rank uploads "
                              "1/world of the replicated Gaussian set + its own camera and targets, the library's allgather "
                              "kernel completes the set over NVLink on the copy stream") + ("; the step's device work replayed from a CUDA graph (graph.CapturedIteration), "
                                                     "instance-capacity overflow checked after the loop"
                                                     if a.impl == "ours" else ""),
               "loss": "mean|color - target| + mean|depth - target|: " +
                       ("fused loss+cotangent kernel (wgs_photometric_loss)" if wl.fused_loss is not None
                        else "torch expression + autograd (what the reference's callers run)")}

    roof = stages = counts = None
    if a.impl == "ours" and rank == 0:
        roof, stages, counts = roofline_pass(wl, min(a.steps, 50), flush, clocks)
    cpu = None
    if rank == 0 and not a.no_cpu_baseline and (a.impl == "ours" and world == 1 or a.impl == "reference"):
        cpu = cpu_baseline(a.workload)

    bucket_floats = wl.bucket_floats
    peer_desc = ("peer-memory allreduce kernel (" + ("NVLS multimem" if wl.peer.multicast else "P2P two-shot") + ")") if wl.peer is not None else "NCCL allreduce"
    window = parity = checks = None
    if a.impl == "ours" and world > 1:
        # the 2-GPU tests of tests/test_gpu_peer.py in a form that runs here, where the ranks exist: the exchange kernel
        # (both paths) against ncclAllReduce, sharded against replicated Adam across an append and a prune
        from wildgs_slam_blackwell_b200.selfcheck import multi_gpu_checks
        checks = multi_gpu_checks(dev)
    if not a.no_window and a.workload == "config2":
        del wl                      # free the headline workload's buffers (and its symmetric bucket) first
        torch.cuda.empty_cache()
        if a.impl == "ours":
            window, parity = window_entry(a, dev, world, rank, flush)
        else:
            window = reference_window(mod, dev, max(2, a.window_steps // 2), 2, flush)

    if rank == 0:
        P, W, H, deg, _ = scenes.CONFIGS[a.workload]
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": "frames/s", "n_gpus": a.gpus if a.impl == "ours" else 1,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(total_ms / a.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "impl": a.impl,
            # identical in both arms (the driver compares it); how each arm ran is under "run"
            "config": {"workload": f"{a.workload}: {P} Gaussians, {W}x{H}, SH degree {deg}, fwd+bwd with rho/theta "
                                   "pose gradients, one view per GPU per step",
                       "l2": "flushed between steps (512 MiB write)"},
            "run": {"frames_per_step": world,
                    "forward": "no-wait (wgs_set_async), capacity overflow checked after the loop; the blocking-forward "
                               "figure is under blocking_forward" if a.impl == "ours" else "stock (blocking)",
                    "sharding": "views across ranks, replicated Gaussians, " + peer_desc +
                                f" of the {bucket_floats * 4} B gradient bucket per step" if world > 1 else "single GPU",
                    "p50_ms": round(statistics.median(per), 4), "p10_ms": round(sorted(per)[len(per) // 10], 4),
                    "p90_ms": round(sorted(per)[(len(per) * 9) // 10], 4)},
            "clocks": clocks,
            # our kernels inside the timed region, per step: preprocess_fwd, tile_scan, scatter, tile_sort_small,
            # tile_sort_large, render_fwd, render_bwd, preprocess_bwd (+ peer_allreduce at N > 1)
            "gpu_launches": (8 + (1 if world > 1 and "peer" in peer_desc else 0)) * a.steps if a.impl == "ours" else 0,
        }
        if blocking:
            line["blocking_forward"] = blocking
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
        # the mapping-window half of the metric and its multi-GPU parity check close the line, so that a record which
        # keeps only the line's tail still carries them
        if window:
            line["window"] = window
        if parity:
            line["multi_gpu_parity"] = parity
        if checks:
            line["multi_gpu_checks"] = checks
        print(json.dumps(line), flush=True)
    rc = 0
    if rank == 0 and parity is not None and not parity.get("ok", False):
        sys.stderr.write(f"bench.py: multi-GPU gradient parity FAILED: {json.dumps(parity)}\n")
        rc = 3
    if rank == 0 and checks is not None and not checks.get("ok", False):
        sys.stderr.write(f"bench.py: multi-GPU self-checks FAILED: {json.dumps(checks)}\n")
        rc = 3
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
