"""Capacity-headroom arena for the Gaussian model's parameters, Adam moments and densification statistics
("fused Adam over the SoA parameter arena (and sharded Adam after reduce-scatter), densify/prune
that compacts the arena in place instead of re-creating nn.Parameters").

The reference grows and prunes the model by re-creating every tensor: `densification_postfix` /
`cat_tensors_to_optimizer` torch.cat six parameters and twelve Adam moments
(thirdparty/gaussian_splatting/scene/gaussian_model.py:566-644), `prune_points` / `_prune_optimizer` boolean-index
them (:526-564); `densify_and_prune` (:728-743) does both inside the mapping loop (src/mapper.py:1185-1198).  Every
such step changes every data pointer, so workspaces, a peer-memory gradient bucket and captured CUDA graphs keyed on
them die with it.

Here every leaf owns a SEGMENT of one flat buffer sized for `capacity` rows and uses its first P:

    params / exp_avg / exp_avg_sq : three buffers with the SAME segment layout (and the gradient bucket of the mapping
                                    window, window.GradBucket(capacity=...), a fourth) -> one offset addresses a
                                    parameter, its gradient and its moments
    append  (clone / split)       : rows written behind row P with zero moments, P += n, statistics of all rows
                                    restart at zero (densification_postfix)                        -- no allocation
    prune                         : surviving rows gathered (one mask scan + one gather launch over all 21 tensors,
                                    compaction.compact_rows) and copied back to the front -- pointers unchanged
    capacity class exceeded       : buffers of the next class (power of two) are allocated once and the rows move over;
                                    between two class changes nothing is allocated and a peer bucket is not re-rendezvoused
    Adam                          : ONE launch over the used rows of all leaves (wgs_adam_step on segment slices), or
                                    SHARDED: after a reduce-scatter of the gradient bucket each rank updates only the
                                    1/world slice it owns and the updated parameters return by an allgather over NVLink
                                    (peer.PeerBucket.reduce_scatter / all_gather) -- Adam's 28 B/element of HBM traffic
                                    shrink by the world size.
No CPU path: the kernels live in libwgs_rasterizer.so.
"""
import ctypes as C
import math
from typing import Dict, Optional

import torch

from . import _lib
from .adam import AdamTensor, _lib_adam
from .compaction import compact_rows

MIN_CAPACITY = 1 << 14


def capacity_class(rows: int) -> int:
    """Smallest power of two >= rows (>= MIN_CAPACITY): a model growing geometrically changes class O(log) times."""
    c = MIN_CAPACITY
    while c < rows:
        c <<= 1
    return c


class GaussianArena:
    STATS = ("xyz_gradient_accum", "denom", "max_radii2D")

    def __init__(self, leaves: Dict[str, torch.Tensor], capacity: Optional[int] = None, group=None, peer_mode: str = "auto",
                 betas=(0.9, 0.999), eps: float = 1e-15, capturable: bool = False):
        """leaves: name -> (P, ...) float32 CUDA tensor (initial values; e.g. xyz, f_dc, opacity, scaling, rotation).
        group: a torch.distributed group whose ranks share one NVLink domain -> the PARAMETER buffer lives in symmetric
        peer memory, so that a sharded optimiser step can return the updated rows to every rank (sharded_adam_step)."""
        first = next(iter(leaves.values()))
        if not first.is_cuda:
            raise RuntimeError("GaussianArena lives on a CUDA (sm_100) device; there is no CPU path")
        self.device = first.device
        self.names = list(leaves)
        self.tails = {k: tuple(v.shape[1:]) for k, v in leaves.items()}
        self.widths = {k: int(math.prod(v.shape[1:])) for k, v in leaves.items()}
        self.P = int(first.shape[0])
        self.group, self.peer_mode = group, peer_mode
        self.betas, self.eps = betas, float(eps)
        self.step_count = 0
        # capturable (torch's meaning): the step count lives on the device and the launch advances it, so an optimiser
        # step captured in a CUDA graph keeps its bias corrections moving on every replay
        self._step_dev = torch.zeros(1, dtype=torch.float32, device=self.device) if capturable else None
        self._sharded_by = None       # the PeerBucket whose ranks currently each hold only their slice of the moments
        self.allocations = 0          # capacity classes entered so far (each one: buffers + at most one peer rendezvous)
        self.capacity = 0
        self._alloc(capacity_class(max(self.P, capacity or 0)))
        for k, v in leaves.items():
            self._seg(self._param, k, self.P).copy_(v.detach().reshape(self.P, -1).reshape(-1))
        self._refresh()

    # ---- storage ----
    def _sync_moments(self):
        """After sharded steps every rank holds current Adam moments only for the slice it owns; gather them before the
        slices move (the row count changes) or a replicated step needs them all."""
        if self._sharded_by is not None:
            segs = self.segments()
            self._sharded_by.all_gather(self.peer_m, segs)
            self._sharded_by.all_gather(self.peer_v, segs)
            self._sharded_by = None

    def _alloc(self, capacity):
        old = (self._param, self._m, self._v, self._stats, self.offsets, self.P) if self.capacity else None
        self.capacity = int(capacity)
        self.offsets, off = {}, 0
        for k in self.names:
            self.offsets[k] = off
            off += self.capacity * self.widths[k]
        self._floats = off
        self.peer_params = self.peer_m = self.peer_v = None
        if self.group is not None:
            import torch.distributed as dist
            from .peer import PeerMemory
            world = dist.get_world_size(self.group)
            mc = self.peer_mode == "nvls" or (self.peer_mode == "auto" and world >= 4)
            # parameters AND moments in peer memory: a sharded step returns the updated parameter rows to every rank
            # each step; the moments stay with their owner and are gathered only when the ownership is about to change
            # (append / prune / capacity class / a replicated step) -- _sync_moments
            self.peer_params = PeerMemory(off, self.device, self.group, use_multicast=mc)
            self.peer_m = PeerMemory(off, self.device, self.group, use_multicast=mc)
            self.peer_v = PeerMemory(off, self.device, self.group, use_multicast=mc)
            self._param, self._m, self._v = self.peer_params.flat, self.peer_m.flat, self.peer_v.flat
        else:
            self._param = torch.zeros(off, dtype=torch.float32, device=self.device)
            self._m = torch.zeros(off, dtype=torch.float32, device=self.device)
            self._v = torch.zeros(off, dtype=torch.float32, device=self.device)
        self._stats = {n: torch.zeros(self.capacity, dtype=torch.float32, device=self.device) for n in self.STATS}
        self.allocations += 1
        if old is not None:
            o_param, o_m, o_v, o_stats, o_off, P = old
            for k in self.names:
                w = self.widths[k]
                for new, prev in ((self._param, o_param), (self._m, o_m), (self._v, o_v)):
                    new[self.offsets[k]:self.offsets[k] + P * w].copy_(prev[o_off[k]:o_off[k] + P * w])
            for n in self.STATS:
                self._stats[n][:P].copy_(o_stats[n][:P])

    def _seg(self, buf, k, rows, start=0):
        w = self.widths[k]
        o = self.offsets[k] + start * w
        return buf[o:o + rows * w]

    def _refresh(self):
        """New leaf views over the first P rows (the storage does not move)."""
        P = self.P
        self.params = {k: self._seg(self._param, k, P).view((P,) + self.tails[k]).detach().requires_grad_(True)
                       for k in self.names}
        self.xyz_gradient_accum = self._stats["xyz_gradient_accum"][:P].view(P, 1)
        self.denom = self._stats["denom"][:P].view(P, 1)
        self.max_radii2D = self._stats["max_radii2D"][:P]

    def moments(self, k):
        P = self.P
        return (self._seg(self._m, k, P).view((P,) + self.tails[k]), self._seg(self._v, k, P).view((P,) + self.tails[k]))

    def segments(self):
        """[(offset, length)] in floats of the used rows of every leaf (lengths rounded up to 4: the exchange kernel's
        granularity; the pad lies inside the segment's capacity)."""
        return [(self.offsets[k], (self.P * self.widths[k] + 3) // 4 * 4) for k in self.names]

    # ---- densify / prune ----
    def reserve(self, rows):
        """Make room for `rows` rows; returns True if a new capacity class had to be allocated (the caller then re-binds
        whatever held pointers: WindowTrainer.rebind(arena.params, capacity=arena.capacity))."""
        if rows <= self.capacity:
            return False
        self._sync_moments()
        self._alloc(capacity_class(rows))
        self._refresh()
        return True

    @torch.no_grad()
    def append(self, rows: Dict[str, torch.Tensor]):
        """densification_postfix + cat_tensors_to_optimizer (gaussian_model.py:566-644): the new Gaussians' rows go
        behind row P and their Adam moments start at zero; the densification statistics of ALL rows restart at zero,
        as the reference re-creates xyz_gradient_accum, denom and max_radii2D after every densify (:635-637), also when
        nothing was added.  A WindowTrainer feeding add_window_statistics keeps a running maximum of its own:
        trainer.reset_statistics() goes with this call.  Returns True if the capacity class changed."""
        n = int(next(iter(rows.values())).shape[0])
        if n == 0:
            self.reset_statistics()
            return False
        self._sync_moments()
        grew = self.reserve(self.P + n)
        P = self.P
        for k in self.names:
            self._seg(self._param, k, n, start=P).copy_(rows[k].detach().reshape(-1))
            self._seg(self._m, k, n, start=P).zero_()
            self._seg(self._v, k, n, start=P).zero_()
        self.P = P + n
        self.reset_statistics()
        self._refresh()
        return grew

    @torch.no_grad()
    def reset_statistics(self):
        """xyz_gradient_accum, denom, max_radii2D <- 0 for every row: a new densification interval begins
        (gaussian_model.py:635-637)."""
        for s in self.STATS:
            self._stats[s][:self.P].zero_()

    @torch.no_grad()
    def prune(self, remove_mask: torch.Tensor):
        """prune_points + _prune_optimizer (gaussian_model.py:526-564): rows where `remove_mask` is set disappear, the
        survivors keep their order and move to the front of every segment -- one mask scan and one gather launch for all
        21 tensors, then a copy back; no pointer changes.  Returns the number of surviving rows (one host sync)."""
        self._sync_moments()
        P = self.P
        keep = ~remove_mask
        srcs = []
        for k in self.names:
            for buf in (self._param, self._m, self._v):
                srcs.append(self._seg(buf, k, P).view(P, self.widths[k]))
        for s in self.STATS:
            srcs.append(self._stats[s][:P].view(P, 1))
        outs = compact_rows(keep, srcs)
        n_keep = int(outs[0].shape[0])
        for src, out in zip(srcs, outs):
            src[:n_keep].copy_(out)
        self.P = n_keep
        self._refresh()
        return n_keep

    # ---- optimiser ----
    def _adam(self, pieces, grad_base: torch.Tensor, lrs: Dict[str, float]):
        """pieces: [(offset, length)] in floats, each inside one leaf's segment; grad_base: the gradient bucket's buffer
        (same segment layout).  One wgs_adam_step launch."""
        step = self.step_count + 1
        b1, b2 = self.betas
        descs = []
        bounds = [(self.offsets[k], self.offsets[k] + self.capacity * self.widths[k], k) for k in self.names]
        for off, length in pieces:
            k = next(name for lo, hi, name in bounds if lo <= off < hi)
            n = min(length, self.offsets[k] + self.P * self.widths[k] - off)      # (drop the 4-float rounding pad)
            if n <= 0:
                continue
            t = AdamTensor()
            t.param = self._param.data_ptr() + 4 * off
            t.grad = grad_base.data_ptr() + 4 * off
            t.exp_avg, t.exp_avg_sq = self._m.data_ptr() + 4 * off, self._v.data_ptr() + 4 * off
            t.n = n
            if self._step_dev is not None:
                t.step_dev = self._step_dev.data_ptr()
                t.step_size = float(lrs[k])
            else:
                t.step_size = float(lrs[k]) / (1.0 - b1 ** step)
                t.bias_correction2_sqrt = math.sqrt(1.0 - b2 ** step)
            descs.append(t)
        if descs:
            arr = (AdamTensor * len(descs))(*descs)
            with torch.cuda.device(self.device):
                rc = _lib_adam().wgs_adam_step(arr, len(descs), b1, b2, self.eps,
                                               C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
            if rc != _lib.WGS_OK:
                raise RuntimeError(f"wgs_adam_step failed ({rc}): {_lib.last_error()}")
        elif self._step_dev is not None:
            self._step_dev += 1.0       # (a rank that owns nothing still counts the step)

    def _check_bucket(self, bucket):
        if bucket.capacity != self.capacity or bucket.names != self.names or any(bucket.offsets[k] != self.offsets[k] for k in self.names):
            raise RuntimeError("the gradient bucket must share the arena's segment layout: "
                               "WindowTrainer(arena.params, capacity=arena.capacity) / rebind(arena.params, capacity=...)")

    @torch.no_grad()
    def adam_step(self, bucket, lrs: Dict[str, float]):
        """Replicated step: every rank updates all used rows from the (all-reduced) gradient bucket."""
        self._check_bucket(bucket)
        self._sync_moments()
        self._adam([(self.offsets[k], self.P * self.widths[k]) for k in self.names], bucket._all, lrs)
        self.step_count += 1

    @torch.no_grad()
    def sharded_adam_step(self, bucket, lrs: Dict[str, float]):
        """Sharded step (call INSTEAD of the bucket's allreduce: WindowTrainer(exchange='reduce_scatter')): the gradient
        slice this rank owns has been summed over the ranks by bucket.peer.reduce_scatter(arena.segments()); Adam runs on
        that slice only; the updated parameter rows return to every rank by an allgather over NVLink."""
        self._check_bucket(bucket)
        if bucket.peer is None or self.peer_params is None:
            raise RuntimeError("sharded_adam_step needs the bucket and the arena in peer memory (group=...)")
        from .peer import owned_pieces
        segs = self.segments()
        self._adam(owned_pieces(segs, bucket.peer.rank, bucket.peer.world), bucket._all, lrs)
        bucket.peer.all_gather(self.peer_params, segs)
        self._sharded_by = bucket.peer
        self.step_count += 1

    @torch.no_grad()
    def add_window_statistics(self, trainer, reduce_max=False):
        """Fold the window's summed densification statistics (WindowTrainer(densify_stats=True)) into the arena's
        running ones -- GaussianModel.add_densification_stats over all views of the window.  The screen-space radius is a
        MAX: this rank's views by default (capturable), over all ranks with reduce_max=True (an NCCL MAX; do it once,
        before the radii are used)."""
        self.xyz_gradient_accum += trainer.grad_accum
        self.denom += trainer.denom
        torch.maximum(self.max_radii2D, trainer.reduce_max_radii() if reduce_max else trainer.max_radii, out=self.max_radii2D)
