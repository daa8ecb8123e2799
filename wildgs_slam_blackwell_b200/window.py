"""Multi-GPU mapping window: views sharded over ranks, Gaussian gradients summed with one allreduce.

New capability (the reference renders ONE keyframe per optimiser step on one GPU,
src/mapper.py:1083-1093, and has no distributed code).  Semantics: the window gradient
is the sum over views of the per-view gradient the reference would have produced; per-view pose
gradients (rho/theta of each keyframe) are not summed, they are returned per view to every rank.

One process per GPU (torchrun); Gaussian parameters are replicated; `torch.distributed` (NCCL over
NVLink on GPUs, gloo in the CPU tests) is the plumbing.  Nothing here touches the rasterizer's
arithmetic: `render_loss` is whatever the caller renders with (the drop-in GaussianRasterizer).
"""
from typing import Callable, Dict, List, Optional, Sequence

import torch
import torch.distributed as dist


def assign_views(costs: Sequence[float], world: int) -> List[List[int]]:
    """Longest-processing-time-first assignment of views to ranks.  `costs[v]` is any per-view cost
    estimate (the previous step's num_rendered is a good one); ties and equal loads resolve to the
    lowest rank / lowest view index, so every rank computes the same assignment."""
    world = max(int(world), 1)
    order = sorted(range(len(costs)), key=lambda v: (-float(costs[v]), v))
    load = [0.0] * world
    out: List[List[int]] = [[] for _ in range(world)]
    for v in order:
        r = min(range(world), key=lambda i: (load[i], i))
        out[r].append(v)
        load[r] += max(float(costs[v]), 1e-9)
    for lst in out:
        lst.sort()
    return out


def assign_strips(costs: Sequence[float], world: int, tile_rows: int) -> List[List[tuple]]:
    """Sub-frame sharding: lay the views end to end (view v weighs costs[v], spread evenly over its
    `tile_rows` rows of 16 pixels), cut the line into `world` pieces of equal weight at tile-row
    granularity, and hand piece r to rank r.  Returns per rank a list of (view, row_begin, row_end);
    every (view, row) appears exactly once, a view is split between at most two neighbouring ranks
    whenever a rank's share is at least one view, and every rank computes the same plan.  Needed when
    the window does not divide by the world size (10 keyframes on 8 GPUs cap at 5x with whole views,
    SURVEY.md §8e)."""
    world = max(int(world), 1)
    n, rows = len(costs), int(tile_rows)
    w = [max(float(c), 1e-9) / rows for c in costs]           # weight of one tile row of view v
    total = sum(w[v] * rows for v in range(n))
    out: List[List[tuple]] = [[] for _ in range(world)]
    acc, r, begin_v, begin_row = 0.0, 0, 0, 0
    for v in range(n):
        begin_row = 0
        for row in range(rows):
            acc += w[v]
            # close rank r's piece after this row once it has reached its share (the last rank takes the rest)
            if r < world - 1 and acc >= total * (r + 1) / world - 1e-12 * total:
                out[r].append((v, begin_row, row + 1))
                begin_row = row + 1
                r += 1
        if begin_row < rows:
            out[r].append((v, begin_row, rows))
    return out


def assign_balanced(costs: Sequence[float], world: int, tile_rows: int) -> List[List[tuple]]:
    """Whole views first, strips only for the remainder.  With n views on `world` ranks every rank gets
    floor(n / world) WHOLE views (longest-processing-time first); the n mod world cheapest views are laid end to end
    and cut at tile-row granularity into one piece per rank, sized so that the ranks' total loads come out equal.
    Returns per rank a list of (view, row_begin, row_end), whole views first; every (view, row) appears exactly once
    and every rank computes the same plan.

    Against assign_strips (all views end to end) the number of work units is the same, but the partial units are all
    SMALL strips (10 keyframes on 8 GPUs: one whole view + a quarter of another per rank) -- the per-Gaussian stages of
    a small strip reject most Gaussians early (block-local compaction in preprocess_bwd), and a rank can run its strip
    on a second stream next to its whole view (WindowTrainer(streams=2))."""
    world = max(int(world), 1)
    n, rows = len(costs), int(tile_rows)
    c = [max(float(x), 1e-9) for x in costs]
    k, r = divmod(n, world)
    order = sorted(range(n), key=lambda v: (-c[v], v))
    whole_views, split_views = order[:n - r], sorted(order[n - r:])
    out: List[List[tuple]] = [[] for _ in range(world)]
    load = [0.0] * world
    for v in whole_views:                                   # LPT with exactly k views per rank
        i = min((i for i in range(world) if len(out[i]) < k), key=lambda i: (load[i], i))
        out[i].append((v, 0, rows))
        load[i] += c[v]
    for lst in out:
        lst.sort()
    if r:
        total = sum(c)
        target = [max(total / world - load[i], 0.0) for i in range(world)]      # what rank i still has to take
        scale = sum(c[v] for v in split_views) / max(sum(target), 1e-30)        # (targets sum to the remainder)
        target = [t * scale for t in target]
        i, acc = 0, 0.0
        for v in split_views:
            begin = 0
            w = c[v] / rows
            for row in range(rows):
                acc += w
                if i < world - 1 and acc >= target[i] - 1e-12 * total:
                    out[i].append((v, begin, row + 1))
                    begin, acc = row + 1, acc - target[i]
                    i += 1
                    while i < world - 1 and target[i] <= 1e-12 * total:         # (a rank whose whole views fill its share)
                        i += 1
            if begin < rows:
                out[i].append((v, begin, rows))
        loads = [sum(c[v] * (e - b) / rows for v, b, e in units) for units in out]
        if max(loads) > 1.08 * sum(c) / world:              # very uneven views: whole views cannot balance them
            return assign_strips(costs, world, tile_rows)
    return out


class GradBucket:
    """One fp32 buffer holding the gradients of all replicated leaves, plus per-leaf views into it.  Installing the
    views as `.grad` makes autograd accumulate straight into the buffer, so the exchange needs no packing copy.

    capacity=None: the leaves sit back to back (one flat range).  capacity=C (rows, >= P): every leaf owns a SEGMENT
    sized for C rows and uses its first P -- a model that grows by densification or shrinks by pruning re-targets the
    views (`retarget`) without re-allocating, so a bucket in symmetric peer memory is rendezvoused once per capacity
    class, not once per densification; only the used rows are exchanged (segment list of the peer kernel)."""

    def __init__(self, params: Dict[str, torch.Tensor], peer_group=None, peer_mode: str = "auto", side_floats: int = 0,
                 stat_floats: int = 0, capacity: Optional[int] = None):
        """stat_floats: floats between the gradients and the side block (`self.stats`), summed by the same allreduce --
        the window's densification statistics (2 per Gaussian; with a capacity: 2 segments of `capacity` floats).
        side_floats: extra floats behind the gradients (`self.side`) that ride in the same allreduce -- the window's
        per-view losses / costs / pose gradients, so that a step is ONE collective.
        peer_group: a process group whose ranks share one NVLink domain -> the buffer is allocated in
        symmetric peer memory (peer.PeerBucket) and all_reduce() runs the library's own kernel instead of
        ncclAllReduce.  peer_mode: 'nvls' | 'p2p' | 'auto' (multicast from 4 ranks up)."""
        self.names = list(params)
        first = next(iter(params.values()))
        P = first.shape[0]
        self.widths = {k: (p.numel() // P if P else int(torch.tensor(p.shape[1:]).prod())) for k, p in params.items()}
        self.tails = {k: tuple(p.shape[1:]) for k, p in params.items()}
        for k, p in params.items():
            if p.dtype != torch.float32:
                raise TypeError(f"GradBucket: leaf {k} must be float32")
        self.capacity = None if capacity is None else max((int(capacity) + 3) // 4 * 4, (P + 3) // 4 * 4)
        self._with_stats = stat_floats > 0
        rows = P if self.capacity is None else self.capacity
        self.offsets, off = {}, 0
        for k in self.names:
            self.offsets[k] = off
            off += rows * self.widths[k]
        self._grad_floats = off
        self._stat_off = off
        stat_total = (2 * rows if self.capacity is not None else stat_floats) if self._with_stats else 0
        self._side_off = off + stat_total
        total = self._side_off + side_floats
        self.peer = None
        if peer_group is not None and first.is_cuda:
            from .peer import PeerBucket
            world = dist.get_world_size(peer_group)
            mc = peer_mode == "nvls" or (peer_mode == "auto" and world >= 4)
            self.peer = PeerBucket(total, first.device, peer_group, use_multicast=mc)
            self._all = self.peer.flat
        else:
            self._all = torch.zeros(total, dtype=torch.float32, device=first.device)
        self.side = self._all[self._side_off:]
        self.views: Dict[str, torch.Tensor] = {}
        self.retarget(params)

    def retarget(self, params: Dict[str, torch.Tensor]):
        """Point the views at the first P rows of every segment (P = rows of `params`; same leaves, P <= capacity)."""
        P = next(iter(params.values())).shape[0]
        if list(params) != self.names or any(tuple(params[k].shape[1:]) != self.tails[k] for k in params) or \
                (self.capacity is not None and P > self.capacity) or (self.capacity is None and self.views and P != self.P):
            raise ValueError("GradBucket.retarget: other leaves, or more rows than the bucket's capacity")
        self.params, self.P = params, P
        for k in self.names:
            o = self.offsets[k]
            self.views[k] = self._all[o:o + P * self.widths[k]].view((P,) + self.tails[k])
        if self.capacity is None:
            self.flat = self._all[:self._grad_floats]
            self.stats = self._all[self._stat_off:self._side_off]
            self.stat_views = (self.stats[:P], self.stats[P:2 * P]) if self._with_stats else None
            self._segments = None
        else:
            self.flat = self._all[:self._grad_floats]      # (includes the unused rows of every segment: zero up to the
                                                           #  4-float pad, stale beyond it -- reduce over `views`, not `flat`)
            r4 = lambda n: (n + 3) // 4 * 4
            self._segments = [(self.offsets[k], r4(P * self.widths[k])) for k in self.names]
            if self._with_stats:
                c = self.capacity
                self.stat_views = (self._all[self._stat_off:self._stat_off + P], self._all[self._stat_off + c:self._stat_off + c + P])
                self._segments += [(self._stat_off, r4(P)), (self._stat_off + c, r4(P))]
            else:
                self.stat_views = None
            self.stats = self._all[self._stat_off:self._side_off]
            if self.side.numel():
                self._segments.append((self._side_off, r4(self.side.numel())))
            # The exchange works on whole 4-float groups, so up to 3 floats behind row P of every segment are summed
            # over the ranks on every step.  After a prune they would hold a dead row's gradients (and grow by the
            # world size per step): zero them here, and let zero_grads / zero_stats clear the padded length.
            n_leaf = len(self.names)
            self._zero_views = [self._all[o:o + n] for o, n in self._segments[:n_leaf]]
            self._zero_stat_views = [self._all[o:o + n] for o, n in self._segments[n_leaf:n_leaf + 2]] if self._with_stats else []
            with torch.no_grad():
                for (o, n), k in zip(self._segments[:n_leaf], self.names):
                    if n > P * self.widths[k]:
                        self._all[o + P * self.widths[k]:o + n].zero_()
                for o, n in (self._segments[n_leaf:n_leaf + 2] if self._with_stats else []):
                    if n > P:
                        self._all[o + P:o + n].zero_()

    @property
    def device(self):
        return self._all.device

    def matches(self, params: Dict[str, torch.Tensor]) -> bool:
        """True if `params` can use this bucket as it is laid out (same leaves; the same P, or P within the capacity)."""
        if list(params) != self.names or any(tuple(params[k].shape[1:]) != self.tails[k] or params[k].device != self._all.device
                                             for k in params):
            return False
        P = next(iter(params.values())).shape[0]
        return P == self.P if self.capacity is None else P <= self.capacity

    def zero_grads(self):
        if self.capacity is None:
            self.flat.zero_()
        else:
            for v in self._zero_views:       # the P rows and the 4-float rounding pad behind them
                v.zero_()

    def zero_stats(self):
        if self.stat_views is not None:
            for v in (self.stat_views if self.capacity is None else self._zero_stat_views):
                v.zero_()

    def install(self, zero=True):
        """Point every leaf's .grad at its slice (optionally zeroing the buffer first)."""
        if zero:
            self.zero_grads()
        for k, p in self.params.items():
            p.grad = self.views[k]

    @property
    def nbytes(self) -> int:
        return sum(v.numel() for v in self.views.values()) * 4

    def reduce_scatter(self):
        """First half of a sharded optimiser step (arena.GaussianArena.sharded_adam_step): the slice of the leaf
        gradients this rank owns receives the sum over the ranks; the small statistics / side segments are all-reduced
        (every rank reads them).  Needs a capacity-segmented bucket in peer memory."""
        if self.peer is None or self._segments is None:
            raise RuntimeError("GradBucket.reduce_scatter needs peer memory and a capacity (segmented) layout")
        n = len(self.names)
        self.peer.reduce_scatter(self._segments[:n])
        if len(self._segments) > n:
            self.peer.all_reduce(self._segments[n:])

    def all_reduce(self, group=None, async_op=False):
        """Sum the bucket over the ranks: the peer-memory kernel when allocated there, else NCCL/gloo."""
        if self.peer is not None:
            self.peer.all_reduce(self._segments)
            return None
        return dist.all_reduce(self._all, group=group, async_op=async_op)


class WindowTrainer:
    """Drives one window step.

        trainer = WindowTrainer(params)                  # params: name -> leaf tensor (requires_grad)
        out = trainer.step(n_views, render_loss)         # render_loss(v) -> (loss, pose_grad_fn)

    `render_loss(v)` renders view v against the replicated parameters and returns a scalar loss plus
    either None or a callable returning that view's pose gradient (6,) AFTER backward.  After step():
    every leaf's .grad holds the gradient summed over ALL views of the window (identical on every
    rank), `out["losses"]` (n_views,), `out["pose_grads"]` (n_views, 6), `out["assignment"]`.
    With densify_stats=True the backward kernels also accumulate the densification statistics of the window
    (rasterizer.densification_stats): after step(), `trainer.grad_accum` / `trainer.denom` (P,1) hold the sums over ALL
    views of ||dL_dmeans2D[:, :2]|| and of the visibility counts (add them to GaussianModel.xyz_gradient_accum / denom),
    `trainer.max_radii` the running maximum of this rank's views (reduce_max_radii() before it is used).
    """

    SIDE_FLOATS = 8 * 64   # (loss, cost, pose gradient) of up to 64 views ride behind the gradients in the bucket

    def _side(self, n_views):
        """(n_views, 8) zeroed side block: inside the bucket when it fits (one collective per step), else its own."""
        b = self.bucket
        if 8 * n_views <= b.side.numel():
            b.side.zero_()
            return b.side[:8 * n_views].view(n_views, 8), True
        return torch.zeros(n_views, 8, dtype=torch.float32, device=b.device), False

    def _finish(self, side, fused, cost_fn, assignment):
        b = self.bucket
        if self._exchange and self.exchange == "reduce_scatter":
            b.reduce_scatter()
            work = None
        elif self._exchange:
            work = b.all_reduce(self.group, async_op=True)
        if self._exchange:
            if not fused:
                dist.all_reduce(side, group=self.group)
            if work is not None:
                work.wait()
        side = side.clone() if fused else side
        if cost_fn is not None:   # measured costs feed the next assignment: the step's one host synchronisation
            self.costs = [float(c) for c in side[:, 1].tolist()]
        return {"losses": side[:, 0], "pose_grads": side[:, 2:8], "assignment": assignment, "bucket_bytes": b.nbytes}

    def __init__(self, params: Dict[str, torch.Tensor], group: Optional[dist.ProcessGroup] = None,
                 sink_names: Optional[Dict[str, str]] = None, peer: bool = False, peer_mode: str = "auto",
                 densify_stats: bool = False, streams: int = 1, emulate: Optional[tuple] = None,
                 capacity: Optional[int] = None, exchange: str = "allreduce"):
        """`sink_names` (optional): leaf name -> rasterizer gradient name ('means3D', 'sh', 'opacity',
        'scales', 'rotations', ...).  When given, the rasterizer's backward kernel writes / accumulates
        straight into the bucket (rasterizer.grad_sink) instead of autograd allocating per-view
        gradient tensors and adding them; requires that the leaves are passed to the rasterizer
        unmodified (no activation in between).
        streams=2 (needs sink_names): a rank's work units run on two CUDA streams, all of them accumulating
        into the bucket with reduction atomics -- the small kernels of one unit (scan, sort, per-Gaussian stages) and
        the tail waves of its compositing kernels overlap the other unit's work; inside a CUDA graph the two branches
        are scheduled by the device.  The bucket sum is then order-dependent in the last bit.
        emulate=(rank, world): measurement aid -- run the units rank `rank` of a `world`-rank job would run, in this
        single process, without the exchange (the per-rank compute time of an N-GPU step on one GPU)."""
        self.sink_names = sink_names
        self.group = group
        self.distributed = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if self.distributed else 1
        self.rank = dist.get_rank(group) if self.distributed else 0
        self._exchange = self.world > 1
        if emulate is not None:
            if self.distributed:
                raise RuntimeError("emulate=(rank, world) is for a single, non-distributed process")
            self.rank, self.world = int(emulate[0]), int(emulate[1])
        self.exchange = exchange      # 'allreduce' | 'reduce_scatter' (leaf gradients only summed on their owner rank)
        self.streams = int(streams) if sink_names is not None else 1
        self._side_stream = None
        # peer=True: bucket in NVLink peer memory, summed by wgs_peer_allreduce (single node, CUDA only)
        self._peer_group = (group if group is not None else dist.group.WORLD) if (peer and self.world > 1) else None
        self._peer_mode = peer_mode
        self._stats = bool(densify_stats)
        self.bucket = self._make_bucket(params, capacity)
        self.costs: List[float] = []

    def _make_bucket(self, params, capacity=None):
        P = next(iter(params.values())).shape[0]
        b = GradBucket(params, self._peer_group, self._peer_mode, self.SIDE_FLOATS, 2 * P if self._stats else 0,
                       capacity=capacity)
        self._bind_stats(b, P)
        return b

    def _bind_stats(self, b, P):
        if self._stats:
            # window sums of ||dL_dmeans2D[:, :2]|| and of the visibility counts ride in the bucket; the
            # screen-space radius needs a MAX: kept per rank, reduced on demand (reduce_max_radii)
            self.grad_accum, self.denom = b.stat_views[0].view(P, 1), b.stat_views[1].view(P, 1)
            old = getattr(self, "max_radii", None)
            if old is None or old.shape[0] != P:
                self.max_radii = torch.zeros(P, dtype=torch.float32, device=b.device)

    def _stats_ctx(self):
        if not self._stats:
            return _null()
        from . import rasterizer
        return rasterizer.densification_stats(self.grad_accum, self.denom, self.max_radii)

    @torch.no_grad()
    def reset_statistics(self):
        """Start a new densification interval: the running screen-space radius maximum (and the window sums, which
        zero_grads clears anyway) return to zero -- densification_postfix re-creates max_radii2D as zeros after every
        densify (gaussian_model.py:635-637); call it with GaussianArena.append / reset_statistics."""
        if self._stats:
            self.max_radii.zero_()
            self.grad_accum.zero_()
            self.denom.zero_()

    def reduce_max_radii(self):
        """max over the ranks of the running screen-space radius (GaussianModel.max_radii2D's update, mapper.py:1176-1179)"""
        if self.world > 1:
            dist.all_reduce(self.max_radii, op=dist.ReduceOp.MAX, group=self.group)
        return self.max_radii

    def broadcast_params(self, src: int = 0):
        """Replicate the parameter tensors from rank `src` (after densify / prune changed them there)."""
        if self.world > 1:
            for p in self.bucket.params.values():
                dist.broadcast(p.data, src=src, group=self.group)

    def rebind(self, params: Dict[str, torch.Tensor], capacity: Optional[int] = None):
        """Call after densify/prune changed the number of Gaussians.  A bucket with capacity headroom is re-targeted in
        place while the new row count fits (no allocation, no peer rendezvous); otherwise a new one is made, with
        `capacity` rows of headroom when given."""
        if self.bucket.matches(params):
            self.bucket.retarget(params)
            self._bind_stats(self.bucket, next(iter(params.values())).shape[0])
        else:
            self.bucket = self._make_bucket(params, capacity)

    def plan(self, n_views: int, tile_rows: Optional[int] = None, plan: Optional[str] = None) -> List[List[tuple]]:
        """Work units (view, row_begin, row_end) per rank; a whole view is (view, 0, tile_rows) -- or (view, 0, 0) when
        no tile_rows were given -- and is rendered without a strip.  plan: 'views' (whole views, LPT), 'strips' (all views end to end, cut per rank), 'balanced' (whole
        views + strips of the remainder); None = 'views', or 'strips' when tile_rows is given and the window does not
        divide by the world size (the behaviour before `plan` existed)."""
        if len(self.costs) != n_views:
            self.costs = [1.0] * n_views
        divides = self.world <= 1 or n_views % self.world == 0
        if plan is None:
            plan = "views" if (tile_rows is None or divides) else "strips"
        if plan != "views" and tile_rows is None:
            raise ValueError(f"plan='{plan}' needs tile_rows (= ceil(H / 16))")
        rows = int(tile_rows) if tile_rows is not None else 0
        if plan == "views" or divides:
            return [[(v, 0, rows) for v in views] for views in assign_views(self.costs, self.world)]
        return assign_strips(self.costs, self.world, rows) if plan == "strips" else assign_balanced(self.costs, self.world, rows)

    def step(self, n_views: int, render_loss: Callable, cost_fn: Optional[Callable[[int], float]] = None,
             tile_rows: Optional[int] = None, plan: Optional[str] = None):
        """`tile_rows` (= ceil(H/16)), when given and the window does not divide by the world size, switches to
        sub-frame units (see plan()): render_loss(v) is then called inside rasterizer.tile_rows(b, e) and
        must return a loss that is a sum over pixels (images are zero outside the strip); per-view losses,
        costs and pose gradients are summed over the strips of a view."""
        from . import rasterizer
        units = self.plan(n_views, tile_rows, plan)
        mine = units[self.rank]
        rows = int(tile_rows) if tile_rows is not None else 0
        is_whole = lambda u: u[1] == 0 and u[2] == rows
        strips = any(not is_whole(u) for lst in units for u in lst)
        if strips and self._stats:
            raise RuntimeError("densify_stats needs whole views (the gradient norm is per view): call step() without tile_rows")
        b = self.bucket
        use_sink = self.sink_names is not None and len(mine) > 0
        two = use_sink and self.streams > 1 and len(mine) > 1
        if use_sink:
            sink_bufs = {g: b.views[k] for k, g in self.sink_names.items()}
            for p in b.params.values():
                p.grad = None
            if two:
                b.zero_grads()                       # every unit adds atomically
        else:
            b.install()
        side, fused = self._side(n_views)  # loss, cost, pose grad (6)
        if self._stats:
            b.zero_stats()

        def run(unit, first):
            v, r0, r1 = unit
            whole = is_whole(unit)
            acc = "atomic" if two else (not first)
            # the contexts surround the FORWARD as well: the rasterizer binds them to that forward's backward, which
            # keeps units on two streams (and two trainers on two threads) apart (rasterizer._Scoped)
            with (rasterizer.tile_rows(r0, r1) if not whole else _null()), self._stats_ctx(), \
                    (rasterizer.grad_sink(sink_bufs, accumulate=acc) if use_sink else _null()):
                loss, pose_grad_fn = render_loss(v)
                loss.backward()
            frac = 1.0 if whole else (r1 - r0) / rows
            # (add_/+=: the strips of a view sum; fill-free so that nothing is copied from the host under capture)
            side[v, 0] += loss.detach()
            side[v, 1].add_(float(cost_fn(v)) * frac if cost_fn is not None else frac)
            if pose_grad_fn is not None:
                side[v, 2:8] += pose_grad_fn().reshape(6)

        if two:
            # longest-processing-time split of the rank's units over the two streams
            w = lambda u: self.costs[u[0]] * (1.0 if is_whole(u) else (u[2] - u[1]) / rows)
            lanes, load = ([], []), [0.0, 0.0]
            for u in sorted(mine, key=lambda u: -w(u)):
                i = 0 if load[0] <= load[1] else 1
                lanes[i].append(u)
                load[i] += w(u)
            dev = b.device
            if self._side_stream is None:
                self._side_stream = torch.cuda.Stream(dev)
            main = torch.cuda.current_stream(dev)
            self._side_stream.wait_stream(main)
            from . import graph
            # chained (programmatic dependent) launches off: the waiting grid of one stream would sit on the SM slots
            # the other stream's kernels are meant to fill
            with graph.chained_launch(False):
                with torch.cuda.stream(self._side_stream):
                    for u in lanes[1]:
                        run(u, False)
                for u in lanes[0]:
                    run(u, False)
            main.wait_stream(self._side_stream)
        else:
            for n_done, u in enumerate(mine):
                run(u, n_done == 0)
        if use_sink:
            b.install(zero=False)
        return self._finish(side, fused, cost_fn, units if strips else [[v for v, _, _ in u] for u in units])


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False
