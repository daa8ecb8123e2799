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


class GradBucket:
    """One flat fp32 buffer holding the gradients of all replicated leaves back to back, plus
    per-leaf views into it.  Installing the views as `.grad` makes autograd accumulate straight into
    the buffer, so the allreduce needs no packing copy."""

    def __init__(self, params: Dict[str, torch.Tensor], peer_group=None, peer_mode: str = "auto", side_floats: int = 0,
                 stat_floats: int = 0):
        """stat_floats: floats between the gradients and the side block (`self.stats`), summed by the same allreduce --
        the window's densification statistics.
        side_floats: extra floats behind the gradients (`self.side`) that ride in the same allreduce -- the window's
        per-view losses / costs / pose gradients, so that a step is ONE collective.
        peer_group: a process group whose ranks share one NVLink domain -> the buffer is allocated in
        symmetric peer memory (peer.PeerBucket) and all_reduce() runs the library's own kernel instead of
        ncclAllReduce.  peer_mode: 'nvls' | 'p2p' | 'auto' (multicast from 4 ranks up)."""
        self.names = list(params)
        self.params = params
        n = sum(p.numel() for p in params.values())
        first = next(iter(params.values()))
        self.peer = None
        if peer_group is not None and first.is_cuda:
            from .peer import PeerBucket
            world = dist.get_world_size(peer_group)
            mc = peer_mode == "nvls" or (peer_mode == "auto" and world >= 4)
            self.peer = PeerBucket(n + side_floats + stat_floats, first.device, peer_group, use_multicast=mc)
            self._all = self.peer.flat
        else:
            self._all = torch.zeros(n + side_floats + stat_floats, dtype=torch.float32, device=first.device)
        self.flat, self.stats, self.side = self._all[:n], self._all[n:n + stat_floats], self._all[n + stat_floats:]
        self.views: Dict[str, torch.Tensor] = {}
        off = 0
        for k, p in params.items():
            if p.dtype != torch.float32:
                raise TypeError(f"GradBucket: leaf {k} must be float32")
            self.views[k] = self.flat[off:off + p.numel()].view(p.shape)
            off += p.numel()

    def matches(self, params: Dict[str, torch.Tensor]) -> bool:
        return list(params) == self.names and all(params[k].shape == self.views[k].shape and
                                                  params[k].device == self.flat.device for k in params)

    def install(self, zero=True):
        """Point every leaf's .grad at its slice (optionally zeroing the buffer first)."""
        if zero:
            self.flat.zero_()
        for k, p in self.params.items():
            p.grad = self.views[k]

    @property
    def nbytes(self) -> int:
        return self.flat.numel() * 4

    def all_reduce(self, group=None, async_op=False):
        """Sum the bucket over the ranks: the peer-memory kernel when allocated there, else NCCL/gloo."""
        if self.peer is not None:
            self.peer.all_reduce()
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
        return torch.zeros(n_views, 8, dtype=torch.float32, device=b.flat.device), False

    def _finish(self, side, fused, cost_fn, assignment):
        b = self.bucket
        if self.world > 1:
            work = b.all_reduce(self.group, async_op=True)
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
                 densify_stats: bool = False):
        """`sink_names` (optional): leaf name -> rasterizer gradient name ('means3D', 'sh', 'opacity',
        'scales', 'rotations', ...).  When given, the rasterizer's backward kernel writes / accumulates
        straight into the bucket (rasterizer.grad_sink) instead of autograd allocating per-view
        gradient tensors and adding them; requires that the leaves are passed to the rasterizer
        unmodified (no activation in between)."""
        self.sink_names = sink_names
        self.group = group
        self.distributed = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if self.distributed else 1
        self.rank = dist.get_rank(group) if self.distributed else 0
        # peer=True: bucket in NVLink peer memory, summed by wgs_peer_allreduce (single node, CUDA only)
        self._peer_group = (group if group is not None else dist.group.WORLD) if (peer and self.world > 1) else None
        self._peer_mode = peer_mode
        self._stats = bool(densify_stats)
        self.bucket = self._make_bucket(params)
        self.costs: List[float] = []

    def _make_bucket(self, params):
        P = next(iter(params.values())).shape[0]
        b = GradBucket(params, self._peer_group, self._peer_mode, self.SIDE_FLOATS, 2 * P if self._stats else 0)
        if self._stats:
            # window sums of ||dL_dmeans2D[:, :2]|| and of the visibility counts ride in the bucket; the
            # screen-space radius needs a MAX: kept per rank, reduced on demand (reduce_max_radii)
            self.grad_accum, self.denom = b.stats[:P].view(P, 1), b.stats[P:].view(P, 1)
            self.max_radii = torch.zeros(P, dtype=torch.float32, device=b.flat.device)
        return b

    def _stats_ctx(self):
        if not self._stats:
            return _null()
        from . import rasterizer
        return rasterizer.densification_stats(self.grad_accum, self.denom, self.max_radii)

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

    def rebind(self, params: Dict[str, torch.Tensor]):
        """Call after densify/prune changed the number of Gaussians."""
        if not self.bucket.matches(params):
            self.bucket = self._make_bucket(params)
        else:
            self.bucket.params = params

    def step(self, n_views: int, render_loss: Callable, cost_fn: Optional[Callable[[int], float]] = None,
             tile_rows: Optional[int] = None):
        """`tile_rows` (= ceil(H/16)), when given and the window does not divide by the world size, switches to
        sub-frame units (assign_strips): render_loss(v) is then called inside rasterizer.tile_rows(b, e) and
        must return a loss that is a sum over pixels (images are zero outside the strip); per-view losses,
        costs and pose gradients are summed over the strips of a view."""
        if len(self.costs) != n_views:
            self.costs = [1.0] * n_views
        if tile_rows is not None and self.world > 1 and n_views % self.world != 0:
            return self._step_strips(n_views, render_loss, cost_fn, int(tile_rows))
        assignment = assign_views(self.costs, self.world)
        mine = assignment[self.rank]
        b = self.bucket
        use_sink = self.sink_names is not None and len(mine) > 0
        if use_sink:
            from . import rasterizer
            sink_bufs = {g: b.views[k] for k, g in self.sink_names.items()}
            for p in b.params.values():
                p.grad = None
        else:
            b.install()
        side, fused = self._side(n_views)  # loss, cost, pose grad (6)
        if self._stats:
            b.stats.zero_()
        for n_done, v in enumerate(mine):
            # the contexts surround the FORWARD as well: the rasterizer binds them to that forward's backward, which
            # keeps two trainers on two threads / streams of one process apart (rasterizer._Scoped)
            with self._stats_ctx(), (rasterizer.grad_sink(sink_bufs, accumulate=n_done > 0) if use_sink else _null()):
                loss, pose_grad_fn = render_loss(v)
                loss.backward()
            side[v, 0] = loss.detach()
            side[v, 1].fill_(float(cost_fn(v)) if cost_fn is not None else 1.0)  # fill_: no H2D copy (graph capture)
            if pose_grad_fn is not None:
                side[v, 2:8] = pose_grad_fn().reshape(6)
        if use_sink:
            b.install(zero=False)
        return self._finish(side, fused, cost_fn, assignment)

    def _step_strips(self, n_views, render_loss, cost_fn, tile_rows):
        from . import rasterizer
        if self._stats:
            raise RuntimeError("densify_stats needs whole views (the gradient norm is per view): call step() without tile_rows")
        plan = assign_strips(self.costs, self.world, tile_rows)
        mine = plan[self.rank]
        b = self.bucket
        use_sink = self.sink_names is not None and len(mine) > 0
        if use_sink:
            sink_bufs = {g: b.views[k] for k, g in self.sink_names.items()}
            for p in b.params.values():
                p.grad = None
        else:
            b.install()
        side, fused = self._side(n_views)
        for n_done, (v, r0, r1) in enumerate(mine):
            whole = r0 == 0 and r1 == tile_rows
            with (rasterizer.tile_rows(r0, r1) if not whole else _null()), \
                    (rasterizer.grad_sink(sink_bufs, accumulate=n_done > 0) if use_sink else _null()):
                loss, pose_grad_fn = render_loss(v)
                loss.backward()
            side[v, 0] += loss.detach()
            side[v, 1].add_(float(cost_fn(v)) if cost_fn is not None else (r1 - r0) / tile_rows)
            if pose_grad_fn is not None:
                side[v, 2:8] += pose_grad_fn().reshape(6)
        if use_sink:
            b.install(zero=False)
        return self._finish(side, fused, cost_fn, plan)


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False
