"""Buffers in NVLink peer memory + the exchange kernel over them.

`PeerMemory(n_floats, device, group)` is a flat fp32 buffer in a symmetric allocation that every rank of the
(single-node) group maps (torch.distributed._symmetric_memory does the handle exchange -- plumbing).
`PeerBucket` adds the flag area of the exchange kernel (csrc/peer_allreduce.cu, C ABI wgs_peer_exchange):
`all_reduce()` sums the bucket over the ranks with ONE kernel of libwgs_rasterizer.so -- NVSwitch multimem reduce +
broadcast when the fabric offers a multicast mapping, peer loads/stores otherwise; `reduce_scatter()` /
`all_gather()` are its two halves for a sharded optimiser step (the summed gradient of 1/world of the model lands
on its owner, Adam runs there, the updated parameters -- a second PeerMemory of the same layout -- return to all
ranks).  All three take a segment list, so a capacity-headroom arena only moves its used rows.  The backward kernels
write the per-view gradients straight into `bucket.flat` (rasterizer.grad_sink): backward -> exchange, no packing
copy, no NCCL launch.  Results are bit-identical on all ranks.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import _lib

ALLREDUCE, REDUCE_SCATTER, ALLGATHER, RENDEZVOUS = 0, 1, 2, 3


class PeerMemory:
    def __init__(self, n_floats, device, group=None, use_multicast=True, extra_floats=0):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.device = torch.device(device)
        self.n = int(n_floats)
        self.n_pad = (self.n + 3) // 4 * 4
        self._extra_off = (self.n_pad + 63) // 64 * 64          # floats; 256-B aligned
        total = self._extra_off + int(extra_floats)
        self._t = symm.empty(total, dtype=torch.float32, device=self.device)
        self._h = symm.rendezvous(self._t, self.group.group_name)
        self._t.zero_()
        self.flat = self._t[:self.n]
        bufs = [int(p) for p in self._h.buffer_ptrs]
        self._buf_ptrs = bufs
        self._bufs = (C.c_uint64 * self.world)(*bufs)
        mc = int(getattr(self._h, "multicast_ptr", 0) or 0)
        self.multicast = bool(mc) and use_multicast
        self._mc = mc if self.multicast else 0
        torch.cuda.synchronize(self.device)
        dist.barrier(self.group)                               # every rank's copy is zero before the first use

    @property
    def nbytes(self):
        return self.n * 4


def configure(blocks=None, unroll=None):
    """Launch shape of the exchange kernel (wgs_peer_configure); every rank must call it with the same values.  Defaults:
    environment WGS_PEER_BLOCKS / WGS_PEER_UNROLL, else (0) the library's measured defaults per path."""
    import os
    b = int(blocks if blocks is not None else os.environ.get("WGS_PEER_BLOCKS", 0) or 0)
    u = int(unroll if unroll is not None else os.environ.get("WGS_PEER_UNROLL", 0) or 0)
    if b or u:
        rc = _lib.load().wgs_peer_configure(b, u)
        if rc != _lib.WGS_OK:
            raise RuntimeError(f"wgs_peer_configure failed ({rc}): {_lib.last_error()}")


def owned_pieces(segments, rank, world):
    """The part of the virtual array (the `segments` = [(offset, length)] in floats laid end to end) rank `rank` owns in a
    reduce-scatter / allgather: a list of (offset, length) pieces of the buffer.  Mirrors the kernel's split (float4
    granularity, ceil(n4 / world) per rank)."""
    n4 = sum(l // 4 for _, l in segments)
    per = (n4 + world - 1) // world
    lo, hi = min(n4, per * rank), min(n4, per * rank + per)
    out, run = [], 0
    for off, length in segments:
        l4 = length // 4
        a, b = max(lo, run), min(hi, run + l4)
        if b > a:
            out.append((off + 4 * (a - run), 4 * (b - a)))
        run += l4
    return out


class PeerBucket(PeerMemory):
    _configured = False

    def __init__(self, n_floats, device, group=None, use_multicast=True):
        flag_floats = int(_lib.load().wgs_peer_flags_bytes()) // 4
        if not PeerBucket._configured:
            PeerBucket._configured = True
            configure()                                        # environment overrides, once per process
        super().__init__(n_floats, device, group, use_multicast, extra_floats=flag_floats)
        self._flags = (C.c_uint64 * self.world)(*[p + 4 * self._extra_off for p in self._buf_ptrs])

    def _exchange(self, mode, segments, target=None):
        t = self if target is None else target
        if segments is None:
            segments = [(0, t.n_pad)]
        if any((o % 4) or (l % 4) for o, l in segments):
            raise ValueError("peer exchange: segment offsets and lengths must be multiples of 4 floats")
        offs = (C.c_uint64 * len(segments))(*[int(o) for o, _ in segments])
        lens = (C.c_uint64 * len(segments))(*[int(l) for _, l in segments])
        with torch.cuda.device(self.device):
            rc = _lib.load().wgs_peer_exchange(t._bufs, self._flags, t._mc, self.rank, self.world, offs, lens, len(segments),
                                               mode, C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
        if rc != _lib.WGS_OK:
            raise RuntimeError(f"wgs_peer_exchange failed ({rc}): {_lib.last_error()}")

    def all_reduce(self, segments=None):
        """In-place sum of `flat` (or of the given [(offset, length)] pieces of it, in floats) over the group, enqueued on
        the current stream."""
        self._exchange(ALLREDUCE, segments)
        return self.flat

    def reduce_scatter(self, segments=None):
        """The slice of the (segmented) data this rank owns -- owned_pieces() -- receives the sum over the group."""
        self._exchange(REDUCE_SCATTER, segments)

    def rendezvous(self):
        """One-block peer barrier on the current stream (no data): see wgs_peer_exchange mode 3."""
        self._exchange(RENDEZVOUS, [(0, 4)])

    def all_gather(self, target=None, segments=None):
        """Every rank's owned slice of `target` (another PeerMemory with the same segment layout; default: this bucket)
        is written into all ranks' copies."""
        self._exchange(ALLGATHER, segments, target)
