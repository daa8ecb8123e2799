"""Gradient bucket in NVLink peer memory + its in-place allreduce kernel.

`PeerBucket(n_floats, device, group)` allocates the flat fp32 gradient bucket of the mapping window in a
symmetric allocation that every rank of the (single-node) group maps (torch.distributed._symmetric_memory
does the handle exchange -- plumbing), and `all_reduce()` sums it over the ranks with ONE kernel of
libwgs_rasterizer.so (csrc/peer_allreduce.cu): NVSwitch multimem reduce + broadcast when the fabric offers a
multicast mapping, peer loads/stores otherwise.  The backward kernels write the per-view gradients straight
into `bucket.flat` (rasterizer.grad_sink), so the step is: backward -> wgs_peer_allreduce, no packing copy,
no NCCL launch.  Results are bit-identical on all ranks.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import _lib


class PeerBucket:
    def __init__(self, n_floats, device, group=None, use_multicast=True):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.device = torch.device(device)
        self.n = int(n_floats)
        self.n_pad = (self.n + 3) // 4 * 4
        lib = _lib.load()
        flag_floats = int(lib.wgs_peer_flags_bytes()) // 4
        self._flag_off = (self.n_pad + 63) // 64 * 64          # floats; 256-B aligned
        total = self._flag_off + flag_floats
        self._t = symm.empty(total, dtype=torch.float32, device=self.device)
        self._h = symm.rendezvous(self._t, self.group.group_name)
        self._t.zero_()
        self.flat = self._t[:self.n]
        bufs = [int(p) for p in self._h.buffer_ptrs]
        self._bufs = (C.c_uint64 * self.world)(*bufs)
        self._flags = (C.c_uint64 * self.world)(*[p + 4 * self._flag_off for p in bufs])
        mc = int(getattr(self._h, "multicast_ptr", 0) or 0)
        self.multicast = bool(mc) and use_multicast
        self._mc = mc if self.multicast else 0
        torch.cuda.synchronize(self.device)
        dist.barrier(self.group)                               # every rank's flags are zero before the first call

    @property
    def nbytes(self):
        return self.n * 4

    def all_reduce(self):
        """In-place sum of `flat` over the group, enqueued on the current stream."""
        with torch.cuda.device(self.device):
            rc = _lib.load().wgs_peer_allreduce(self._bufs, self._flags, self._mc, self.rank, self.world, self.n_pad,
                                                C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
        if rc != _lib.WGS_OK:
            raise RuntimeError(f"wgs_peer_allreduce failed ({rc}): {_lib.last_error()}")
        return self.flat
