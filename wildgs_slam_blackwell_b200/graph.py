"""CUDA-graph capture of an optimisation iteration and no-wait operation of the rasterizer.

The pose-refinement loop (src/mapper.py:885-909: up to 100 iterations of render -> loss -> backward on
fixed shapes) and the mapping loop are launch-bound once the kernels are fast: ~15 launches plus
Python/autograd bookkeeping per iteration, and one host wait per forward (the reference blocks on a
D2H copy of num_rendered, rasterizer_impl.cu:283-284).  `CapturedIteration` records the whole
iteration once (torch.cuda.CUDAGraph; the C ABI detects the capture and skips its host wait) and
replays it with one launch; `async_forward` is the lighter form for eager loops.

Overflow contract: without the host wait nobody can grow the instance buffer mid-flight.  The captured
capacity is 1.5x the instance count seen during warm-up.  A forward that needs more renders the background
only (zero gradients in its backward) and bumps a STICKY per-device record; `check()` (one synchronisation)
raises if any forward of any replay since the previous check overflowed -- recapture then.
"""
import ctypes as C

import torch

from . import _lib


def _index(device):
    """CUDA device index of `device` ('cuda' without an index means the current device)."""
    device = torch.device(device)
    return device.index if device.index is not None else torch.cuda.current_device()


def _context(device):
    idx = _index(device)
    with torch.cuda.device(idx):       # wgs_context_create selects the device it is created on
        return _lib.context(idx)


def last_num_rendered(device):
    """(num_rendered, capacity) of the most recent forward on `device`; the caller has synchronised."""
    n, cap = C.c_int64(0), C.c_int64(0)
    rc = _lib.load().wgs_last_num_rendered(_context(device), C.byref(n), C.byref(cap))
    if rc not in (_lib.WGS_OK, _lib.WGS_ERR_CAPACITY):
        raise RuntimeError(_lib.last_error())
    return int(n.value), int(cap.value)


def overflow_status(device, clear=True):
    """(forwards that exceeded their instance capacity since the record was last cleared, largest instance count
    any of them needed); the caller has synchronised.  The record is sticky (a device word only this call
    clears), so one check covers every forward of every replay since the previous check."""
    cnt, mx = C.c_int64(0), C.c_int64(0)
    rc = _lib.load().wgs_overflow_status(_context(device), C.byref(cnt), C.byref(mx), int(bool(clear)))
    if rc not in (_lib.WGS_OK, _lib.WGS_ERR_CAPACITY):
        raise RuntimeError(_lib.last_error())
    return int(cnt.value), int(mx.value)


def check_overflow(device):
    """Synchronise and raise if ANY forward on `device` since the last check overflowed its instance capacity
    (such a forward rendered the background only and its backward produced zero gradients).  The exception
    carries `.needed`, the largest instance count seen, so the caller can recapture with enough capacity."""
    torch.cuda.synchronize(_index(device))
    cnt, mx = overflow_status(device, clear=True)
    if cnt:
        from . import rasterizer
        rasterizer.raise_capacity_floor(_index(device), int(mx * 1.25) + 1024)
        err = RuntimeError(f"rasterizer instance capacity exceeded by {cnt} forward(s) in no-wait mode (largest "
                           f"needed {mx} instances); those forwards rendered the background only -- rerun them "
                           "outside async/graph mode, or recapture (the capacity hint has been raised)")
        err.needed = mx
        raise err
    return last_num_rendered(device)[0]


class async_forward:
    """Context manager: forwards issued inside do not wait for num_rendered (see wgs_set_async).  Nests: the
    mode that was active on entry is restored on exit."""
    _depth = {}

    def __init__(self, device):
        self.device = device

    def __enter__(self):
        idx = _index(self.device)
        async_forward._depth[idx] = async_forward._depth.get(idx, 0) + 1
        _lib.load().wgs_set_async(_context(self.device), 1)
        return self

    def __exit__(self, *exc):
        idx = _index(self.device)
        async_forward._depth[idx] -= 1
        if async_forward._depth[idx] == 0:
            _lib.load().wgs_set_async(_context(self.device), 0)
        return False


class chained_launch:
    """Context manager around wgs_set_chained_launch: the programmatic dependent launch of the forward / backward
    kernels (on by default) is switched for the launches and graph captures made inside.  It pays on one stream (no gap
    between the kernels of a frame); with a second stream that could use the SM slots the waiting grid holds it costs
    more than it saves, so window.WindowTrainer turns it off around its two-stream section.  Process-wide setting;
    the previous one is restored on exit."""

    def __init__(self, enabled):
        self.enabled = bool(enabled)

    def __enter__(self):
        self._before = _lib.load().wgs_set_chained_launch(int(self.enabled))
        return self

    def __exit__(self, *exc):
        _lib.load().wgs_set_chained_launch(self._before)
        return False


class CapturedIteration:
    """Capture `step()` -- a whole iteration over tensors with fixed storage, backward included -- into a
    CUDA graph.  `step` must set the .grad of its leaves to None before its backward (so the captured
    backward allocates them in the graph's pool); whatever it returns is kept and refreshed by replay().

        it = CapturedIteration(step, "cuda:0")
        for _ in range(100):
            loss = it.replay()          # one launch; read leaf.grad / loss afterwards
            optimizer.step()            # in-place updates of the captured leaves are seen by the next replay
        it.check()                      # covers all 100 replays (sticky overflow record)
    """

    def __init__(self, step, device, warmup=3, pool=None, collect=2):
        """collect: generation handed to gc.collect before the warm-up (see below); a full collection walks every Python
        object of the process (20-35 ms beside a 2 M-row model, measured: profiles/r03d_capture_breakdown.json).  A loop
        that re-captures over FRESH leaf tensors, never stepped on the default stream, has no stale accumulator to drop and
        may pass 0 (young objects only) -- if it was wrong the capture fails loudly, it cannot record a wrong graph.
        pool: a memory-pool handle (`previous.pool()`) to capture into instead of a fresh private pool -- for a loop that
        RE-captures its step (the row count changed) and no longer replays the previous graph: the blocks that graph's
        capture allocated are reused instead of paying cudaMalloc / cudaFree per capture."""
        import gc
        from . import rasterizer
        self.device = torch.device("cuda", _index(device))
        # autograd binds a leaf's gradient accumulator to the stream that was current when it was created and keeps
        # it while any graph references it; drop dead graphs from earlier eager iterations so that the warm-up below
        # (on the side stream) creates fresh ones -- a live graph built on the default stream would make the captured
        # backward launch on that stream and invalidate the capture
        gc.collect(collect)
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):       # eager warm-up: contexts, workspaces, capacity hint
            for _ in range(warmup):
                step()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        rasterizer.take_capture_pins()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, pool=pool):
            self.out = step()
        # the recorded kernels hold raw pointers into the gradient-accumulator workspaces: keep those alive as long
        # as the graph (the cache in rasterizer.py may evict them)
        self._pins = rasterizer.take_capture_pins()

    def pool(self):
        return self.graph.pool()

    def replay(self):
        self.graph.replay()
        return self.out

    def check(self):
        return check_overflow(self.device)
