"""CUDA-graph capture of an optimisation iteration and no-wait operation of the rasterizer.

The pose-refinement loop (src/mapper.py:885-909: up to 100 iterations of render -> loss -> backward on
fixed shapes) and the mapping loop are launch-bound once the kernels are fast: ~15 launches plus
Python/autograd bookkeeping per iteration, and one host wait per forward (the reference blocks on a
D2H copy of num_rendered, rasterizer_impl.cu:283-284).  `CapturedIteration` records the whole
iteration once (torch.cuda.CUDAGraph; the C ABI detects the capture and skips its host wait) and
replays it with one launch; `async_forward` is the lighter form for eager loops.

Overflow contract: without the host wait nobody can grow the instance buffer mid-flight.  The captured
capacity is 1.5x the instance count seen during warm-up; `check()` (one synchronisation) raises if
the most recent forward needed more -- recapture then.
"""
import ctypes as C

import torch

from . import _lib


def last_num_rendered(device):
    """(num_rendered, capacity) of the most recent forward on `device`; the caller has synchronised."""
    device = torch.device(device)
    n, cap = C.c_int64(0), C.c_int64(0)
    rc = _lib.load().wgs_last_num_rendered(_lib.context(device.index), C.byref(n), C.byref(cap))
    if rc not in (_lib.WGS_OK, _lib.WGS_ERR_CAPACITY):
        raise RuntimeError(_lib.last_error())
    return int(n.value), int(cap.value)


def check_overflow(device):
    """Synchronise and raise if the most recent forward overflowed its instance capacity."""
    torch.cuda.synchronize(device)
    n, cap = last_num_rendered(device)
    if n > cap:
        raise RuntimeError(f"rasterizer instance capacity exceeded in no-wait mode: {n} > {cap}; outputs of that "
                           "forward are undefined -- rerun it outside async/graph mode (or recapture)")
    return n


class async_forward:
    """Context manager: forwards issued inside do not wait for num_rendered (see wgs_set_async)."""

    def __init__(self, device):
        self.device = torch.device(device)

    def __enter__(self):
        _lib.load().wgs_set_async(_lib.context(self.device.index), 1)
        return self

    def __exit__(self, *exc):
        _lib.load().wgs_set_async(_lib.context(self.device.index), 0)
        return False


class CapturedIteration:
    """Capture `step()` -- a whole iteration over tensors with fixed storage, backward included -- into a
    CUDA graph.  `step` must set the .grad of its leaves to None before its backward (so the captured
    backward allocates them in the graph's pool); whatever it returns is kept and refreshed by replay().

        it = CapturedIteration(step, "cuda:0")
        for _ in range(100):
            loss = it.replay()          # one launch; read leaf.grad / loss afterwards
            optimizer.step()            # in-place updates of the captured leaves are seen by the next replay
        it.check()
    """

    def __init__(self, step, device, warmup=3):
        import gc
        self.device = torch.device(device)
        # autograd binds a leaf's gradient accumulator to the stream that was current when it was created and keeps
        # it while any graph references it; drop dead graphs from earlier eager iterations so that the warm-up below
        # (on the side stream) creates fresh ones -- a live graph built on the default stream would make the captured
        # backward launch on that stream and invalidate the capture
        gc.collect()
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):       # eager warm-up: contexts, workspaces, capacity hint
            for _ in range(warmup):
                step()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = step()

    def replay(self):
        self.graph.replay()
        return self.out

    def check(self):
        return check_overflow(self.device)
