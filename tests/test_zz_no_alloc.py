"""GPU: wgs_compact_plan_static (include/wgs_compact.h), the prune plan for hosts that may not allocate.

Some users of this library cannot hand wgs_compact_plan the workspace it wants: a plugin that runs inside another
process's fixed device-memory budget, a real-time mapping thread that must never wait on the allocator (cudaMalloc
and cudaFree can synchronise the whole device), or a caller with no allocator on that GPU at all.  For them the
library keeps the block sums in a static device array of its own (room for 2**24 rows) and the entry point allocates
nothing.  The array is shared by every call on the device and keeps whatever the last call wrote, so this test makes
two different calls back to back -- the second with fewer rows than the first, so stale words sit behind the ones it
uses -- and wants both to agree exactly (integer work: bit for bit, like tests/test_gpu_compact.py) with
wgs_compact_plan on a caller-owned workspace.  Everything goes through the C ABI with ctypes.

This file sorts last on purpose: `pytest -x` stops at the first failure, and a problem in this extra entry point
must not hide the parity tests.  It holds no state of its own, so it passes alone and after the rest of the suite."""
import ctypes as C
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wildgs_slam_blackwell_b200 import _lib, compaction  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _stream():
    return C.c_void_p(torch.cuda.current_stream(DEV).cuda_stream)


def _plan_with_workspace(lib, keep):
    """The existing entry point: the caller allocates the scan workspace."""
    P = keep.numel()
    dst = torch.full((max(P, 1),), -7, dtype=torch.int32, device=DEV)
    count = torch.full((1,), -7, dtype=torch.int32, device=DEV)
    ws = torch.empty(max(int(lib.wgs_compact_workspace_bytes(P)), 256), dtype=torch.uint8, device=DEV)
    rc = lib.wgs_compact_plan(P, keep.data_ptr(), dst.data_ptr(), count.data_ptr(), ws.data_ptr(), _stream())
    assert rc == _lib.WGS_OK, _lib.last_error()
    return dst[:P], count


def _mask(P, frac, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return (torch.rand(P, device=DEV, generator=g) < frac).to(torch.uint8)


def test_static_plan_matches_workspace_plan_on_two_calls_in_a_row():
    lib = compaction._lib_compact()
    with torch.cuda.device(DEV):
        # a big call, then a smaller one that reuses the front of the array, then the edge sizes
        cases = [(1_500_003, 0.9, 1), (70_001, 0.35, 2), (1025, 0.5, 3), (1, 1.0, 4), (0, 0.5, 5), (4096, 0.0, 6),
                 (300_000, 1.0, 7)]
        masks = [_mask(P, frac, seed) for P, frac, seed in cases]
        want = [_plan_with_workspace(lib, k) for k in masks]
        outs = [(torch.full((max(k.numel(), 1),), -9, dtype=torch.int32, device=DEV),
                 torch.full((1,), -9, dtype=torch.int32, device=DEV)) for k in masks]
        torch.cuda.synchronize()
        for i, (k, (dst, count)) in enumerate(zip(masks, outs)):
            if i == 1:      # first call did the per-device set-up; from here on free memory must not move
                torch.cuda.synchronize()
                free_before = torch.cuda.mem_get_info()[0]
                torch_before = torch.cuda.memory_allocated(DEV)
            rc = lib.wgs_compact_plan_static(k.numel(), k.data_ptr(), dst.data_ptr(), count.data_ptr(), _stream())
            assert rc == _lib.WGS_OK, _lib.last_error()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(DEV) == torch_before
        assert torch.cuda.mem_get_info()[0] == free_before, "the static entry point must not allocate device memory"
        for k, (dst, count), (dst_ref, count_ref) in zip(masks, outs, want):
            P = k.numel()
            assert torch.equal(count, count_ref) and int(count) == int(k.sum())
            assert torch.equal(dst[:P], dst_ref)
        # the indices drive the gather exactly like the existing plan's
        k, (dst, count) = masks[1], outs[1]
        rows = torch.randn(k.numel(), 5, device=DEV)
        got = torch.empty(int(count), 5, device=DEV)
        desc = (compaction.CompactTensor * 1)()
        desc[0].src, desc[0].dst, desc[0].row_bytes = rows.data_ptr(), got.data_ptr(), 20
        assert lib.wgs_compact_rows(k.numel(), dst.data_ptr(), desc, 1, _stream()) == _lib.WGS_OK
        assert torch.equal(got, rows[k.bool()])


def test_static_plan_same_mask_twice_gives_the_same_answer():
    lib = compaction._lib_compact()
    with torch.cuda.device(DEV):
        k = _mask(999_999, 0.6, 11)
        dst_ref, count_ref = _plan_with_workspace(lib, k)
        for _ in range(2):
            dst = torch.full((k.numel(),), -9, dtype=torch.int32, device=DEV)
            count = torch.full((1,), -9, dtype=torch.int32, device=DEV)
            assert lib.wgs_compact_plan_static(k.numel(), k.data_ptr(), dst.data_ptr(), count.data_ptr(),
                                               _stream()) == _lib.WGS_OK, _lib.last_error()
            assert torch.equal(dst, dst_ref) and torch.equal(count, count_ref)


def test_static_plan_refuses_sizes_beyond_its_array():
    lib = compaction._lib_compact()
    with torch.cuda.device(DEV):
        small = torch.zeros(64, dtype=torch.int32, device=DEV)       # never dereferenced: the call is refused first
        count = torch.full((1,), 123, dtype=torch.int32, device=DEV)
        rc = lib.wgs_compact_plan_static(compaction.STATIC_MAX_ROWS + 1, small.data_ptr(), small.data_ptr(),
                                         count.data_ptr(), _stream())
        assert rc == _lib.WGS_ERR_CAPACITY
        assert "WGS_COMPACT_STATIC_MAX_ROWS" in _lib.last_error()
        rc = lib.wgs_compact_plan_static(-1, small.data_ptr(), small.data_ptr(), count.data_ptr(), _stream())
        assert rc == _lib.WGS_ERR_INVALID_ARGUMENT
        torch.cuda.synchronize()
        assert int(count) == 123 and int(small.abs().sum()) == 0     # nothing was launched


def test_static_plan_at_its_maximum_size():
    lib = compaction._lib_compact()
    with torch.cuda.device(DEV):
        k = _mask(compaction.STATIC_MAX_ROWS, 0.5, 12)
        dst_ref, count_ref = _plan_with_workspace(lib, k)
        dst = torch.empty(k.numel(), dtype=torch.int32, device=DEV)
        count = torch.empty(1, dtype=torch.int32, device=DEV)
        assert lib.wgs_compact_plan_static(k.numel(), k.data_ptr(), dst.data_ptr(), count.data_ptr(),
                                           _stream()) == _lib.WGS_OK, _lib.last_error()
        assert torch.equal(dst, dst_ref) and torch.equal(count, count_ref)
