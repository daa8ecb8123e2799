"""CPU (no GPU): the C-ABI library builds/loads and exports every symbol include/wgs_rasterizer.h declares;
the ctypes structs mirror the header; host-side validation behaves like the reference's front-end
(RAST/diff_gaussian_rasterization/__init__.py:204-228); there is no CPU fallback."""
import ctypes as C
import os
import re
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wildgs_slam_blackwell_b200 import _lib, rasterizer  # noqa: E402

HEADER = open(os.path.join(ROOT, "include", "wgs_rasterizer.h")).read()


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    declared = set(re.findall(r"\b(wgs_[a-z_]+)\s*\(", HEADER))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    m = re.search(r"#define WGS_ABI_VERSION (\d+)", HEADER)
    assert lib.wgs_abi_version() == int(m.group(1)) == _lib.ABI_VERSION


def _struct_fields(name):
    body = re.search(r"typedef struct " + name + r" \{(.*?)\} " + name + ";", HEADER, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):
            names.append(re.findall(r"[A-Za-z_0-9]+", part)[-1])
    return names


@pytest.mark.parametrize("cname,cls", [("WgsForwardArgs", _lib.ForwardArgs), ("WgsBackwardArgs", _lib.BackwardArgs),
                                       ("WgsWorkspaceSizes", _lib.WorkspaceSizes)])
def test_ctypes_structs_mirror_header(cname, cls):
    assert _struct_fields(cname) == [f[0] for f in cls._fields_]


def test_workspace_sizes_and_argument_errors_without_gpu():
    ws = _lib.workspace_sizes(300_000, 640, 480, 1_200_000)
    assert ws.geom_bytes >= 300_000 * 57 and ws.binning_bytes >= 1_200_000 * 8
    assert ws.image_bytes >= 640 * 480 * 8 and ws.grad_bytes >= 300_000 * 48 and ws.tile_bytes >= 1200 * 20
    out = _lib.WorkspaceSizes()
    assert _lib.load().wgs_workspace_sizes(-1, 640, 480, 0, C.byref(out)) == _lib.WGS_ERR_INVALID_ARGUMENT
    assert b"invalid" in _lib.load().wgs_last_error()
    assert _lib.load().wgs_forward(None, None, None, None) == _lib.WGS_ERR_INVALID_ARGUMENT
    assert _lib.load().wgs_backward(None, None, None) == _lib.WGS_ERR_INVALID_ARGUMENT
    if not torch.cuda.is_available():
        assert not _lib.load().wgs_context_create(0)          # no device -> null + message, never a fallback
        assert _lib.last_error()


def _settings():
    z = torch.zeros(4, 4)
    return rasterizer.GaussianRasterizationSettings(16, 16, 1.0, 1.0, torch.zeros(3), 1.0, z, z, z, 0, torch.zeros(3),
                                                    False, False)


def test_settings_field_order_matches_reference():
    assert rasterizer.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "projmatrix_raw", "sh_degree", "campos", "prefiltered", "debug")


def test_front_end_validation_messages():
    r = rasterizer.GaussianRasterizer(_settings())
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="Please provide excatly one of either SHs or precomputed colors!"):
        r(m, m, torch.zeros(4, 1), scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="Please provide excatly one"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=m, rotations=torch.zeros(4, 4),
          cov3D_precomp=torch.zeros(4, 6))


def test_no_cpu_fallback():
    r = rasterizer.GaussianRasterizer(_settings())
    m = torch.zeros(4, 3)
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(RuntimeError, match="must have dimensions"):
        rasterizer.forward_raw(torch.zeros(4, 2), None, None, None, None, None, None, _settings())
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        r.markVisible(m)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "wildgs_slam_blackwell_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("the oracle", ""), f
    assert "oracle" not in open(os.path.join(ROOT, "diff_gaussian_rasterization", "__init__.py")).read()


def _integration_block(marker):
    """The fenced python block of INTEGRATION.md that starts with `# --- <marker>`."""
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    for block in re.findall(r"```python\n(.*?)```", doc, re.S):
        if block.startswith("# --- " + marker):
            return block
    raise AssertionError(f"INTEGRATION.md has no '{marker}' block")


def test_integration_md_structs_match_the_library():
    """The reference-side binding INTEGRATION.md section 3 prints is the layout the library reads: same field names,
    offsets, sizes as the package's own ctypes mirror (which test_ctypes_structs_mirror_header ties to the header)."""
    ns = {}
    exec(_integration_block("wgs-abi-structs"), ns)
    for doc_name, cls in (("WgsForwardArgs", _lib.ForwardArgs), ("WgsBackwardArgs", _lib.BackwardArgs),
                          ("WgsWorkspaceSizes", _lib.WorkspaceSizes)):
        doc_cls = ns[doc_name]
        assert C.sizeof(doc_cls) == C.sizeof(cls), doc_name
        assert [f[0] for f in doc_cls._fields_] == [f[0] for f in cls._fields_], doc_name
        for name, _ in cls._fields_:
            assert getattr(doc_cls, name).offset == getattr(cls, name).offset, (doc_name, name)
            assert getattr(doc_cls, name).size == getattr(cls, name).size, (doc_name, name)
    compile(_integration_block("wgs-abi-calls"), "INTEGRATION.md", "exec")      # at least syntactically valid here


def test_stale_struct_layout_is_rejected():
    """ABI v10: a binding built for an older, shorter WgsForwardArgs is refused instead of read past its end."""
    lib = _lib.load()
    a = _lib.ForwardArgs()
    assert a.struct_size == C.sizeof(_lib.ForwardArgs)
    a.struct_size -= 24                                   # the ABI-v8 layout INTEGRATION.md used to print
    n = C.c_int64(0)
    fake_ctx = C.c_void_p(1)                              # never dereferenced: the size check comes first
    assert lib.wgs_forward(fake_ctx, C.byref(a), None, C.byref(n)) == _lib.WGS_ERR_INVALID_ARGUMENT
    assert b"struct_size" in lib.wgs_last_error()
    b = _lib.BackwardArgs()
    b.struct_size = 8
    assert lib.wgs_backward(fake_ctx, C.byref(b), None) == _lib.WGS_ERR_INVALID_ARGUMENT
    assert b"struct_size" in lib.wgs_last_error()


def test_scoped_contexts_are_per_thread_on_the_forward_side():
    """tile_rows / grad_sink / densification_stats: what a forward sees is per thread (contextvars); a backward that
    was not bound by its forward falls back to the innermost instance of the process."""
    import threading
    seen = {}

    def worker(name, rows):
        with rasterizer.tile_rows(*rows):
            barrier.wait()
            seen[name] = (rasterizer.tile_rows.current().begin, rasterizer.tile_rows.current().end)
            barrier.wait()

    barrier = threading.Barrier(2)
    ts = [threading.Thread(target=worker, args=("a", (0, 3))), threading.Thread(target=worker, args=("b", (5, 9)))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert seen == {"a": (0, 3), "b": (5, 9)}
    assert rasterizer.tile_rows.current() is None
    sink = rasterizer.grad_sink({})
    assert rasterizer.grad_sink.current_for_backward(None) is None
    with sink:
        assert rasterizer.grad_sink.current() is sink
        assert rasterizer.grad_sink.current_for_backward(None) is sink      # backward-only wrapping
        other = rasterizer.grad_sink({})
        assert rasterizer.grad_sink.current_for_backward(other) is other    # the forward's binding wins
    assert rasterizer.grad_sink.current_for_backward(None) is None


def test_backward_only_scope_refuses_to_guess_between_threads():
    """Two threads each hold a grad_sink and a backward arrives whose forward ran outside both: the autograd worker
    cannot know whose it is.  It used to take the most recently entered one (silently another thread's buffers); it
    raises now.  Nesting on ONE thread stays legal (innermost wins), and a forward-bound instance is never ambiguous."""
    import threading
    errors, ready, done = [], threading.Barrier(3), threading.Event()

    def holder():
        with rasterizer.grad_sink({}):
            ready.wait()
            done.wait()

    ts = [threading.Thread(target=holder) for _ in range(2)]
    [t.start() for t in ts]
    ready.wait()
    try:
        with pytest.raises(RuntimeError, match="several threads"):
            rasterizer.grad_sink.current_for_backward(None)
        bound = rasterizer.grad_sink({})
        assert rasterizer.grad_sink.current_for_backward(bound) is bound
    finally:
        done.set()
        [t.join() for t in ts]
    with rasterizer.grad_sink({}) as outer, rasterizer.grad_sink({}) as inner:
        assert rasterizer.grad_sink.current_for_backward(None) is inner
    assert rasterizer.grad_sink.current_for_backward(None) is None and not errors
