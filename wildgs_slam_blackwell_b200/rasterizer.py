"""Host side of the drop-in: same names, argument order, outputs and error behaviour as the
reference's thirdparty/diff-gaussian-rasterization-w-pose/diff_gaussian_rasterization/__init__.py
(:21-46 rasterize_gaussians, :48-171 _RasterizeGaussians, :173-186 settings, :188-244 module),
calling libwgs_rasterizer.so through ctypes instead of the pybind `_C` module.

PyTorch is used for device memory, streams and autograd plumbing only.
"""
import collections
import contextvars
import ctypes as C
import threading
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _lib


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    projmatrix_raw: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def cpu_deep_copy_tuple(input_tuple):
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


# ---- host state: instance-capacity hint per (device, W, H, P) and the self-cleaning gradient workspaces ----
_capacity_hint = collections.OrderedDict()      # (device, W, H, P) -> instances to reserve; least recently used beyond _HINT_KEEP go
_capacity_ratio = {}        # (device, W, H) -> instances per Gaussian of the last forward that read its count back
_HINT_KEEP = 64
_capacity_floor = {}        # device -> instance count a no-wait forward was seen to need (graph.check_overflow)
# (device, stream, P) -> accumulator workspace.  One per STREAM: two backwards in flight on two streams of a device
# never share accumulators / the pose ticket; one per P: a masked render or a densified model next to the full one
# does not evict (and free) a workspace a captured CUDA graph still points at.  Least-recently-used entries beyond
# _GRAD_WS_KEEP per device are dropped unless a live CapturedIteration pinned them.
_grad_ws = collections.OrderedDict()
_GRAD_WS_KEEP = 8
_grad_ws_lock = threading.Lock()
_capture_pins = []          # workspaces handed out while a stream capture was active (graph.CapturedIteration takes them)


def _ptr(t):
    """Device pointer of a tensor; empty tensor == absent == NULL (reference convention)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32c(t, device):
    """float32, contiguous, on `device`; empty stays empty."""
    if t is None:
        return None
    if t.numel() == 0:
        return t
    if t.dtype != torch.float32 or t.device != device:
        t = t.to(device=device, dtype=torch.float32)
    return t.contiguous()


def _empty_u8(nbytes, device):
    return torch.empty(int(nbytes), dtype=torch.uint8, device=device)


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _check(rc, what):
    if rc != _lib.WGS_OK:
        msg = _lib.last_error()
        if rc == _lib.WGS_ERR_NO_COLOR_SOURCE:
            raise RuntimeError("For non-RGB, provide precomputed Gaussian colors! (" + msg + ")")
        raise RuntimeError(f"{what} failed ({rc}): {msg}")


class ForwardState:
    """What one forward leaves behind for its backward / for the parity exports."""
    __slots__ = ("args", "geom", "tile", "binning", "image", "num_rendered", "keep", "sink", "stats")


def forward_raw(means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, s, sh_rest=None,
                activations=False):
    """Runs the forward through the C ABI.  Returns (color, radii, depth, opacity, n_touched, state).
    sh_rest / activations: raw-parameter mode (WgsForwardArgs::activations, ::sh_rest)."""
    if means3D.ndim != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("wildgs_slam_blackwell_b200: tensors must live on a CUDA (sm_100) device; "
                           "there is no CPU path")
    device = means3D.device
    lib = _lib.load()
    with torch.cuda.device(device):
        ctx = _lib.context(device.index if device.index is not None else torch.cuda.current_device())
        P = means3D.shape[0]
        H, W = int(s.image_height), int(s.image_width)
        means3D = _f32c(means3D, device)
        sh = _f32c(sh, device)
        colors_precomp = _f32c(colors_precomp, device)
        opacities = _f32c(opacities, device)
        scales = _f32c(scales, device)
        rotations = _f32c(rotations, device)
        cov3Ds_precomp = _f32c(cov3Ds_precomp, device)
        bg = _f32c(s.bg, device)
        view = _f32c(s.viewmatrix, device)
        proj = _f32c(s.projmatrix, device)
        campos = _f32c(s.campos, device)
        M = int(sh.shape[1]) if (sh is not None and sh.numel() != 0) else 0
        sh_rest = _f32c(sh_rest, device)
        if sh_rest is not None and sh_rest.numel() == 0:
            sh_rest = None                  # degree 0: DC coefficients only
        if sh_rest is not None and M:
            if sh.shape[1] != 1:
                raise RuntimeError("with sh_rest, sh must be the (P,1,3) DC coefficients")
            M = 1 + int(sh_rest.shape[1])

        strip = tile_rows.current()
        new_img = torch.zeros if strip is not None else torch.empty   # outside a strip nothing is written
        color = new_img((3, H, W), dtype=torch.float32, device=device)
        depth = new_img((1, H, W), dtype=torch.float32, device=device)
        opacity = new_img((1, H, W), dtype=torch.float32, device=device)
        radii = torch.empty((P,), dtype=torch.int32, device=device)
        n_touched = torch.empty((P,), dtype=torch.int32, device=device)

        key = (device.index, W, H, P)
        # A densify or prune changes P: the instances-per-Gaussian ratio seen at this resolution carries the learned
        # size over to the new row count, so a no-wait forward at the new P does not start from the 4 P guess.
        carried = int(_capacity_ratio.get(key[:3], 0.0) * P * 1.5)
        cap = max(_capacity_hint.get(key, 0), carried, _capacity_floor.get(device.index, 0), 4 * P, 1 << 16)
        st = ForwardState()
        a = _lib.ForwardArgs()
        a.P, a.sh_degree, a.M, a.width, a.height = P, int(s.sh_degree), M, W, H
        a.tanfovx, a.tanfovy, a.scale_modifier = float(s.tanfovx), float(s.tanfovy), float(s.scale_modifier)
        a.prefiltered, a.debug = int(bool(s.prefiltered)), int(bool(s.debug))
        a.background, a.means3D, a.sh, a.colors_precomp = _ptr(bg), _ptr(means3D), _ptr(sh), _ptr(colors_precomp)
        a.opacities, a.scales, a.rotations, a.cov3D_precomp = _ptr(opacities), _ptr(scales), _ptr(rotations), _ptr(cov3Ds_precomp)
        a.viewmatrix, a.projmatrix, a.campos = _ptr(view), _ptr(proj), _ptr(campos)
        a.out_color, a.out_depth, a.out_opacity = color.data_ptr(), depth.data_ptr(), opacity.data_ptr()
        a.radii, a.n_touched = _ptr(radii), _ptr(n_touched)
        if strip is not None:
            if not (0 <= strip.begin < strip.end <= (H + 15) // 16):
                raise RuntimeError(f"tile_rows({strip.begin}, {strip.end}) outside the {(H + 15) // 16} tile rows")
            a.tile_row_begin, a.tile_row_end = strip.begin, strip.end
        a.activations, a.sh_rest = int(bool(activations)), (_ptr(sh_rest) if M > 1 else None)
        R = C.c_int64(0)
        stream = _stream_ptr(device)
        while True:
            ws = _lib.workspace_sizes(P, W, H, cap)
            st.geom = _empty_u8(ws.geom_bytes, device)
            st.tile = _empty_u8(ws.tile_bytes, device)
            st.binning = _empty_u8(ws.binning_bytes, device)
            st.image = _empty_u8(ws.image_bytes, device)
            a.geom, a.tile, a.binning, a.image = (st.geom.data_ptr(), st.tile.data_ptr(), st.binning.data_ptr(),
                                                  st.image.data_ptr())
            a.capacity = cap
            rc = lib.wgs_forward(ctx, C.byref(a), stream, C.byref(R))
            if rc == _lib.WGS_ERR_CAPACITY:
                cap = int(R.value * 1.25) + 1024
                continue
            _check(rc, "wgs_forward")
            break
        if R.value >= 0:   # (-1: graph capture / async mode, R is not read back)
            old = _capacity_hint.get(key, 0)
            if not (R.value * 1.2 <= old <= R.value * 4):
                _capacity_hint[key] = int(R.value * 1.5)
            if key in _capacity_hint:
                _capacity_hint.move_to_end(key)
                while len(_capacity_hint) > _HINT_KEEP:
                    _capacity_hint.popitem(last=False)
            if P > 0 and strip is None:
                _capacity_ratio[key[:3]] = R.value / P
        st.args = a
        st.num_rendered = int(R.value)
        # what is active around the FORWARD is bound to its backward (the backward runs on autograd's worker thread)
        st.sink, st.stats = grad_sink.current(), densification_stats.current()
        # Keep every INPUT the argument block points at alive as long as the state lives.  The
        # outputs are deliberately not held: under autograd they carry grad_fn -> ctx -> state, and
        # holding them here would close a reference cycle that only the cyclic GC frees (workspaces
        # would pile up and every forward would hit cudaMalloc).
        st.keep = (bg, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, view, proj, campos,
                   sh_rest)
    return color, radii, depth, opacity, n_touched, st


def _grad_workspace(device, P):
    """The accumulator workspace of (device, current stream, P): zeroed once, the kernels re-zero it."""
    stream = torch.cuda.current_stream(device)
    key = (device.index, int(stream.cuda_stream), int(P))
    with _grad_ws_lock:
        ws = _grad_ws.get(key)
        if ws is None:
            nbytes = _lib.workspace_sizes(P, 16, 16, 0).grad_bytes
            ws = torch.zeros(int(nbytes), dtype=torch.uint8, device=device)
            _grad_ws[key] = ws
            mine = [k for k in _grad_ws if k[0] == device.index]
            for k in mine[:max(0, len(mine) - _GRAD_WS_KEEP)]:
                del _grad_ws[k]      # (a pinned tensor stays alive through its CapturedIteration)
        else:
            _grad_ws.move_to_end(key)
        if torch.cuda.is_current_stream_capturing():
            _capture_pins.append(ws)
    return ws


def raise_capacity_floor(device_index, needed):
    """Every later forward on the device starts with room for `needed` instances (no-wait forwards cannot retry)."""
    _capacity_floor[device_index] = max(_capacity_floor.get(device_index, 0), int(needed))


def _drop_grad_workspace(device, P):
    stream = torch.cuda.current_stream(device)
    with _grad_ws_lock:
        _grad_ws.pop((device.index, int(stream.cuda_stream), int(P)), None)


def take_capture_pins():
    """Workspaces whose raw pointers went into kernels recorded during a stream capture since the last call; the
    owner of the graph keeps them so that their lifetime is at least the graph's."""
    with _grad_ws_lock:
        pins = list(_capture_pins)
        del _capture_pins[:]
    return pins


class _Scoped:
    """Context manager whose innermost active instance is found through a contextvars.ContextVar (per thread / per
    asyncio task) -- the forward binds what is active around IT to its autograd node -- and, for callers that wrap
    only `loss.backward()`, through a per-device registry: autograd runs a CUDA node's backward on its own worker
    thread, which never sees the caller's context variables, so the backward-only form cannot be thread-local; it
    is safe with one backward at a time per device, the forward-bound form with any number of threads/streams."""
    _var = None          # ContextVar, one per subclass
    _stack = None        # process-wide stack of active instances, one list per subclass

    def __enter__(self):
        self._token = self._var.set(self)
        self._thread = threading.get_ident()
        self._stack.append(self)
        return self

    def __exit__(self, *exc):
        self._var.reset(self._token)
        self._stack.remove(self)
        return False

    @classmethod
    def current(cls):
        """innermost instance active in the calling context (forward side)"""
        return cls._var.get()

    @classmethod
    def current_for_backward(cls, bound):
        """`bound` (what the forward saw) wins; else the innermost instance entered anywhere in the process -- unless
        instances entered by different threads are active: the worker thread that runs this backward cannot tell whose
        it is, and guessing would hand one thread's buffers to another thread's backward, so that raises."""
        if bound is not None:
            return bound
        stack = list(cls._stack)
        if not stack:
            return None
        if len({inst._thread for inst in stack}) > 1:
            raise RuntimeError(f"{cls.__name__}: instances entered by several threads are active and this backward's "
                               f"forward ran outside all of them; issue the forward inside `with {cls.__name__}(...)` "
                               "(that binds it to its backward), or run one backward-only block at a time")
        return stack[-1]


class tile_rows(_Scoped):
    """Context manager (new capability: sub-frame sharding of a keyframe window): forwards issued
    while active render only tile rows [begin, end) (rows of 16 pixels).  Output images have the full
    (H, W) shape and are ZERO outside the strip, so images, n_touched and -- after backward --
    every gradient of the strips of a partition sum to the whole-image results; radii keep their
    whole-image meaning.  The matching backward uses the same strip automatically.

        with tile_rows(0, 15):  color_top, *_ = rasterizer(...)
    """
    _var = contextvars.ContextVar("wgs_tile_rows", default=None)
    _stack = []

    def __init__(self, begin, end):
        self.begin, self.end = int(begin), int(end)


class grad_sink(_Scoped):
    """Context manager (new capability, used by the multi-GPU mapping window): the backward writes the
    Gaussian-parameter gradients straight into the caller's buffers
    (`buffers`: any of 'means3D', 'sh', 'colors', 'opacity', 'scales', 'rotations', 'cov3D' ->
    float32 CUDA tensor of the gradient's shape, e.g. slices of one flat allreduce bucket) and
    autograd receives None for them; with accumulate=True (or "atomic") the kernel adds to the buffers instead
    of overwriting, so a window of views sums in place.  Pose gradients still flow through autograd.

        with grad_sink(bucket_views, accumulate=(k > 0)):
            loss_k = render_and_loss(k)      # the forward binds the sink to its backward (thread-safe form)
            loss_k.backward()

    Wrapping only `loss.backward()` works too (one backward at a time per process; see _Scoped).
    """
    _var = contextvars.ContextVar("wgs_grad_sink", default=None)
    _stack = []

    def __init__(self, buffers, accumulate=False):
        """accumulate: False (overwrite), True (add; one backward at a time per buffer) or "atomic" (add with reduction
        atomics: backwards running concurrently on several streams may share the buffers; zero them first)"""
        self.buffers = dict(buffers)
        self.accumulate = 2 if accumulate == "atomic" else int(bool(accumulate))


class densification_stats(_Scoped):
    """Context manager: every backward of a forward issued while active (or every backward run while active) also does
    what the mapper runs after it on the same view
    (src/mapper.py:1176-1182, GaussianModel.add_densification_stats): for the Gaussians with radii > 0,
    `xyz_gradient_accum += ||dL_dmeans2D[:, :2]||`, `denom += 1`, `max_radii2D = max(max_radii2D, radii)` -- inside
    the per-Gaussian backward kernel, so the window needs neither the (P,3) dL_dmeans2D tensor nor the ~12 indexing
    kernels per view.  Tensors: float32 CUDA, P elements each (any may be None).  Whole-image backward only."""
    _var = contextvars.ContextVar("wgs_densification_stats", default=None)
    _stack = []

    def __init__(self, xyz_gradient_accum=None, denom=None, max_radii2D=None):
        self.bufs = (xyz_gradient_accum, denom, max_radii2D)


def backward_raw(st, s, means3D, sh, colors_precomp, scales, rotations, cov3Ds_precomp, radii,
                 grad_out_color, grad_out_depth, sink=None, stats=None):
    """Runs the backward through the C ABI.  Returns dict of gradient tensors (+ 'tau' (6,))."""
    device = means3D.device
    lib = _lib.load()
    with torch.cuda.device(device):
        ctx = _lib.context(device.index if device.index is not None else torch.cuda.current_device())
        P = means3D.shape[0]
        H, W = int(s.image_height), int(s.image_width)
        fa = st.args
        M = fa.M
        use_sh = M > 0 and not fa.colors_precomp
        grad_out_color = _f32c(grad_out_color, device)
        grad_out_depth = _f32c(grad_out_depth, device)
        proj_raw = _f32c(s.projmatrix_raw, device)
        def e(name, *shape):
            if sink is not None and name in sink.buffers:
                t = sink.buffers[name]
                if (tuple(t.shape) != shape or t.dtype != torch.float32 or t.device != device or
                        not t.is_contiguous()):
                    raise RuntimeError(f"grad_sink buffer '{name}' must be a contiguous float32 {shape} tensor on {device}")
                return t
            if sink is not None and sink.accumulate and name not in ("means2D", "tau"):
                raise RuntimeError(f"grad_sink(accumulate=True) needs a buffer for '{name}'")
            return torch.empty(shape, dtype=torch.float32, device=device)
        g = dict(means3D=e("means3D", P, 3), means2D=None, opacity=e("opacity", P, 1), tau=e("tau", 6))
        # dL_dmeans2D only feeds densification statistics; in accumulate mode it is skipped (null)
        if sink is None or not sink.accumulate or "means2D" in sink.buffers:
            g["means2D"] = e("means2D", P, 3)
        split = bool(fa.sh_rest)
        g["sh"] = e("sh", P, 1 if split else M, 3) if use_sh else None
        g["sh_rest"] = e("sh_rest", P, M - 1, 3) if (use_sh and split) else None
        g["colors"] = e("colors", P, 3) if fa.colors_precomp else None
        g["scales"] = e("scales", P, 3) if not fa.cov3D_precomp else None
        g["rotations"] = e("rotations", P, 4) if not fa.cov3D_precomp else None
        g["cov3D"] = e("cov3D", P, 6) if fa.cov3D_precomp else None
        b = _lib.BackwardArgs()
        b.P, b.sh_degree, b.M, b.width, b.height = P, fa.sh_degree, M, W, H
        b.tanfovx, b.tanfovy, b.scale_modifier, b.debug = fa.tanfovx, fa.tanfovy, fa.scale_modifier, fa.debug
        b.accumulate = sink.accumulate if sink is not None else 0
        b.background, b.means3D, b.sh, b.colors_precomp = fa.background, fa.means3D, fa.sh, fa.colors_precomp
        b.scales, b.rotations, b.cov3D_precomp = fa.scales, fa.rotations, fa.cov3D_precomp
        b.viewmatrix, b.projmatrix, b.projmatrix_raw, b.campos = fa.viewmatrix, fa.projmatrix, _ptr(proj_raw), fa.campos
        b.radii = _ptr(radii)
        b.dL_dout_color, b.dL_dout_depth = _ptr(grad_out_color), _ptr(grad_out_depth)
        b.dL_dmeans3D, b.dL_dmeans2D, b.dL_dopacity = _ptr(g["means3D"]), _ptr(g["means2D"]), _ptr(g["opacity"])
        b.dL_dsh, b.dL_dcolors = _ptr(g["sh"]), _ptr(g["colors"])
        b.dL_dscales, b.dL_drotations, b.dL_dcov3D = _ptr(g["scales"]), _ptr(g["rotations"]), _ptr(g["cov3D"])
        b.dL_dtau = _ptr(g["tau"])
        b.geom, b.tile, b.binning, b.image = fa.geom, fa.tile, fa.binning, fa.image
        b.capacity = fa.capacity
        b.tile_row_begin, b.tile_row_end = fa.tile_row_begin, fa.tile_row_end
        b.activations, b.sh_rest, b.dL_dsh_rest = fa.activations, fa.sh_rest, _ptr(g["sh_rest"])
        stats = densification_stats.current_for_backward(stats)
        if stats is not None:
            for t in stats.bufs:
                if t is not None and (t.numel() != P or t.dtype != torch.float32 or t.device != device or not t.is_contiguous()):
                    raise RuntimeError(f"densification_stats buffers must be contiguous float32 tensors of {P} elements on {device}")
            b.stat_grad_accum, b.stat_denom, b.stat_max_radii = (_ptr(t) for t in stats.bufs)
        if P > 0:
            b.grad = _grad_workspace(device, P).data_ptr()
        rc = lib.wgs_backward(ctx, C.byref(b), _stream_ptr(device))
        if rc != _lib.WGS_OK:
            _drop_grad_workspace(device, P)  # accumulators may be dirty
        _check(rc, "wgs_backward")
    return g


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        theta, rho, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, theta, rho, raster_settings)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho,
                raster_settings):
        args = (means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, raster_settings)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args[:-1] + tuple(raster_settings))
            try:
                color, radii, depth, opacity, n_touched, st = forward_raw(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            color, radii, depth, opacity, n_touched, st = forward_raw(*args)
        ctx.raster_settings = raster_settings
        # radii / opacity / n_touched cotangents are never used (reference ignores them too):
        # do not let autograd zero-fill them
        ctx.set_materialize_grads(False)
        ctx.pose_shapes = (theta.shape, rho.shape)
        ctx.num_rendered = st.num_rendered
        ctx.state = st
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh)
        ctx.mark_non_differentiable(radii, n_touched)
        return color, radii, depth, opacity, n_touched

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_radii, grad_out_depth, grad_out_opacity, grad_n_touched):
        # grad_out_opacity is ignored exactly like the reference (__init__.py:108-138)
        s = ctx.raster_settings
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh = ctx.saved_tensors
        if grad_out_color is None:
            grad_out_color = torch.zeros((3, s.image_height, s.image_width), device=means3D.device)
        if grad_out_depth is None:
            grad_out_depth = torch.zeros((1, s.image_height, s.image_width), device=means3D.device)
        sink = grad_sink.current_for_backward(ctx.state.sink)
        args = (ctx.state, s, means3D, sh, colors_precomp, scales, rotations, cov3Ds_precomp, radii, grad_out_color,
                grad_out_depth, sink, ctx.state.stats)
        if s.debug:
            cpu_args = cpu_deep_copy_tuple(args[2:-2] + tuple(s))
            try:
                g = backward_raw(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            g = backward_raw(*args)
        # The reference returns (1,3) for both and lets autograd sum_to_size them onto the leaves;
        # hand autograd the leaf's own shape when it has 3 elements (same .grad, no reduce kernel).
        th_shape, rho_shape = ctx.pose_shapes
        grad_rho = g["tau"][:3].view(rho_shape if len(rho_shape) and rho_shape.numel() == 3 else (1, 3))
        grad_theta = g["tau"][3:].view(th_shape if len(th_shape) and th_shape.numel() == 3 else (1, 3))
        if sink is not None:   # sunk gradients do not go through autograd
            for k in sink.buffers:
                g[k] = None
        return (g["means3D"], g["means2D"], g["sh"], g["colors"], g["opacity"], g["scales"], g["rotations"],
                g["cov3D"], grad_theta, grad_rho, None)


class _RasterizeGaussianModel(torch.autograd.Function):
    """Raw-parameter mode: takes GaussianModel's leaf tensors (_xyz, _features_dc,
    _features_rest, _opacity, _scaling, _rotation) and applies sigmoid / exp / normalize and the dc|rest
    concatenation inside the per-Gaussian kernels, forward and backward.  Replaces the torch kernels the
    reference's render() launches around every rasterizer call
    (thirdparty/gaussian_splatting/gaussian_renderer/__init__.py:77-110 via gaussian_model.py:77-97)."""

    @staticmethod
    def forward(ctx, xyz, means2D, f_dc, f_rest, opacity_raw, scaling_raw, rotation_raw, theta, rho, raster_settings):
        empty = torch.Tensor([])
        color, radii, depth, opacity, n_touched, st = forward_raw(xyz, f_dc, empty, opacity_raw, scaling_raw,
                                                                  rotation_raw, empty, raster_settings,
                                                                  sh_rest=f_rest, activations=True)
        ctx.raster_settings = raster_settings
        ctx.set_materialize_grads(False)
        ctx.pose_shapes = (theta.shape, rho.shape)
        ctx.state = st
        ctx.save_for_backward(xyz, scaling_raw, rotation_raw, radii, f_dc)
        ctx.mark_non_differentiable(radii, n_touched)
        return color, radii, depth, opacity, n_touched

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_radii, grad_out_depth, grad_out_opacity, grad_n_touched):
        s = ctx.raster_settings
        xyz, scaling_raw, rotation_raw, radii, f_dc = ctx.saved_tensors
        if grad_out_color is None:
            grad_out_color = torch.zeros((3, s.image_height, s.image_width), device=xyz.device)
        if grad_out_depth is None:
            grad_out_depth = torch.zeros((1, s.image_height, s.image_width), device=xyz.device)
        sink = grad_sink.current_for_backward(ctx.state.sink)
        empty = torch.Tensor([])
        g = backward_raw(ctx.state, s, xyz, f_dc, empty, scaling_raw, rotation_raw, empty, radii, grad_out_color,
                         grad_out_depth, sink, ctx.state.stats)
        th_shape, rho_shape = ctx.pose_shapes
        grad_rho = g["tau"][:3].view(rho_shape if len(rho_shape) and rho_shape.numel() == 3 else (1, 3))
        grad_theta = g["tau"][3:].view(th_shape if len(th_shape) and th_shape.numel() == 3 else (1, 3))
        if sink is not None:
            for k in sink.buffers:
                g[k] = None
        return (g["means3D"], g["means2D"], g["sh"], g["sh_rest"], g["opacity"], g["scales"], g["rotations"],
                grad_theta, grad_rho, None)


def rasterize_gaussian_model(xyz, means2D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw,
                             theta, rho, raster_settings):
    """Same outputs as rasterize_gaussians(get_xyz, means2D, get_features, [], get_opacity, get_scaling,
    get_rotation, [], theta, rho, settings) of the reference, from the RAW GaussianModel parameters."""
    if theta is None:
        theta = torch.Tensor([])
    if rho is None:
        rho = torch.Tensor([])
    if features_rest is None:
        features_rest = torch.Tensor([])
    return _RasterizeGaussianModel.apply(xyz, means2D, features_dc, features_rest, opacity_raw, scaling_raw,
                                         rotation_raw, theta, rho, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        # reference: __init__.py:193-202 -> _C.mark_visible (rasterize_points.cu:213-232)
        with torch.no_grad():
            s = self.raster_settings
            if not positions.is_cuda:
                raise RuntimeError("wildgs_slam_blackwell_b200: positions must be a CUDA tensor")
            device = positions.device
            pos = _f32c(positions, device)
            view, proj = _f32c(s.viewmatrix, device), _f32c(s.projmatrix, device)
            P = pos.shape[0]
            visible = torch.zeros((P,), dtype=torch.bool, device=device)
            if P:
                with torch.cuda.device(device):
                    rc = _lib.load().wgs_mark_visible(P, pos.data_ptr(), view.data_ptr(), proj.data_ptr(),
                                                      visible.data_ptr(), _stream_ptr(device))
                _check(rc, "wgs_mark_visible")
        return visible

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, theta=None, rho=None):
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])
        if theta is None:
            theta = torch.Tensor([])
        if rho is None:
            rho = torch.Tensor([])
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   theta, rho, raster_settings)
