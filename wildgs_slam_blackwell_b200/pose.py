"""Fused pose step (include/wgs_pose.h): host-side mirror of the reference's
update_pose (src/utils/pose_utils.py:81-100, same name / arguments / return) and of the
`pose_optimizer.step(); update_pose(viewpoint)` pair of the pose-refinement loop (src/mapper.py:861-907) as
`FusedPoseOptimizer`.  One single-warp launch each; R, T and the camera matrices render() uses are updated in
place, the Adam step counter lives on the device, so a whole tracking iteration
(render -> loss -> backward -> pose step) can be one CUDA graph (graph.CapturedIteration)."""
import ctypes as C

import torch

from . import _lib
from . import renderer

_fp = C.c_void_p
EXPORTS = ("wgs_pose_step",)


class PoseAdam(C.Structure):
    _fields_ = [("exp_avg", _fp), ("exp_avg_sq", _fp), ("step", _fp), ("grad_rot", _fp), ("grad_trans", _fp),
                ("grad_exposure_a", _fp), ("grad_exposure_b", _fp),
                ("lr_rot", C.c_double), ("lr_trans", C.c_double), ("lr_exposure_a", C.c_double),
                ("lr_exposure_b", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double)]


class PoseStepArgs(C.Structure):
    _fields_ = [("cam_rot_delta", _fp), ("cam_trans_delta", _fp), ("exposure_a", _fp), ("exposure_b", _fp),
                ("R", _fp), ("T", _fp), ("converged_threshold", C.c_float), ("converged", _fp),
                ("adam", C.POINTER(PoseAdam)), ("projection_matrix", _fp), ("viewmatrix", _fp), ("projmatrix", _fp),
                ("campos", _fp)]


_proto_done = False


def _lib_pose():
    global _proto_done
    lib = _lib.load()
    if not _proto_done:
        lib.wgs_pose_step.argtypes = [C.POINTER(PoseStepArgs), C.c_void_p]
        _proto_done = True
    return lib


def _inplace(t, name, n):
    if not (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == n):
        raise RuntimeError(f"fused pose step: camera.{name} must be a contiguous float32 CUDA tensor of {n} elements "
                           "(it is updated in place)")
    return t


def _bump(*tensors):
    inc = getattr(torch.autograd.graph, "increment_version", None)
    for t in tensors:
        if inc is not None:
            inc(t)


def _launch(camera, threshold, adam, converged):
    R, T = _inplace(camera.R, "R", 9), _inplace(camera.T, "T", 3)
    rot = _inplace(camera.cam_rot_delta.data, "cam_rot_delta", 3)
    trans = _inplace(camera.cam_trans_delta.data, "cam_trans_delta", 3)
    device = R.device
    a = PoseStepArgs()
    a.cam_rot_delta, a.cam_trans_delta, a.R, a.T = rot.data_ptr(), trans.data_ptr(), R.data_ptr(), T.data_ptr()
    a.converged_threshold, a.converged = float(threshold), converged.data_ptr()
    if adam is not None:
        a.adam = C.pointer(adam)
        ea, eb = getattr(camera, "exposure_a", None), getattr(camera, "exposure_b", None)
        a.exposure_a = ea.data.data_ptr() if ea is not None else None
        a.exposure_b = eb.data.data_ptr() if eb is not None else None
    mats = None
    Pt = getattr(camera, "projection_matrix", None)
    if torch.is_tensor(Pt) and Pt.is_cuda and Pt.dtype == torch.float32 and Pt.is_contiguous():
        ent = getattr(camera, "_wgs_matrices", None)
        mats = ent[1] if ent is not None else renderer.camera_matrices(R, T, Pt)   # tensors render() reads
        a.projection_matrix = Pt.data_ptr()
        a.viewmatrix, a.projmatrix, a.campos = mats[0].data_ptr(), mats[1].data_ptr(), mats[2].data_ptr()
    with torch.cuda.device(device):
        rc = _lib_pose().wgs_pose_step(C.byref(a), C.c_void_p(torch.cuda.current_stream(device).cuda_stream))
    if rc != _lib.WGS_OK:
        raise RuntimeError(f"wgs_pose_step failed ({rc}): {_lib.last_error()}")
    _bump(R, T, rot, trans)
    if mats is not None:   # the cached matrices are current for the new R/T: re-key the cache instead of recomputing
        try:
            camera._wgs_matrices = ((id(R), R._version, id(T), T._version, id(Pt)), mats, (R, T, Pt))
        except AttributeError:
            pass


def update_pose(camera, converged_threshold=1e-4):
    """Drop-in for pose_utils.update_pose: returns a 0-dim bool tensor on the device (`if converged:` synchronises,
    exactly like the reference's `tau.norm() < threshold`)."""
    flag = torch.empty(1, dtype=torch.int32, device=camera.R.device)
    _launch(camera, converged_threshold, None, flag)
    return flag[0] != 0


class FusedPoseOptimizer:
    """`pose_optimizer = torch.optim.Adam([rot, trans, exposure_a, exposure_b groups]); ...; pose_optimizer.step();
    converged = update_pose(viewpoint)` as one launch.  `step()` returns nothing; `converged` is an int32 device
    tensor refreshed by every step (read it every few iterations: reading synchronises)."""

    def __init__(self, camera, lr_rot, lr_trans, lr_exposure_a=0.01, lr_exposure_b=0.01, betas=(0.9, 0.999), eps=1e-8,
                 converged_threshold=1e-4):
        self.camera, self.threshold = camera, converged_threshold
        dev = camera.R.device
        self.state = torch.zeros(16, dtype=torch.float32, device=dev)     # exp_avg (8) | exp_avg_sq (8)
        self.step_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self.converged = torch.zeros(1, dtype=torch.int32, device=dev)
        self.hyper = dict(lr_rot=lr_rot, lr_trans=lr_trans, lr_exposure_a=lr_exposure_a, lr_exposure_b=lr_exposure_b,
                          beta1=betas[0], beta2=betas[1], eps=eps)

    def reset(self, R=None, T=None):
        """Start a new refinement with the same (possibly graph-captured) buffers: Adam state and step counter
        zeroed, optionally a new initial pose copied INTO camera.R / camera.T, deltas zeroed, camera matrices
        refreshed in place.  Lets one captured iteration serve every frame of a sequence."""
        cam = self.camera
        with torch.no_grad():
            self.state.zero_()
            self.step_count.zero_()
            if R is not None:
                cam.R.copy_(R)
            if T is not None:
                cam.T.copy_(T)
            cam.cam_rot_delta.zero_()
            cam.cam_trans_delta.zero_()
        update_pose(cam, self.threshold)

    def zero_grad(self):
        cam = self.camera
        for t in (cam.cam_rot_delta, cam.cam_trans_delta, getattr(cam, "exposure_a", None), getattr(cam, "exposure_b", None)):
            if t is not None:
                t.grad = None

    def step(self):
        cam = self.camera
        g = lambda t: None if (t is None or t.grad is None) else t.grad
        gr, gt = g(cam.cam_rot_delta), g(cam.cam_trans_delta)
        if gr is None or gt is None:
            raise RuntimeError("FusedPoseOptimizer.step: cam_rot_delta / cam_trans_delta have no gradient")
        ga, gb = g(getattr(cam, "exposure_a", None)), g(getattr(cam, "exposure_b", None))
        ad = PoseAdam()
        ad.exp_avg, ad.exp_avg_sq, ad.step = self.state.data_ptr(), self.state.data_ptr() + 32, self.step_count.data_ptr()
        self._keep = tuple(x.contiguous() if x is not None else None for x in (gr, gt, ga, gb))
        ad.grad_rot, ad.grad_trans = self._keep[0].data_ptr(), self._keep[1].data_ptr()
        ad.grad_exposure_a = self._keep[2].data_ptr() if ga is not None else None
        ad.grad_exposure_b = self._keep[3].data_ptr() if gb is not None else None
        for k, v in self.hyper.items():
            setattr(ad, k, float(v))
        _launch(cam, self.threshold, ad, self.converged)
