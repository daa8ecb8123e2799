"""render(): host-side mirror of the reference's
thirdparty/gaussian_splatting/gaussian_renderer/__init__.py:24-153 (same name, arguments, returned
dict), the one caller of the rasterizer in mapping and tracking (src/mapper.py:887-889, 1106, 1277).
SURVEY.md §8(f) row 1.

What differs from the reference wrapper, not in results but in launches per call:
  * GaussianModel's activations (sigmoid / exp / normalize) and the dc|rest concatenation of
    get_features run inside the rasterizer's per-Gaussian kernels (raw-parameter mode of the C ABI)
    instead of as 4 forward + ~8 backward torch kernels and a (P,M,3) copy;
  * the three camera matrices come from one 16-thread kernel and are cached on the camera until its
    pose tensors change, instead of ~15 tiny kernels (two linalg.inv, bmm, inverse) per call.
Paths the fused kernels do not cover (pipe.compute_cov3D_python, pipe.convert_SHs_python,
override_color, isotropic scaling, a Gaussian mask) run the reference's own composition on top of
the drop-in GaussianRasterizer.
"""
import ctypes as C
import math

import torch

from . import _lib
from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussian_model

_RAW = ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation")


def camera_matrices(R, T, projection_matrix):
    """(viewmatrix, projmatrix, campos) of a W2C pose (R (3,3), T (3,)) and the stored projection_matrix = P^T:
    what Camera.world_view_transform / full_proj_transform / camera_center return (camera_utils.py:137-151)."""
    device = projection_matrix.device
    if device.type != "cuda":
        raise RuntimeError("camera_matrices runs on a CUDA (sm_100) device only")
    f = lambda t: t.detach().to(device=device, dtype=torch.float32).contiguous()
    R, T, Pt = f(R), f(T), f(projection_matrix)
    out = torch.empty(35, dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        rc = _lib.load().wgs_camera_matrices(R.data_ptr(), T.data_ptr(), Pt.data_ptr(), out.data_ptr(),
                                             out.data_ptr() + 64, out.data_ptr() + 128,
                                             C.c_void_p(torch.cuda.current_stream(device).cuda_stream))
    if rc != _lib.WGS_OK:
        raise RuntimeError(f"wgs_camera_matrices failed ({rc}): {_lib.last_error()}")
    return out[:16].view(4, 4), out[16:32].view(4, 4), out[32:35]


def _cached_matrices(cam):
    """Matrices of `cam`, recomputed only when its pose tensors were replaced or written in place."""
    R, T, Pt = cam.R, cam.T, cam.projection_matrix
    key = (id(R), R._version, id(T), T._version, id(Pt))
    ent = getattr(cam, "_wgs_matrices", None)
    if ent is None or ent[0] != key:
        ent = (key, camera_matrices(R, T, Pt), (R, T, Pt))
        try:
            cam._wgs_matrices = ent
        except AttributeError:
            pass
    return ent[1]


def _camera(cam):
    if all(hasattr(cam, a) for a in ("R", "T", "projection_matrix")) and torch.is_tensor(cam.R) and \
            cam.projection_matrix.is_cuda:
        return _cached_matrices(cam)
    return cam.world_view_transform, cam.full_proj_transform, cam.camera_center


def render(viewpoint_camera, pc, pipe, bg_color, scaling_modifier=1.0, override_color=None, mask=None):
    """Render the scene.  Background tensor (bg_color) must be on GPU!"""
    if pc.get_xyz.shape[0] == 0:
        return None
    xyz = pc.get_xyz
    screenspace_points = torch.zeros_like(xyz, requires_grad=True)  # gradient sink for the 2D means
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    viewmatrix, projmatrix, campos = _camera(viewpoint_camera)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=tanfovx, tanfovy=tanfovy, bg=bg_color, scale_modifier=scaling_modifier, viewmatrix=viewmatrix,
        projmatrix=projmatrix, projmatrix_raw=viewpoint_camera.projection_matrix, sh_degree=pc.active_sh_degree,
        campos=campos, prefiltered=False, debug=False)
    theta, rho = viewpoint_camera.cam_rot_delta, viewpoint_camera.cam_trans_delta

    fused = (mask is None and override_color is None and not pipe.compute_cov3D_python
             and not pipe.convert_SHs_python and all(hasattr(pc, a) for a in _RAW)
             and pc._scaling.shape[-1] == 3)
    if fused:
        rendered_image, radii, depth, opacity, n_touched = rasterize_gaussian_model(
            pc._xyz, screenspace_points, pc._features_dc, pc._features_rest, pc._opacity, pc._scaling,
            pc._rotation, theta, rho, raster_settings)
    else:
        # the reference's composition (gaussian_renderer/__init__.py:77-141) on the drop-in rasterizer
        rasterizer = GaussianRasterizer(raster_settings=raster_settings)
        opacity_in = pc.get_opacity
        scales = rotations = cov3D_precomp = shs = colors_precomp = None
        if pipe.compute_cov3D_python:
            cov3D_precomp = pc.get_covariance(scaling_modifier)
        else:
            scales = pc.get_scaling.repeat(1, 3) if pc.get_scaling.shape[-1] == 1 else pc.get_scaling
            rotations = pc.get_rotation
        if override_color is None:
            if pipe.convert_SHs_python:
                raise NotImplementedError("convert_SHs_python: evaluate SH in Python and pass override_color")
            shs = pc.get_features
        else:
            colors_precomp = override_color
        means2D = screenspace_points
        if mask is not None:
            # the reference unpacks four of the five outputs on this branch and would raise
            # (gaussian_renderer/__init__.py:116-128); here the branch works
            sel = lambda t: None if t is None else t[mask]
            xyz, means2D, shs, colors_precomp, opacity_in = xyz[mask], means2D[mask], sel(shs), sel(colors_precomp), opacity_in[mask]
            scales, rotations, cov3D_precomp = sel(scales), sel(rotations), sel(cov3D_precomp)
        rendered_image, radii, depth, opacity, n_touched = rasterizer(
            means3D=xyz, means2D=means2D, shs=shs, colors_precomp=colors_precomp, opacities=opacity_in,
            scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp, theta=theta, rho=rho)
    return {
        "render": rendered_image,
        "viewspace_points": screenspace_points,
        "visibility_filter": radii > 0,
        "radii": radii,
        "depth": depth,
        "opacity": opacity,
        "n_touched": n_touched,
    }
