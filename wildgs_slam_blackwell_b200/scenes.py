"""Seeded synthetic scenes for the rasterizer path.

numpy only (PCG64 is platform-independent), so the same seed gives the same bytes in this
container, on the GPU box, in the oracle, the tests and bench.py.  Camera conventions follow
the reference callers: `viewmatrix = W2C^T`, `projmatrix = viewmatrix @ P^T` with P from
getProjectionMatrix2 (thirdparty/gaussian_splatting/utils/graphics_utils.py:72-93, transposed
as in src/mapper.py:111-121), pose increments through SE3_exp (src/utils/pose_utils.py:66-98).
"""
import math

import numpy as np

SH_C0 = 0.28209479177387814

# name -> (P, W, H, sh_degree, n_views)   (BASELINE.json configs / BASELINE.md §4)
CONFIGS = {
    "config1": (10_000, 320, 240, 0, 1),
    "config2": (300_000, 640, 480, 0, 1),
    "config3": (1_000_000, 640, 480, 0, 10),
    "config4": (3_000_000, 1920, 1080, 3, 1),
    "config5": (2_000_000, 640, 480, 0, 16),
}


def _skew(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]], dtype=np.float64)


def se3_exp(tau):
    """tau = (rho, theta) -> 4x4, float64 (src/utils/pose_utils.py:28-79)."""
    tau = np.asarray(tau, dtype=np.float64)
    rho, theta = tau[:3], tau[3:]
    W = _skew(theta)
    W2 = W @ W
    a = np.linalg.norm(theta)
    I = np.eye(3)
    if a < 1e-5:
        R = I + W + 0.5 * W2
        V = I + 0.5 * W + W2 / 6.0
    else:
        R = I + (math.sin(a) / a) * W + ((1 - math.cos(a)) / a**2) * W2
        V = I + W * ((1 - math.cos(a)) / a**2) + W2 * ((a - math.sin(a)) / a**3)
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = V @ rho
    return T


def projection_matrix2(znear, zfar, cx, cy, fx, fy, W, H):
    left = ((2 * cx - W) / W - 1.0) * W / 2.0
    right = ((2 * cx - W) / W + 1.0) * W / 2.0
    top = ((2 * cy - H) / H + 1.0) * H / 2.0
    bottom = ((2 * cy - H) / H - 1.0) * H / 2.0
    left, right = znear / fx * left, znear / fx * right
    top, bottom = znear / fy * top, znear / fy * bottom
    P = np.zeros((4, 4), dtype=np.float64)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(W, H, seed, tau=None, bg=(0.0, 0.0, 0.0)):
    """Per-view constants in the layout GaussianRasterizationSettings wants (float32)."""
    rng = np.random.default_rng(seed)
    if tau is None:
        tau = np.concatenate([rng.normal(0, 0.05, 3), rng.normal(0, 0.03, 3)])
    fx = fy = 0.82 * W
    cx, cy = W / 2.0, H / 2.0
    w2c = se3_exp(tau)
    Pm = projection_matrix2(0.01, 100.0, cx, cy, fx, fy, W, H)
    view = w2c.T.astype(np.float32)                      # viewmatrix = W2C^T
    proj_raw = Pm.T.astype(np.float32)                   # projmatrix_raw = P^T
    proj = (view.astype(np.float64) @ proj_raw.astype(np.float64)).astype(np.float32)
    campos = np.linalg.inv(w2c)[:3, 3].astype(np.float32)
    return dict(width=W, height=H, fx=fx, fy=fy, tanfovx=W / (2 * fx), tanfovy=H / (2 * fy),
                viewmatrix=np.ascontiguousarray(view), projmatrix=np.ascontiguousarray(proj),
                projmatrix_raw=np.ascontiguousarray(proj_raw), campos=campos,
                bg=np.asarray(bg, dtype=np.float32), w2c=w2c, tau=np.asarray(tau))


def make_gaussians(P, W, H, sh_degree, seed, cam0=None):
    """Gaussians drawn in camera-0's frustum (enlarged 7.5 % per side), mapped to world."""
    rng = np.random.default_rng(seed)
    if cam0 is None:
        cam0 = make_camera(W, H, seed)
    fx, fy = cam0["fx"], cam0["fy"]
    z = rng.uniform(0.5, 6.0, P)
    near = rng.random(P) < 0.03
    z = np.where(near, rng.uniform(-1.0, 0.2, P), z)
    u = rng.uniform(-0.075 * W, 1.075 * W, P)
    v = rng.uniform(-0.075 * H, 1.075 * H, P)
    zs = np.where(near, np.abs(z) + 0.1, z)
    xc = (u - W / 2.0) * zs / fx
    yc = (v - H / 2.0) * zs / fy
    pc = np.stack([xc, yc, z, np.ones(P)], 1)
    pw = (np.linalg.inv(cam0["w2c"]) @ pc.T).T[:, :3]
    sigma_px = np.exp(rng.normal(math.log(1.5), 0.5, P))
    scales = (sigma_px * zs / fx)[:, None] * np.exp(rng.normal(0, 0.3, (P, 3)))
    q = rng.normal(0, 1, (P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    opac = 1.0 / (1.0 + np.exp(-rng.normal(0.5, 1.5, (P, 1))))
    M = (sh_degree + 1) ** 2
    sh = np.zeros((P, M, 3))
    sh[:, 0, :] = (rng.uniform(0, 1, (P, 3)) - 0.5) / SH_C0
    if M > 1:
        sh[:, 1:, :] = rng.normal(0, 0.1, (P, M - 1, 3))
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    return dict(means3D=f32(pw), scales=f32(scales), rotations=f32(q), opacities=f32(opac),
                sh=f32(sh), sh_degree=sh_degree)


def make_cotangents(W, H, seed):
    """Fixed-seed dL/dcolor (3,H,W) and dL/ddepth (1,H,W) for parity runs."""
    rng = np.random.default_rng(seed + 77)
    N = W * H
    dc = rng.normal(0, 1, (3, H, W)) / (3 * N)
    dd = rng.normal(0, 1, (1, H, W)) / N
    return np.ascontiguousarray(dc, np.float32), np.ascontiguousarray(dd, np.float32)


def make_scene(name_or_P, W=None, H=None, sh_degree=0, seed=None, n_views=1, bg=(0.0, 0.0, 0.0)):
    """Returns (gaussians, [cameras], [(dL_dcolor, dL_ddepth)])."""
    if isinstance(name_or_P, str):
        P, W, H, sh_degree, n_views = CONFIGS[name_or_P]
        if seed is None:
            seed = 1000 + int(name_or_P[-1])
    else:
        P = int(name_or_P)
        if seed is None:
            seed = 1000
    cams = [make_camera(W, H, seed + k, bg=bg) for k in range(n_views)]
    g = make_gaussians(P, W, H, sh_degree, seed, cams[0])
    cots = [make_cotangents(W, H, seed + k) for k in range(n_views)]
    return g, cams, cots
