#!/usr/bin/env python
"""Generate golden vectors by running the UNMODIFIED reference rasterizer (oracle/_ref,
built for sm_100 by oracle/ref_cuda/build_ref.py) on seeded synthetic scenes.

Run on a B200:  gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'
then copy gpurun_out/golden/*.npz into tests/golden/.  The scenes are regenerated from their
seeds (wildgs_slam_blackwell_b200/scenes.py), so only the reference OUTPUTS are stored.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.ref_cuda import ref_api  # noqa: E402
from wildgs_slam_blackwell_b200 import scenes  # noqa: E402

# name -> dict(P, W, H, deg, seed, bg, precomp)
CASES = {
    "g0_deg0": dict(P=2000, W=160, H=96, deg=0, seed=11, bg=(0.0, 0.0, 0.0), precomp=False),
    "g1_deg3_bg_ragged": dict(P=1500, W=150, H=100, deg=3, seed=12, bg=(0.3, 0.5, 0.7), precomp=False),
    "g2_precomp": dict(P=1000, W=96, H=64, deg=0, seed=13, bg=(0.1, 0.2, 0.3), precomp=True),
    "g3_deg1_dense": dict(P=4000, W=64, H=48, deg=1, seed=14, bg=(0.0, 0.0, 0.0), precomp=False),
}


def precomp_inputs(g, seed):
    """colors_precomp (P,3) and cov3D_precomp (P,6) derived deterministically from the scene."""
    rng = np.random.default_rng(seed + 500)
    P = g["means3D"].shape[0]
    colors = rng.uniform(0, 1, (P, 3)).astype(np.float32)
    s = g["scales"].astype(np.float64)
    q = g["rotations"].astype(np.float64)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                  2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                  2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).reshape(P, 3, 3)
    Sg = R @ (s[:, :, None] ** 2 * np.transpose(R, (0, 2, 1)))
    cov = np.stack([Sg[:, 0, 0], Sg[:, 0, 1], Sg[:, 0, 2], Sg[:, 1, 1], Sg[:, 1, 2], Sg[:, 2, 2]], 1)
    return colors, np.ascontiguousarray(cov, np.float32)


def build_case(c):
    cam = scenes.make_camera(c["W"], c["H"], c["seed"], bg=c["bg"])
    g = scenes.make_gaussians(c["P"], c["W"], c["H"], c["deg"], c["seed"], cam)
    cot = scenes.make_cotangents(c["W"], c["H"], c["seed"])
    colors = cov = None
    if c["precomp"]:
        colors, cov = precomp_inputs(g, c["seed"])
    return g, cam, cot, colors, cov


def main():
    outdir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(outdir, exist_ok=True)
    dev = "cuda:0"
    for name, c in CASES.items():
        g, cam, cot, colors, cov = build_case(c)
        gt, ct = ref_api.to_cuda(g, dev), ref_api.to_cuda(cam, dev)
        cp = None if colors is None else torch.from_numpy(colors).to(dev)
        cv = None if cov is None else torch.from_numpy(cov).to(dev)
        ref = ref_api.RefRasterizer()
        out = ref.forward(gt, ct, colors_precomp=cp, cov3D_precomp=cv)
        st = ref.export_state()
        gr = ref.backward(torch.from_numpy(cot[0]).to(dev), torch.from_numpy(cot[1]).to(dev))
        rec = {}
        for k in ("color", "depth", "opacity", "radii", "n_touched"):
            rec[k] = out[k].cpu().numpy()
        rec["num_rendered"] = np.int64(out["num_rendered"])
        vis = rec["radii"] > 0
        keys = ["depths", "means2D", "conic_opacity", "tiles_touched"]
        if not c["precomp"]:      # the reference leaves these unwritten on the precomputed path
            keys += ["rgb", "clamped", "cov3D"]
        for k in keys:
            a = st[k].cpu().numpy()
            a[~vis] = 0          # culled entries are stale/uninitialised in the reference
            rec["st_" + k] = a
        for k in ("point_list_keys", "point_list", "ranges", "n_contrib", "final_T"):
            rec["st_" + k] = st[k].cpu().numpy()
        for k, v in gr.items():
            if k in ("dL_dtau",):
                continue        # (P,6): only its sum is part of the API
            rec["g_" + k] = v.cpu().numpy()
        np.savez_compressed(os.path.join(outdir, name + ".npz"), **rec)
        print(name, "R =", out["num_rendered"], "visible =", int(vis.sum()),
              "tau =", gr["dL_dtau_sum"].cpu().numpy())


if __name__ == "__main__":
    main()
