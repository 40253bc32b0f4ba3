#!/usr/bin/env python
"""Generate tests/golden/reference_math.npz by IMPORTING THE REFERENCE (/root/reference) in the build
container.  The reference cannot travel to the GPU box, so the vectors are committed; this script is the
record of how they were made.

What the reference holds for this path: the Python math around the (absent) rasterizer --
SH evaluation, quaternion -> rotation, covariance build + packing, camera matrices.  Functions that
hard-code device="cuda" (utils/general_utils.py:73,104,120) are run on the CPU by temporarily mapping
"cuda" -> "cpu" in torch.zeros; their arithmetic is untouched.

    python tests/golden/make_golden.py
"""
import math
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)

_zeros = torch.zeros


def _zeros_cpu(*a, **k):
    if str(k.get("device", "")).startswith("cuda"):
        k["device"] = "cpu"
    return _zeros(*a, **k)


def main():
    from utils import sh_utils, graphics_utils, general_utils
    g = torch.Generator().manual_seed(1234)
    out = {}
    # eval_sh, degrees 0..3 (utils/sh_utils.py:57-112) + the +0.5 / clamp of gaussian_renderer/__init__.py:90
    N = 64
    sh = torch.randn(N, 3, 16, generator=g) * 0.4
    dirs = torch.randn(N, 3, generator=g)
    dirs = dirs / dirs.norm(dim=1, keepdim=True)
    out["sh_coeffs"] = sh.numpy(); out["sh_dirs"] = dirs.numpy()
    for d in range(4):
        r = sh_utils.eval_sh(d, sh, dirs)
        out[f"sh_eval_deg{d}"] = r.numpy()
        out[f"sh_color_deg{d}"] = torch.clamp_min(r + 0.5, 0.0).numpy()
    out["RGB2SH"] = sh_utils.RGB2SH(torch.linspace(0, 1, 5)).numpy()
    # rotation / covariance (utils/general_utils.py:72-128, scene/gaussian_model.py:27-32)
    q = torch.randn(N, 4, generator=g)
    s = torch.exp(torch.randn(N, 3, generator=g) * 0.5)
    torch.zeros = _zeros_cpu
    try:
        R = general_utils.build_rotation(q)
        L = general_utils.build_scaling_rotation(1.7 * s, q)
        cov = general_utils.strip_symmetric(L @ L.transpose(1, 2))
    finally:
        torch.zeros = _zeros
    out["quat"] = q.numpy(); out["scale"] = s.numpy(); out["scale_mod"] = np.float32(1.7)
    out["rotation"] = R.numpy(); out["L"] = L.numpy(); out["cov6"] = cov.numpy()
    # camera matrices (utils/graphics_utils.py:49-87,97-133; scene/cameras.py:103-121)
    th = 0.7
    Rm = np.array([[math.cos(th), 0, math.sin(th)], [0, 1, 0], [-math.sin(th), 0, math.cos(th)]])
    T = np.array([0.3, -0.2, 4.5])
    fovx, fovy = 1.2, 0.8
    wvt = torch.tensor(graphics_utils.getWorld2View2(Rm, T, np.array([0.0, 0.0, 0.0]), 1.0)).transpose(0, 1)
    proj = graphics_utils.getProjectionMatrix(znear=0.01, zfar=100.0, fovX=fovx, fovY=fovy).transpose(0, 1)
    full = wvt.unsqueeze(0).bmm(proj.unsqueeze(0)).squeeze(0)
    out["cam_R"] = Rm; out["cam_T"] = T; out["cam_fov"] = np.array([fovx, fovy])
    out["world_view_transform"] = wvt.numpy(); out["projection_matrix"] = proj.numpy()
    out["full_proj_transform"] = full.numpy(); out["camera_center"] = wvt.inverse()[3, :3].numpy()
    pts = torch.randn(16, 3, generator=g)
    out["points"] = pts.numpy()
    out["points_ndc"] = graphics_utils.geom_transform_points(pts, full).numpy()   # :25-32
    out["fov2focal"] = np.array([graphics_utils.fov2focal(fovx, 1920)])
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_math.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, {k: v.shape for k, v in out.items() if hasattr(v, "shape")})


if __name__ == "__main__":
    main()
