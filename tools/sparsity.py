#!/usr/bin/env python
"""Fraction of Gaussians with a non-zero masked colour gradient per view (bench scene): how sparse the
view-parallel exchange's [P,3] payload is."""
import importlib, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3dgs_study_b200"); pkg.ensure_on_path()
synth = importlib.import_module("3dgs_study_b200.synth")
import diff_gaussian_rasterization as dgr
C = dgr._C
P, W, H = 3_000_000, 1920, 1080
dev = torch.device("cuda:0")
sc = synth.make_scene(P, 3, seed=0).to(dev)
dL = synth.make_grad_image(W, H).to(dev)
bg = torch.zeros(3, device=dev)
union = torch.zeros(P, dtype=torch.bool, device=dev)
for k in range(8):
    cam = synth.make_camera(W, H, yaw=2.0 * np.pi * k / 8).to(dev)
    R, color, radii, geom, binning, img, _ = C.rasterize_gaussians(bg, sc.means3D, None, sc.opacities, sc.scales, sc.rotations, 1.0, None,
        cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy, H, W, sc.shs, 3, cam.camera_center, False, False)
    out = C.rasterize_gaussians_backward(bg, sc.means3D, radii, None, sc.scales, sc.rotations, 1.0, None, cam.world_view_transform,
        cam.full_proj_transform, cam.tanfovx, cam.tanfovy, dL, sc.shs, 3, cam.camera_center, geom, R, binning, img, False,
        opacities=sc.opacities, image_height=H, image_width=W, masked_colors=True)
    dcol = out[1]
    nz = (dcol != 0).any(dim=1)
    union |= nz
    # 32-byte sectors of the dense [P,3] f32 array that hold at least one non-zero row
    rows = torch.nonzero(nz).flatten()
    sect = torch.unique(torch.cat([(rows * 12) // 32, (rows * 12 + 11) // 32]))
    gm = (out[3] != 0).any(dim=1)
    print(f"view {k}: visible {int((radii > 0).sum())}  nonzero dcol rows {int(nz.sum())} ({100 * float(nz.float().mean()):.1f} %)  "
          f"sectors touched {100 * sect.numel() / (P * 12 / 32):.1f} %  nonzero dL_dmeans3D rows {100 * float(gm.float().mean()):.1f} %")
print(f"union over 8 views: {100 * float(union.float().mean()):.1f} %")
