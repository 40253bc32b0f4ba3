#!/usr/bin/env python
"""Workload counters of the render stage for the bench scene (b200splat_render_stats)."""
import importlib, json, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3dgs_study_b200"); pkg.ensure_on_path()
synth = importlib.import_module("3dgs_study_b200.synth")
import diff_gaussian_rasterization as dgr
C = dgr._C
P = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
H = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
dev = torch.device("cuda:0")
sc = synth.make_scene(P, 3, seed=0).to(dev)
for k in range(int(sys.argv[4]) if len(sys.argv) > 4 else 2):
    cam = synth.make_camera(W, H, yaw=2.0 * np.pi * k / 8).to(dev)
    R, color, radii, geom, binning, img, _ = C.rasterize_gaussians(torch.zeros(3, device=dev), sc.means3D, None, sc.opacities, sc.scales, sc.rotations, 1.0, None,
        cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy, H, W, sc.shs, 3, cam.camera_center, False, False)
    st = C.render_stats(P, W, H, R, geom, binning, img)
    st.update(view=k, R=R, visible=int((radii > 0).sum()))
    print(json.dumps(st))
    # per-tile depth of the deepest contributor (= list entries the backward visits) and list lengths
    st2 = C.export_state(P, W, H, R, radii, geom, binning, img)
    nc = st2["n_contrib"].float()
    gy, gx = (H + 15) // 16, (W + 15) // 16
    pad = torch.zeros(gy * 16, gx * 16, device=dev); pad[:H, :W] = nc
    tmax = pad.view(gy, 16, gx, 16).amax(dim=(1, 3)).flatten()
    ln = (st2["ranges"][:, 1] - st2["ranges"][:, 0]).float()
    q = torch.tensor([0.5, 0.9, 0.99, 1.0], device=dev)
    print("tile deepest-contributor quantiles 50/90/99/100:", torch.quantile(tmax, q).tolist(), "mean", float(tmax.mean()))
    print("tile list-length quantiles 50/90/99/100:", torch.quantile(ln, q).tolist(), "mean", float(ln.mean()))
