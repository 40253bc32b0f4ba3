#!/usr/bin/env python
"""Host-side cost of one view-parallel step (cProfile over a few steps, GPU kept busy)."""
import cProfile, importlib, os, pstats, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3dgs_study_b200"); pkg.ensure_on_path()
synth = importlib.import_module("3dgs_study_b200.synth")
vp = importlib.import_module("3dgs_study_b200.view_parallel")
dev = torch.device("cuda:0")
P = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
W, H, D = 1920, 1080, 3
sc = synth.make_scene(P, D, seed=0).to(dev)
cams = [synth.make_camera(W, H, yaw=2.0 * np.pi * k / 8).to(dev) for k in range(8)]
dL = synth.make_grad_image(W, H).to(dev)
st = vp.ViewParallelStep(vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations), D, torch.zeros(3, device=dev))
for i in range(5):
    st.step([cams[i % 8]], [dL])
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(20):
    st.step([cams[i % 8]], [dL])
torch.cuda.synchronize()
print("ms/step (wall):", (time.perf_counter() - t0) / 20 * 1e3)
# tiny scene: the GPU work vanishes, what remains is host time per step
sc2 = synth.make_scene(1000, D, seed=0).to(dev)
st2 = vp.ViewParallelStep(vp.GaussianParams(sc2.means3D, sc2.shs, sc2.opacities, sc2.scales, sc2.rotations), D, torch.zeros(3, device=dev))
cam2 = synth.make_camera(64, 64).to(dev); dL2 = synth.make_grad_image(64, 64).to(dev)
for i in range(5):
    st2.step([cam2], [dL2])
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(50):
    st2.step([cam2], [dL2])
torch.cuda.synchronize()
print("ms/step tiny scene (host-bound):", (time.perf_counter() - t0) / 50 * 1e3)
pr = cProfile.Profile(); pr.enable()
for i in range(50):
    st2.step([cam2], [dL2])
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
