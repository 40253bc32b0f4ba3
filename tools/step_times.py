#!/usr/bin/env python
"""Per-step wall times of the bench step (sync after every step) + allocator statistics: diagnostic."""
import importlib, os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3dgs_study_b200"); pkg.ensure_on_path()
synth = importlib.import_module("3dgs_study_b200.synth")
vp = importlib.import_module("3dgs_study_b200.view_parallel")
dev = torch.device("cuda:0")
P, W, H, D = 3_000_000, 1920, 1080, 3
sc = synth.make_scene(P, D, seed=0).to(dev)
cams = [synth.make_camera(W, H, yaw=2.0 * np.pi * k / 8).to(dev) for k in range(8)]
dL = synth.make_grad_image(W, H).to(dev)
st = vp.ViewParallelStep(vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations), D, torch.zeros(3, device=dev))
ts = []
for i in range(40):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    st.step([cams[i % 8]], [dL])
    torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    if i in (9, 39):
        ms = torch.cuda.memory_stats()
        print(i, "num_alloc_retries", ms["num_alloc_retries"], "segments", ms["segment.all.current"], "device_alloc", ms.get("num_device_alloc"), "device_free", ms.get("num_device_free"),
              "reserved GB", ms["reserved_bytes.all.current"] / 1e9)
print(" ".join(f"{t:.2f}" for t in ts))
t0 = time.perf_counter()
for i in range(40):
    st.step([cams[i % 8]], [dL])
torch.cuda.synchronize()
print("async ms/step", (time.perf_counter() - t0) / 40 * 1e3)
