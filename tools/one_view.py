#!/usr/bin/env python
"""A few fwd+bwd views of the bench workload and nothing else: the command ncu wraps."""
import importlib, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def main():
    views = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 3_000_000
    W = int(sys.argv[3]) if len(sys.argv) > 3 else 1920
    H = int(sys.argv[4]) if len(sys.argv) > 4 else 1080
    pkg = importlib.import_module("3dgs_study_b200"); pkg.ensure_on_path()
    synth = importlib.import_module("3dgs_study_b200.synth")
    vp = importlib.import_module("3dgs_study_b200.view_parallel")
    dev = torch.device("cuda:0")
    D = 3
    sc = synth.make_scene(P, D, seed=0).to(dev)
    cams = [synth.make_camera(W, H, yaw=2.0 * np.pi * k / 8).to(dev) for k in range(8)]
    dL = synth.make_grad_image(W, H).to(dev)
    st = vp.ViewParallelStep(vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations), D, torch.zeros(3, device=dev), exchange="packed")     # the step bench.py times
    for i in range(views):
        st.step([cams[i % 8]], [dL])
    torch.cuda.synchronize()
    print("ok", float(st.grad_checksum()))

if __name__ == "__main__":
    main()
