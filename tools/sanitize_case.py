#!/usr/bin/env python
"""The workload tools/sanitize.sh runs under compute-sanitizer: the smoke() case, an antialiasing + inverse-depth
case, the dense_opaque case (early exits, alpha clamp), a precomputed-colour/covariance case, the packed-row
backward + accumulate kernels, the loss / Adam / knn kernels -- every kernel of libb200splat.so at a small size."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
importlib.import_module("3dgs_study_b200").ensure_on_path()
from _util import Case  # noqa: E402

vp = importlib.import_module("3dgs_study_b200.view_parallel")
synth = importlib.import_module("3dgs_study_b200.synth")
dev = torch.device("cuda:0")

cases = {
    "smoke": dict(P=4000, W=256, H=192, D=3, seed=5, s_med=0.03),
    "aa_invdepth": dict(P=3000, W=200, H=136, D=3, seed=1, s_med=0.04, antialiasing=True),
    "dense_opaque": dict(P=6000, W=128, H=96, D=2, seed=4, s_med=0.25),
    "precomp": dict(P=3000, W=160, H=120, D=0, seed=5, s_med=0.05, precomp_color=True, precomp_cov=True),
    "ragged": dict(P=5000, W=331, H=173, D=3, seed=2, s_med=0.03),
}
for name, kw in cases.items():
    c = Case(**kw)
    c.cuda_run(invdepth_grad=(name == "aa_invdepth"))
    c.cuda_state(want_invdepth=True)
    print("ran", name, flush=True)

sc = synth.make_scene(5000, 3, seed=3, s_med=0.04).to(dev)
cams = [synth.make_camera(200, 120, yaw=0.4 * k).to(dev) for k in range(2)]
dLs = [synth.make_grad_image(200, 120, seed=40 + k).to(dev) for k in range(2)]
params = vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)
for mode in ("packed", "factorized"):
    st = vp.ViewParallelStep(params, 3, torch.zeros(3, device=dev), exchange=mode)
    st.step(cams, dLs); st.step(cams, dLs)
    torch.cuda.synchronize()
    print("ran exchange", mode, flush=True)

import gaussian_training as gt  # noqa: E402
img = torch.rand(3, 120, 200, device=dev, requires_grad=True)
gtimg = torch.rand(3, 120, 200, device=dev)
loss = gt.training_loss(img, gtimg, 0.2)
loss.backward()
from simple_knn._C import distCUDA2  # noqa: E402
d2 = distCUDA2(sc.means3D)
torch.cuda.synchronize()
print("ran loss / knn", float(loss), float(d2.mean()), flush=True)
