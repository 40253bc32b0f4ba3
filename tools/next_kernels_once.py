#!/usr/bin/env python
"""Launches each kernel of the widened rows a few times at BASELINE configs[1] sizes, for an
ncu capture (`ncu --set full -k regex:...`).  Not a benchmark: timings come from tools/bench_next.py."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
importlib.import_module("3dgs_study_b200")
synth = importlib.import_module("3dgs_study_b200.synth")
import gaussian_training as gtr  # noqa: E402
from simple_knn._C import distCUDA2  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
gt = torch.rand(3, 1080, 1920, generator=g).to(dev)
img = (gt + 0.1 * torch.randn(3, 1080, 1920, generator=g).to(dev)).clamp(0, 1).requires_grad_()
for _ in range(2):
    img.grad = None
    loss, _ = gtr.training_loss(img, gt, 0.2)
    loss.backward()

P = 3_000_000
raw = [torch.randn(P, 3, device=dev).requires_grad_(), torch.randn(P, 4, device=dev).requires_grad_(),
       torch.randn(P, 1, device=dev).requires_grad_()]
for _ in range(2):
    s, r, o = gtr.activate(*raw)
    (s.sum() + r.sum() + o.sum()).backward()

feats = torch.nn.Parameter(torch.randn(P, 16, 3, device=dev))
xyz = torch.nn.Parameter(torch.randn(P, 3, device=dev))
opt = gtr.FusedAdam([{"params": [xyz], "lr": 1.6e-4, "name": "xyz"},
                     {"params": [feats], "lr": 0.0025, "lr_tail": 0.0025 / 20, "row_width": 48, "row_split": 3, "name": "features"},
                     {"params": [raw[2]], "lr": 0.05, "name": "opacity"}, {"params": [raw[0]], "lr": 0.005, "name": "scaling"},
                     {"params": [raw[1]], "lr": 0.001, "name": "rotation"}], lr=0.0, eps=1e-15)
for p in (xyz, feats):
    p.grad = torch.randn_like(p) * 1e-3
for _ in range(2):
    opt.step()

pts = synth.make_scene(1_000_000, 0, seed=0).means3D.to(dev).contiguous()
for _ in range(2):
    d = distCUDA2(pts)
torch.cuda.synchronize()
print("ok", float(loss.item()), float(d.mean().item()))
