#!/usr/bin/env python
"""BASELINE.json configs[3] as SURVEY.md 8(d) defines it: a training step of 8 views, P = 3 M SH3 1080p, on G ranks:
rank g renders views {g, g+G, ...} of an 8-camera orbit one after the other, accumulating their gradient rows, then ONE
exchange per step.  Reports views/s (= 8 / step time, max over ranks); efficiency is computed by the reader from the
G = 1 line.  Launch: python tools/bench_config4.py            (G = 1)
                torchrun --nproc-per-node G tools/bench_config4.py [--preset translucent] [--exchange packed|factorized]"""
import argparse, importlib, json, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=12)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--gaussians", type=int, default=3_000_000)
ap.add_argument("--preset", default="opaque")
ap.add_argument("--exchange", default=None)
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
import torch.distributed as dist
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
pkg = importlib.import_module("3dgs_study_b200"); pkg.ensure_on_path()
synth = importlib.import_module("3dgs_study_b200.synth")
vp = importlib.import_module("3dgs_study_b200.view_parallel")
W, H, D, NV = 1920, 1080, 3, 8
sc = synth.make_scene(a.gaussians, D, seed=0, preset=a.preset).to(dev)
cams = [synth.make_camera(W, H, yaw=2.0 * np.pi * k / NV).to(dev) for k in range(NV)]
dL = synth.make_grad_image(W, H).to(dev)
mine = vp.shard_views(NV, rank, world)
st = vp.ViewParallelStep(vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations), D, torch.zeros(3, device=dev),
                         world_size=world, exchange=a.exchange or ("packed" if world == 1 else None))
def step():
    st.step([cams[k] for k in mine], [dL] * len(mine))
for _ in range(a.warmup):
    step()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.steps):
    step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.steps
if world > 1:
    t = torch.tensor([ms], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
touched = int((st.leaves.means3D.grad.abs().sum(dim=1) > 0).sum())
if rank == 0:
    print(json.dumps({"config": "configs[3]: 8 views per step, 3 M SH3 1080p, views sharded over ranks, one exchange per step",
                      "n_gpus": world, "views_per_step": NV, "views_per_rank": len(mine), "exchange": st.exchange, "preset": a.preset,
                      "ms_per_step": round(ms, 4), "views_per_s": round(NV * 1000.0 / ms, 2),
                      "gaussians_with_gradient_after_step": touched}), flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
