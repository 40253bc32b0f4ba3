#!/usr/bin/env python
"""torchrun, N >= 2 GPUs: the exchange flavours (incl. the sparse packed-row exchange over peer memory) of the view-parallel step give the same summed gradients
(allreduce baseline == factorized over NCCL == factorized over NVLink peer memory), to float-summation noise."""
import importlib, os, sys
import numpy as np
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
pkg = importlib.import_module("3dgs_study_b200"); pkg.ensure_on_path()
synth = importlib.import_module("3dgs_study_b200.synth")
vp = importlib.import_module("3dgs_study_b200.view_parallel")
ok = True
for P, W, H, D in ((20000, 320, 200, 3), (4321, 200, 120, 1)):
    sc = synth.make_scene(P, D, seed=0, s_med=0.03).to(dev)
    cam = synth.make_camera(W, H, yaw=2.0 * np.pi * rank / world).to(dev)
    dL = synth.make_grad_image(W, H, seed=5 + rank).to(dev)
    params = vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)
    bg = torch.zeros(3, device=dev)
    res = {}
    for name, kw in (("allreduce", dict(exchange="allreduce")), ("nccl", dict(exchange="factorized", peer_memory=False)),
                     ("peer", dict(exchange="factorized", peer_memory="try")), ("packed", dict(exchange="packed"))):
        st = vp.ViewParallelStep(params, D, bg, world_size=world, **kw)
        for _ in range(3):     # repeated steps exercise buffer reuse / the second barrier
            st.step([cam], [dL])
        torch.cuda.synchronize()
        accum, denom, maxr = st.reduce_densification_stats()     # three steps' worth, summed / maxed over the ranks
        res[name] = ([g.clone() for g in st.grads()], accum, denom, maxr,
                     bool(st._symm and "flat" in st._symm))
    for name in ("nccl", "peer", "packed"):
        for f, a, b in zip(vp.GaussianParams._fields, res["allreduce"][0], res[name][0]):
            err = float((a - b).abs().max()) / max(float(a.abs().max()), 1e-20)
            if err > 5e-4:
                ok = False
            if rank == 0:
                print(f"P={P} {name:5s} {f:10s} rel err {err:.2e}")
        sa = (torch.allclose(res["allreduce"][1], res[name][1], rtol=1e-4, atol=1e-10) and torch.equal(res["allreduce"][2], res[name][2])
              and torch.equal(res["allreduce"][3], res[name][3]))
        ok = ok and sa
        if rank == 0:
            print(f"P={P} {name:5s} stats equal: {sa}; peer memory in use: {res[name][4]}")
t = torch.tensor([1.0 if ok else 0.0], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print("EXCHANGE CHECK", "OK" if t.item() == 1.0 else "FAILED")
dist.barrier(); dist.destroy_process_group()
