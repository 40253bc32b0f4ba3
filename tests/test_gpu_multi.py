"""N = 2 GPUs (skipped on a one-GPU box): the view-parallel exchange flavours -- plain NCCL all-reduce, factorized over
NCCL, and the sparse packed-row exchange over NVLink peer memory (symmetric-memory rendezvous, two barriers, peer
rotation) -- give the same summed gradients and the same densification statistics over several steps.
tools/check_exchange.py is the torchrun flavour of the same check for N = 4 / 8 (output kept in profiles/)."""
import importlib
import os
import tempfile

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _worker(rank, world, initfile, outdir):
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", init_method=f"file://{initfile}", rank=rank, world_size=world, device_id=dev)
    importlib.import_module("3dgs_study_b200").ensure_on_path()
    synth = importlib.import_module("3dgs_study_b200.synth")
    vp = importlib.import_module("3dgs_study_b200.view_parallel")
    report = {}
    for P, W, H, D, nv in ((20000, 320, 200, 3, 1), (4321, 200, 120, 1, 2)):
        sc = synth.make_scene(P, D, seed=0, s_med=0.03).to(dev)
        cams = [synth.make_camera(W, H, yaw=2.0 * np.pi * (rank * nv + k) / (world * nv)).to(dev) for k in range(nv)]
        dLs = [synth.make_grad_image(W, H, seed=5 + rank * nv + k).to(dev) for k in range(nv)]
        params = vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)
        bg = torch.zeros(3, device=dev)
        res = {}
        for name, kw in (("allreduce", dict(exchange="allreduce")), ("factorized", dict(exchange="factorized", peer_memory=False)),
                         ("packed", dict(exchange="packed"))):
            st = vp.ViewParallelStep(params, D, bg, world_size=world, **kw)
            for _ in range(3):                      # repeated steps: buffer reuse, the second barrier, accumulating stats
                st.step(cams, dLs)
            torch.cuda.synchronize()
            assert st.exchange == name, f"{name} fell back to {st.exchange}"
            res[name] = ([g.clone() for g in st.grads()],) + tuple(st.reduce_densification_stats())
        for name in ("factorized", "packed"):
            for f, a, b in zip(vp.GaussianParams._fields, res["allreduce"][0], res[name][0]):
                report[f"P{P}/{name}/{f}"] = float((a - b).abs().max()) / max(float(a.abs().max()), 1e-20)
            report[f"P{P}/{name}/stats"] = float(torch.allclose(res["allreduce"][1], res[name][1], rtol=1e-4, atol=1e-10)
                                                 and torch.equal(res["allreduce"][2], res[name][2])
                                                 and torch.equal(res["allreduce"][3], res[name][3]))
        report[f"P{P}/denom_max"] = float(res["packed"][2].max())
    torch.save(report, os.path.join(outdir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_two_rank_exchanges_agree():
    import torch.multiprocessing as mp
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(2, os.path.join(d, "init"), d), nprocs=2, join=True)
        reps = [torch.load(os.path.join(d, f"r{r}.pt")) for r in range(2)]
    for rep in reps:
        for k, v in rep.items():
            if k.endswith("/stats"):
                assert v == 1.0, (k, rep)
            elif k.endswith("denom_max"):
                assert v >= 3.0, (k, v)             # three accumulated steps
            else:
                assert v <= 5e-4, (k, v, rep)       # float-atomic summation order differs between flavours
