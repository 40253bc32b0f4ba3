"""CPU, world_size 2, gloo: the N>1 host logic of the view-parallel step with the PyTorch
oracle injected as the rasterize function.  Checks: view sharding, sum-allreduce of the parameter
gradients == single-process sum over all views, densification statistics keep the reference's
norm-then-sum semantics (/root/reference/scene/gaussian_model.py:576-581) and max-reduced radii."""
import importlib
import os
import tempfile

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

synth = importlib.import_module("3dgs_study_b200.synth")
vp = importlib.import_module("3dgs_study_b200.view_parallel")

P, W, H, D, NV = 400, 64, 48, 1, 4


def _oracle_rasterize(cam, lv, means2D, sh_degree, bg, mod):
    from oracle import torch_ref as tr
    color, radii, _, _ = tr.rasterize_gaussians(
        image_height=cam.image_height, image_width=cam.image_width, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=bg,
        scale_modifier=mod, viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform, sh_degree=sh_degree,
        campos=cam.camera_center, means3D=lv.means3D, means2D=means2D, opacities=lv.opacities, shs=lv.shs,
        scales=lv.scales, rotations=lv.rotations)
    return color, radii


def _setup():
    sc = synth.make_scene(P, D, seed=21, s_med=0.08)
    cams = [synth.make_camera(W, H, yaw=2 * np.pi * k / NV) for k in range(NV)]
    dLs = [synth.make_grad_image(W, H, seed=30 + k) for k in range(NV)]
    params = vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)
    return params, cams, dLs


def _worker(rank, world, initfile, outdir):
    dist.init_process_group("gloo", init_method=f"file://{initfile}", rank=rank, world_size=world)
    torch.set_num_threads(2)
    params, cams, dLs = _setup()
    mine = vp.shard_views(NV, rank, world)
    st = vp.ViewParallelStep(params, D, torch.zeros(3), world_size=world, rasterize=_oracle_rasterize)
    st.step([cams[k] for k in mine], [dLs[k] for k in mine])
    g = st.grads()
    torch.save({"grads": [t.clone() for t in g], "accum": st.xyz_gradient_accum.clone(), "denom": st.denom.clone(),
                "maxr": st.max_radii2D.clone(), "views": mine}, os.path.join(outdir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_views():
    assert vp.shard_views(8, 1, 4) == [1, 5]
    assert sorted(sum([vp.shard_views(8, r, 3) for r in range(3)], [])) == list(range(8))
    assert vp.GRAD_FLOATS_PER_GAUSSIAN(16) == 59


def test_two_rank_step_matches_single_process():
    with tempfile.TemporaryDirectory() as d:
        initfile = os.path.join(d, "init")
        mp.spawn(_worker, args=(2, initfile, d), nprocs=2, join=True)
        r0, r1 = torch.load(os.path.join(d, "r0.pt")), torch.load(os.path.join(d, "r1.pt"))
    assert r0["views"] == [0, 2] and r1["views"] == [1, 3]
    params, cams, dLs = _setup()
    single = vp.ViewParallelStep(params, D, torch.zeros(3), world_size=1, rasterize=_oracle_rasterize)
    single.step(cams, dLs)
    for a, b, c in zip(r0["grads"], r1["grads"], single.grads()):
        assert torch.equal(a, b)                                      # every rank holds the reduced gradient
        assert float((a - c).abs().max()) <= 1e-3 * max(float(c.abs().max()), 1e-12)
    assert torch.allclose(r0["accum"], single.xyz_gradient_accum, rtol=1e-4, atol=1e-9)
    assert torch.equal(r0["denom"], single.denom) and torch.equal(r0["maxr"], single.max_radii2D)
    assert float(single.denom.max()) > 1                              # some Gaussian seen from several views
