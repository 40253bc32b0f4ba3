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
    # two steps: the densification statistics accumulate over steps (the reference consumes them every ~100
    # iterations, /root/reference/train.py:127-133) and are reduced on demand; gradients are per step
    st.step([cams[k] for k in mine], [dLs[k] for k in mine])
    st.step([cams[k] for k in mine], [dLs[k] for k in mine])
    g = st.grads()
    accum, denom, maxr = st.reduce_densification_stats()
    again = st.reduce_densification_stats()                 # idempotent: the local accumulators are untouched
    assert torch.equal(again[0], accum) and torch.equal(again[1], denom) and torch.equal(again[2], maxr)
    torch.save({"grads": [t.clone() for t in g], "accum": accum, "denom": denom, "maxr": maxr, "views": mine},
               os.path.join(outdir, f"r{rank}.pt"))
    st.reset_densification_stats()
    assert float(st.xyz_gradient_accum.abs().max()) == 0.0 and float(st.denom.max()) == 0.0
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
    single.step(cams, dLs)                                            # statistics of two steps, like the ranks
    for a, b, c in zip(r0["grads"], r1["grads"], single.grads()):
        assert torch.equal(a, b)                                      # every rank holds the reduced gradient
        assert float((a - c).abs().max()) <= 1e-3 * max(float(c.abs().max()), 1e-12)
    assert torch.allclose(r0["accum"], single.xyz_gradient_accum, rtol=1e-4, atol=1e-9)
    assert torch.equal(r0["denom"], single.denom) and torch.equal(r0["maxr"], single.max_radii2D)
    assert float(single.denom.max()) > 2                              # some Gaussian seen from several views, twice
    assert torch.equal(r0["accum"], r1["accum"])


# ---- factorized exchange (all-gather of masked colour gradients + local SH rebuild) ---------------------------
class _TorchBackend:
    """CPU stand-in for the CUDA backend of the factorized exchange, built on the PyTorch oracle."""

    def __init__(self, sh_degree, bg):
        self.D, self.bg = sh_degree, bg

    def view_fwd_bwd(self, cam, dL, lv, out, on_forward=None, on_colors=None):
        leaves = vp.GaussianParams(*[t.detach().clone().requires_grad_(True) for t in lv])
        m2d = torch.zeros_like(leaves.means3D, requires_grad=True)
        color, radii = _oracle_rasterize(cam, leaves, m2d, self.D, self.bg, 1.0)
        torch.autograd.backward([color], [dL])
        out["dL_dmeans3D"].copy_(leaves.means3D.grad); out["dL_dopacity"].copy_(leaves.opacities.grad.view(-1, 1))
        out["dL_dscales"].copy_(leaves.scales.grad); out["dL_drotations"].copy_(leaves.rotations.grad)
        out["dL_dcolors"].copy_(leaves.shs.grad[:, 0, :] / 0.28209479177387814)   # Y_0 is constant: dcol = dL/dsh_0 / Y_0
        if on_colors is not None:
            on_colors()
        return color.detach(), radii, m2d.grad

    def sh_grad_from_views(self, means3D, dcols, campos, M, degree):
        from oracle import torch_ref as tr
        nb = (degree + 1) ** 2
        out = torch.zeros(means3D.shape[0], M, 3)
        for v, dc in enumerate(dcols):
            dirs = means3D - campos[v]
            dirs = dirs / dirs.norm(dim=1, keepdim=True)
            for k in range(nb):
                onehot = torch.zeros(means3D.shape[0], 1, nb); onehot[:, 0, k] = 1.0
                out[:, k, :] += tr.eval_sh(degree, onehot, dirs)[:, :1] * dc
        return out


def _worker_fact(rank, world, initfile, outdir):
    dist.init_process_group("gloo", init_method=f"file://{initfile}", rank=rank, world_size=world)
    torch.set_num_threads(2)
    params, cams, dLs = _setup()
    mine = vp.shard_views(NV, rank, world)
    st = vp.ViewParallelStep(params, D, torch.zeros(3), world_size=world, exchange="factorized",
                             backend=_TorchBackend(D, torch.zeros(3)))
    st.step([cams[k] for k in mine], [dLs[k] for k in mine])
    st.step([cams[k] for k in mine], [dLs[k] for k in mine])
    local_accum = st.xyz_gradient_accum.clone()
    accum, denom, maxr = st.reduce_densification_stats()
    torch.save({"grads": [t.clone() for t in st.grads()], "accum": accum, "local_accum": local_accum,
                "denom": denom, "maxr": maxr}, os.path.join(outdir, f"f{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_factorized_exchange_matches_single_process():
    with tempfile.TemporaryDirectory() as d:
        initfile = os.path.join(d, "init")
        mp.spawn(_worker_fact, args=(2, initfile, d), nprocs=2, join=True)
        r0, r1 = torch.load(os.path.join(d, "f0.pt")), torch.load(os.path.join(d, "f1.pt"))
    params, cams, dLs = _setup()
    single = vp.ViewParallelStep(params, D, torch.zeros(3), world_size=1, rasterize=_oracle_rasterize)
    single.step(cams, dLs)
    single.step(cams, dLs)
    for a, b, c in zip(r0["grads"], r1["grads"], single.grads()):
        assert float((a - b).abs().max()) <= 1e-6 * max(float(c.abs().max()), 1e-12)     # ranks agree
        assert float((a - c).abs().max()) <= 1e-3 * max(float(c.abs().max()), 1e-12)     # == sum over all views
    assert not torch.equal(r0["local_accum"], r0["accum"])                               # stats were rank-local until reduced
    assert torch.allclose(r0["accum"], single.xyz_gradient_accum, rtol=1e-4, atol=1e-9)
    assert torch.equal(r0["denom"], single.denom) and torch.equal(r0["maxr"], single.max_radii2D)
