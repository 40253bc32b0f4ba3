#!/usr/bin/env python
"""Generate tests/golden/next_steps.npz by IMPORTING THE REFERENCE (/root/reference) in the build container
(it cannot travel to the GPU box; the vectors are committed, this script records how they were made).

Covers the steps either side of the rasterizer:
  * utils/loss_utils.py l1_loss (:17-24), ssim (:59-108) and their autograd gradients, and the training
    loss of train.py:103-104;
  * the reference's own optimizer: GaussianModel.training_setup (scene/gaussian_model.py:176-205) builds
    torch.optim.Adam over the six parameter groups, update_learning_rate (:207-214) + optimizer.step()
    (train.py:137-141) are driven for a few iterations.  `plyfile` and `simple_knn` (absent from this image)
    are stubbed for the import only and "cuda" is mapped to "cpu" in torch.zeros; no arithmetic is touched.

    python tests/golden/make_golden_next.py
"""
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)

_zeros = torch.zeros


def _zeros_cpu(*a, **k):
    if str(k.get("device", "")).startswith("cuda"):
        k["device"] = "cpu"
    return _zeros(*a, **k)


def loss_cases(out):
    from utils.loss_utils import l1_loss, ssim, gaussian, create_window
    for ws in (1, 3, 5, 7, 9, 11):
        out[f"window_taps_{ws}"] = gaussian(ws, 1.5).numpy()               # loss_utils.py:35-45
    out["window_2d_11"] = create_window(11, 3).numpy()                      # :46-57, [3,1,11,11]
    g = torch.Generator().manual_seed(4321)
    cases = {"chw": ((3, 37, 53), 11), "tiny": ((3, 7, 9), 11), "batch": ((2, 3, 24, 40), 11), "win7": ((3, 20, 31), 7),
             "wide": ((1, 33, 150), 11)}
    for name, (shape, ws) in cases.items():
        gt = torch.rand(*shape, generator=g)
        img = (gt + 0.15 * torch.randn(*shape, generator=g)).clamp(0, 1)
        img[..., :3, :4] = gt[..., :3, :4]          # exact zeros of the difference (sign(0) = 0)
        a = img.clone().requires_grad_()
        l1 = l1_loss(a, gt)
        (g_l1,) = torch.autograd.grad(l1, a)
        a2 = img.clone().requires_grad_()
        s = ssim(a2, gt, window_size=ws)
        (g_s,) = torch.autograd.grad(s, a2)
        a3 = img.clone().requires_grad_()
        loss = (1.0 - 0.2) * l1_loss(a3, gt) + 0.2 * (1.0 - ssim(a3, gt, window_size=ws))   # train.py:103-104
        (g_loss,) = torch.autograd.grad(loss, a3)
        out[f"loss_{name}_img"] = img.numpy()
        out[f"loss_{name}_gt"] = gt.numpy()
        out[f"loss_{name}_ws"] = np.int32(ws)
        out[f"loss_{name}_l1"] = np.float32(l1.item())
        out[f"loss_{name}_ssim"] = np.float32(s.item())
        out[f"loss_{name}_loss"] = np.float32(loss.item())
        out[f"loss_{name}_g_l1"] = g_l1.numpy()
        out[f"loss_{name}_g_ssim"] = g_s.numpy()
        out[f"loss_{name}_g_loss"] = g_loss.numpy()
        if len(shape) == 4:
            a4 = img.clone().requires_grad_()
            sb = ssim(a4, gt, window_size=ws, size_average=False)
            wgt = torch.tensor([0.3, -1.7])
            (g_sb,) = torch.autograd.grad((sb * wgt).sum(), a4)
            out[f"loss_{name}_ssim_per_batch"] = sb.detach().numpy()
            out[f"loss_{name}_g_ssim_per_batch_w"] = g_sb.numpy()
            out[f"loss_{name}_batch_w"] = wgt.numpy()


def adam_case(out):
    sys.modules.setdefault("plyfile", types.SimpleNamespace(PlyData=None, PlyElement=None))
    sk = types.ModuleType("simple_knn")
    skc = types.ModuleType("simple_knn._C")
    skc.distCUDA2 = None
    sys.modules.setdefault("simple_knn", sk)
    sys.modules.setdefault("simple_knn._C", skc)
    torch.zeros = _zeros_cpu
    try:
        from scene.gaussian_model import GaussianModel
        from arguments import OptimizationParams
        from argparse import ArgumentParser
        opt = OptimizationParams(ArgumentParser())
        g = torch.Generator().manual_seed(99)
        P, M = 41, 16
        gm = GaussianModel(3)
        gm.spatial_lr_scale = 2.5
        init = {"xyz": torch.randn(P, 3, generator=g), "f_dc": torch.randn(P, 1, 3, generator=g) * 0.5,
                "f_rest": torch.randn(P, M - 1, 3, generator=g) * 0.1, "opacity": torch.randn(P, 1, generator=g),
                "scaling": torch.randn(P, 3, generator=g) - 3.0, "rotation": torch.randn(P, 4, generator=g)}
        gm._xyz = torch.nn.Parameter(init["xyz"].clone())
        gm._features_dc = torch.nn.Parameter(init["f_dc"].clone())
        gm._features_rest = torch.nn.Parameter(init["f_rest"].clone())
        gm._opacity = torch.nn.Parameter(init["opacity"].clone())
        gm._scaling = torch.nn.Parameter(init["scaling"].clone())
        gm._rotation = torch.nn.Parameter(init["rotation"].clone())
        gm.training_setup(opt)
        params = {"xyz": gm._xyz, "f_dc": gm._features_dc, "f_rest": gm._features_rest, "opacity": gm._opacity,
                  "scaling": gm._scaling, "rotation": gm._rotation}
        iters = [1, 2, 3, 500, 29999]
        lrs = []
        for k, it in enumerate(iters):
            gm.update_learning_rate(it)
            for n, p in params.items():
                gr = torch.randn(p.shape, generator=g) * (10.0 ** float(torch.randint(-6, 0, (1,), generator=g)))
                if k == 2:
                    gr[::2] = 0.0            # half the Gaussians unseen in this view: exact-zero gradients
                p.grad = gr
                out[f"adam_grad_{n}_{k}"] = gr.numpy().copy()
            lrs.append([grp["lr"] for grp in gm.optimizer.param_groups])
            gm.optimizer.step()
            gm.optimizer.zero_grad(set_to_none=True)
            for n, p in params.items():
                out[f"adam_param_{n}_{k}"] = p.detach().numpy().copy()
        # activations: GaussianModel.get_scaling / get_rotation / get_opacity (gaussian_model.py:106-127) + autograd
        ga = torch.Generator().manual_seed(7)
        Pa = 37
        gm2 = GaussianModel(3)
        raw = {"scaling": torch.randn(Pa, 3, generator=ga) - 2.0, "rotation": torch.randn(Pa, 4, generator=ga) * 3.0,
               "opacity": torch.randn(Pa, 1, generator=ga) * 3.0}
        raw["rotation"][5] = 0.0                       # the clamp_min(1e-12) branch of normalize
        raw["rotation"][6] = torch.tensor([1e-3, 0.0, 0.0, 0.0])
        gm2._scaling = torch.nn.Parameter(raw["scaling"].clone())
        gm2._rotation = torch.nn.Parameter(raw["rotation"].clone())
        gm2._opacity = torch.nn.Parameter(raw["opacity"].clone())
        outs = {"scaling": gm2.get_scaling, "rotation": gm2.get_rotation, "opacity": gm2.get_opacity}
        ups = {k: torch.randn(v.shape, generator=ga) for k, v in outs.items()}
        sum((outs[k] * ups[k]).sum() for k in outs).backward()
        for k, prm in (("scaling", gm2._scaling), ("rotation", gm2._rotation), ("opacity", gm2._opacity)):
            out[f"act_raw_{k}"] = raw[k].numpy()
            out[f"act_out_{k}"] = outs[k].detach().numpy()
            out[f"act_up_{k}"] = ups[k].numpy()
            out[f"act_grad_{k}"] = prm.grad.numpy().copy()
        out["adam_iters"] = np.array(iters, np.int32)
        out["adam_lrs"] = np.array(lrs, np.float64)
        out["adam_group_names"] = np.array([grp["name"] for grp in gm.optimizer.param_groups])
        for n, p in params.items():
            st = gm.optimizer.state[p]
            out[f"adam_init_{n}"] = init[n].numpy()
            out[f"adam_m_{n}"] = st["exp_avg"].numpy().copy()
            out[f"adam_v_{n}"] = st["exp_avg_sq"].numpy().copy()
        out["adam_eps"] = np.float64(gm.optimizer.defaults["eps"])
        out["adam_betas"] = np.array(gm.optimizer.defaults["betas"], np.float64)
        out["adam_spatial_lr_scale"] = np.float64(gm.spatial_lr_scale)
    finally:
        torch.zeros = _zeros


def main():
    out = {}
    loss_cases(out)
    adam_case(out)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "next_steps.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays; torch", torch.__version__)


if __name__ == "__main__":
    main()
