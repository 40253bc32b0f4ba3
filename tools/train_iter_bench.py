#!/usr/bin/env python
"""One full training iteration around the operator, the way /root/reference/train.py:86-141 runs it
(activations of gaussian_model.py:106-127, render() of gaussian_renderer/__init__.py:20-112, loss,
backward, densification statistics, optimizer step), at BASELINE configs[1] size on one GPU, in two arms:

  fused : gaussian_training.activate + training_loss + b200splat_densify_stats + FusedAdam, SH coefficients
          in one [P,M,3] parameter (this repo's 8f pieces)
  torch : the reference's own Python around the same operator -- loss_utils restated with stock torch ops,
          boolean-mask densification statistics, stock torch.optim.Adam(eps=1e-15)

Both arms use this repo's rasterizer (the reference's is absent).  CUDA events on the launch stream; a
different camera every iteration (orbit), inputs far larger than L2.  JSON on stdout and in
gpurun_out/train_iter.json."""
import argparse
import importlib
import json
import math
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
importlib.import_module("3dgs_study_b200")
synth = importlib.import_module("3dgs_study_b200.synth")
import diff_gaussian_rasterization as dgr  # noqa: E402
import gaussian_training as gtr  # noqa: E402
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, _C  # noqa: E402



def torch_ssim(img1, img2, w2):
    C = 3
    x, y = img1[None], img2[None]
    mu1, mu2 = F.conv2d(x, w2, padding=5, groups=C), F.conv2d(y, w2, padding=5, groups=C)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = F.conv2d(x * x, w2, padding=5, groups=C) - mu1_sq
    s2 = F.conv2d(y * y, w2, padding=5, groups=C) - mu2_sq
    s12 = F.conv2d(x * y, w2, padding=5, groups=C) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu1_mu2 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))).mean()


class Model:
    """Raw parameters + activations as GaussianModel keeps them (scene/gaussian_model.py:36-50,106-127)."""

    def __init__(self, sc, dev, fused):
        P = sc.means3D.shape[0]
        self._xyz = torch.nn.Parameter(sc.means3D.to(dev))
        self.fused = fused
        if fused:
            # ONE [P, M, 3] SH parameter (FusedAdam trains its f_dc / f_rest columns at their own rates): no torch.cat
            self._features = torch.nn.Parameter(sc.shs.contiguous().to(dev))
        else:
            self._features_dc = torch.nn.Parameter(sc.shs[:, :1, :].contiguous().to(dev))
            self._features_rest = torch.nn.Parameter(sc.shs[:, 1:, :].contiguous().to(dev))
        self._scaling = torch.nn.Parameter(torch.log(sc.scales).to(dev))
        self._rotation = torch.nn.Parameter(sc.rotations.to(dev))
        op = sc.opacities.clamp(1e-6, 1 - 1e-6)
        self._opacity = torch.nn.Parameter(torch.log(op / (1 - op)).to(dev))
        self.xyz_gradient_accum = torch.zeros(P, 1, device=dev)
        self.denom = torch.zeros(P, 1, device=dev)
        self.max_radii2D = torch.zeros(P, device=dev)
        M = sc.shs.shape[1]
        feats = ([{"params": [self._features], "lr": 0.0025, "lr_tail": 0.0025 / 20.0, "row_width": 3 * M, "row_split": 3,
                   "name": "features"}] if fused else
                 [{"params": [self._features_dc], "lr": 0.0025, "name": "f_dc"},
                  {"params": [self._features_rest], "lr": 0.0025 / 20.0, "name": "f_rest"}])
        groups = [{"params": [self._xyz], "lr": 0.00016, "name": "xyz"}] + feats + [
                  {"params": [self._opacity], "lr": 0.05, "name": "opacity"},
                  {"params": [self._scaling], "lr": 0.005, "name": "scaling"},
                  {"params": [self._rotation], "lr": 0.001, "name": "rotation"}]
        self.optimizer = (gtr.FusedAdam if fused else torch.optim.Adam)(groups, lr=0.0, eps=1e-15)
        self.sched = gtr.expon_lr(0.00016, 0.0000016, lr_delay_mult=0.01, max_steps=30000)


def iteration(m, cam, gt, bg, it, fused, w2):
    for grp in m.optimizer.param_groups:
        if grp["name"] == "xyz":
            grp["lr"] = m.sched(it)
    screen = torch.zeros_like(m._xyz, requires_grad=True)
    rs = GaussianRasterizationSettings(image_height=cam.image_height, image_width=cam.image_width, tanfovx=cam.tanfovx,
                                       tanfovy=cam.tanfovy, bg=bg, scale_modifier=1.0, viewmatrix=cam.world_view_transform,
                                       projmatrix=cam.full_proj_transform, sh_degree=3, campos=cam.camera_center,
                                       prefiltered=False, debug=False)
    if fused:
        shs = m._features
        scales, rotations, opacities = gtr.activate(m._scaling, m._rotation, m._opacity)
    else:
        shs = torch.cat((m._features_dc, m._features_rest), dim=1)
        scales, rotations, opacities = torch.exp(m._scaling), F.normalize(m._rotation), torch.sigmoid(m._opacity)
    image, radii = GaussianRasterizer(rs)(means3D=m._xyz, means2D=screen, shs=shs, colors_precomp=None, opacities=opacities,
                                          scales=scales, rotations=rotations, cov3D_precomp=None)
    if fused:
        loss, _ = gtr.training_loss(image, gt, 0.2)
    else:
        loss = 0.8 * torch.abs(image - gt).mean() + 0.2 * (1.0 - torch_ssim(image, gt, w2))
    loss.backward()
    with torch.no_grad():
        if fused:
            _C.densify_stats(screen.grad, radii, m.xyz_gradient_accum, m.denom, m.max_radii2D)
        else:
            vis = radii > 0
            m.max_radii2D[vis] = torch.max(m.max_radii2D[vis], radii[vis].float())
            m.xyz_gradient_accum[vis] += torch.norm(screen.grad[vis, :2], dim=-1, keepdim=True)
            m.denom[vis] += 1
        m.optimizer.step()
        m.optimizer.zero_grad(set_to_none=True)
    return loss


def measure(gaussians=3_000_000, width=1920, height=1080, iters=24, warmup=4, arms=("fused", "torch")):
    """Time the two arms; returns the result dict (errors of an arm are reported in place, never raised)."""
    legacy = dgr.LEGACY_OUTPUTS
    dgr.LEGACY_OUTPUTS = True     # the reference call site unpacks (color, radii)
    try:
        return _measure(gaussians, width, height, iters, warmup, arms)
    finally:
        dgr.LEGACY_OUTPUTS = legacy


def _measure(gaussians, width, height, iters, warmup, arms):
    dev = torch.device("cuda:0")
    sc = synth.make_scene(gaussians, 3, seed=0)
    cams = [synth.make_camera(width, height, yaw=2 * math.pi * k / 8).to(dev) for k in range(8)]
    g = torch.Generator().manual_seed(5)
    gt = torch.rand(3, height, width, generator=g).to(dev)
    bg = torch.zeros(3, device=dev)
    taps = torch.tensor(list(gtr.loss_utils.window_taps(11)), device=dev)
    w2 = (taps[:, None] * taps[None, :]).expand(3, 1, 11, 11).contiguous()
    res = {"workload": f"{gaussians} Gaussians SH3, {width}x{height}, one view per iteration, 8-camera orbit: activations, "
                       "render, L1+SSIM loss, backward, densification statistics, Adam step",
           "iters": iters, "warmup": warmup}
    for name, fused in (("fused", True), ("torch", False)):
        if name not in arms:
            continue
        try:
            m = Model(sc, dev, fused)
            for i in range(warmup):
                iteration(m, cams[i % 8], gt, bg, i + 1, fused, w2)
            torch.cuda.synchronize()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for i in range(iters):
                loss = iteration(m, cams[i % 8], gt, bg, warmup + i + 1, fused, w2)
            e.record()
            torch.cuda.synchronize()
            ms = s.elapsed_time(e) / iters
            res[name] = {"ms_per_iteration": round(ms, 4), "iterations_per_s": round(1000.0 / ms, 2), "last_loss": float(loss.item())}
            del m
            torch.cuda.empty_cache()
        except Exception as ex:  # report and keep going: the other arm is still worth having
            res[name] = {"error": repr(ex)}
    if "ms_per_iteration" in res.get("fused", {}) and "ms_per_iteration" in res.get("torch", {}):
        res["speedup_fused_vs_torch"] = round(res["torch"]["ms_per_iteration"] / res["fused"]["ms_per_iteration"], 3)
    res["arms"] = {"fused": "gaussian_training.activate + training_loss + b200splat_densify_stats + FusedAdam, SH in one parameter",
                   "torch": "the reference's own torch code around the same operator (cat/exp/sigmoid/normalize, conv2d SSIM, "
                            "masked-index statistics, stock torch.optim.Adam)"}
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--iters", type=int, default=24)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--arms", default="fused,torch")
    a = ap.parse_args()
    res = measure(a.gaussians, a.width, a.height, a.iters, a.warmup, tuple(a.arms.split(",")))
    print(json.dumps(res))
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        json.dump(res, open(os.path.join(out, "train_iter.json"), "w"))


if __name__ == "__main__":
    main()
