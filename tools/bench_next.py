#!/usr/bin/env python
"""Times the steps either side of the rasterizer on one GPU at BASELINE configs[1] sizes
(3 M Gaussians SH3, 1920x1080) with CUDA events on the launch stream, next to what the reference runs on the
same box: its loss and optimizer are pure torch (utils/loss_utils.py, torch.optim.Adam -- restated / called
here with stock torch), its distCUDA2 is oracle/_ref/libsimple_knn_ref.so when that was built.
One JSON line per step on stdout and in gpurun_out/bench_next.jsonl (when that directory exists).
Inputs exceed L2 (126 MB) for Adam; the loss working set (~0.3 GB) does too."""
import argparse
import ctypes
import importlib
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
importlib.import_module("3dgs_study_b200")
import gaussian_training as gtr  # noqa: E402
from simple_knn._C import distCUDA2  # noqa: E402


def timed(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def torch_loss(img, gt, w2):
    C = 3
    x, y = img[None], gt[None]
    mu1, mu2 = F.conv2d(x, w2, padding=5, groups=C), F.conv2d(y, w2, padding=5, groups=C)
    s1 = F.conv2d(x * x, w2, padding=5, groups=C) - mu1.pow(2)
    s2 = F.conv2d(y * y, w2, padding=5, groups=C) - mu2.pow(2)
    s12 = F.conv2d(x * y, w2, padding=5, groups=C) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim = (((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))).mean()
    return 0.8 * torch.abs(img - gt).mean() + 0.2 * (1.0 - ssim)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--knn-points", type=int, default=1_000_000)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = None
    for k in ("hbm_copy_gbs_burst", "hbm_gbs_burst", "hbm_gbs", "hbm_copy_gbs"):
        if isinstance(peaks.get(k), (int, float)):
            hbm = float(peaks[k])
            break
    if hbm is None:
        for k, v in peaks.items():
            if "hbm" in k.lower() and isinstance(v, (int, float)):
                hbm = float(v)
                break
    lines = []

    # ---- loss -------------------------------------------------------------------------------------------
    g = torch.Generator().manual_seed(0)
    gt = torch.rand(3, 1080, 1920, generator=g).to(dev)
    img = (gt + 0.1 * torch.randn(3, 1080, 1920, generator=g).to(dev)).clamp(0, 1).requires_grad_()

    def ours():
        img.grad = None
        loss, _ = gtr.training_loss(img, gt, 0.2)
        loss.backward()

    taps = torch.tensor(list(gtr.loss_utils.window_taps(11)), device=dev)
    w2 = (taps[:, None] * taps[None, :]).expand(3, 1, 11, 11).contiguous()

    def ref():
        img.grad = None
        torch_loss(img, gt, w2).backward()

    t_ours, t_ref = timed(ours, a.iters), timed(ref, a.iters)
    n = img.numel()
    # the two kernels alone, through the C ABI (no autograd bookkeeping): forward(+finalize) and backward
    from gaussian_training import _lib as nl
    L = nl.lib()
    wt = gtr.loss_utils.window_taps(11)
    ai, bi = img.detach().contiguous(), gt.contiguous()
    maps = torch.empty(3, 3, 1080, 1920, device=dev)
    parts = torch.empty(L.b200splat_l1_ssim_partials(3, 1080, 1920), 2, device=dev)
    out3, up, da = torch.empty(3, device=dev), torch.ones(3, device=dev), torch.empty_like(ai)
    sp = nl.stream_ptr(dev)
    k_fwd = timed(lambda: L.b200splat_training_loss_forward(3, 1080, 1920, ai.data_ptr(), bi.data_ptr(), ctypes.addressof(wt), 0.2,
                                                            maps.data_ptr(), parts.data_ptr(), out3.data_ptr(), sp), 50)
    k_bwd = timed(lambda: L.b200splat_training_loss_backward(3, 1080, 1920, ai.data_ptr(), bi.data_ptr(), ctypes.addressof(wt), 0.2,
                                                             maps.data_ptr(), up.data_ptr(), da.data_ptr(), sp), 50)
    algo = 44 * n     # fwd 8 in + 12 out, bwd 20 in + 4 out bytes per pixel and plane
    lines.append({"step": "l1_ssim_loss fwd+bwd", "workload": "3x1080x1920 fp32, lambda 0.2", "ms": t_ours,
                  "torch_reference_ms": t_ref, "speedup_vs_torch": t_ref / t_ours, "algo_bytes": algo,
                  "achieved_gbs": algo / t_ours / 1e6, "hbm_peak_gbs": hbm, "frac": (algo / t_ours / 1e6 / hbm) if hbm else None,
                  "kernels_fwd_ms": k_fwd, "kernels_bwd_ms": k_bwd, "kernels_gbs": algo / (k_fwd + k_bwd) / 1e6,
                  "kernels_frac": (algo / (k_fwd + k_bwd) / 1e6 / hbm) if hbm else None,
                  "note": "ms = through autograd (includes its bookkeeping; host-launch bound when run back to back); "
                          "kernels_* = the C-ABI calls alone, back to back (working set 0.3 GB > L2)"})
    del img, gt, maps

    # ---- Adam -------------------------------------------------------------------------------------------
    P = a.gaussians
    shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}
    lrs = {"xyz": 0.00016, "f_dc": 0.0025, "f_rest": 0.0025 / 20, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}

    def make(cls):
        ps = {k: torch.nn.Parameter(torch.randn(*s, device=dev)) for k, s in shapes.items()}
        opt = cls([{"params": [p], "lr": lrs[k], "name": k} for k, p in ps.items()], lr=0.0, eps=1e-15)
        for p in ps.values():
            p.grad = torch.randn_like(p) * 1e-3
        return ps, opt

    ps, opt = make(gtr.FusedAdam)
    t_ours = timed(opt.step, max(5, a.iters // 2))
    del ps, opt
    torch.cuda.empty_cache()
    ps, opt = make(torch.optim.Adam)
    t_ref = timed(opt.step, max(5, a.iters // 2))
    del ps, opt
    torch.cuda.empty_cache()
    nparam = 59 * P
    algo = 28 * nparam   # p, g, m, v read; p, m, v written
    lines.append({"step": "adam step (6 groups, 1 launch)", "workload": f"{P} Gaussians SH3 = {nparam} fp32 parameters",
                  "ms": t_ours, "torch_reference_ms": t_ref, "speedup_vs_torch": t_ref / t_ours, "algo_bytes": algo,
                  "achieved_gbs": algo / t_ours / 1e6, "hbm_peak_gbs": hbm, "frac": (algo / t_ours / 1e6 / hbm) if hbm else None,
                  "note": "reference = stock torch.optim.Adam(eps=1e-15) with the same groups (its default foreach path)"})

    # ---- distCUDA2 --------------------------------------------------------------------------------------
    synth = importlib.import_module("3dgs_study_b200.synth")
    Pk = a.knn_points
    pts = synth.make_scene(Pk, 0, seed=0).means3D.to(dev).contiguous()
    t_ours = timed(lambda: distCUDA2(pts), 3, warmup=1)
    line = {"step": "distCUDA2 (mean sq. distance to 3 nearest)", "workload": f"{Pk} points of the synthetic scene", "ms": t_ours}
    path = os.path.join(ROOT, "oracle", "_ref", "libsimple_knn_ref.so")
    if os.path.exists(path):
        rl = ctypes.CDLL(path)
        rl.ref_dist_cuda2.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        out = torch.zeros(Pk, device=dev)
        t_ref = timed(lambda: rl.ref_dist_cuda2(Pk, pts.data_ptr(), out.data_ptr()), 2, warmup=1)
        same = bool(torch.equal(out, distCUDA2(pts)))
        line.update({"reference_ms": t_ref, "speedup_vs_reference": t_ref / t_ours, "bit_identical_to_reference": same,
                     "note": "reference = the repo's own simple_knn.cu compiled for sm_100a (oracle/_ref)"})
    lines.append(line)

    outdir = os.path.join(ROOT, "gpurun_out")
    for l in lines:
        print(json.dumps(l))
    if os.path.isdir(outdir):
        with open(os.path.join(outdir, "bench_next.jsonl"), "w") as f:
            for l in lines:
                f.write(json.dumps(l) + "\n")


if __name__ == "__main__":
    main()
