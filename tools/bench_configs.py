#!/usr/bin/env python
"""BASELINE.json configs[2] (densification sweep, per-stage times at 1600x1060) and configs[4] (render-only 4K,
6 M Gaussians SH3, antialiasing + inverse depth, forward fps) on one GPU.  Prints one JSON object per line."""
import importlib, json, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3dgs_study_b200"); pkg.ensure_on_path()
synth = importlib.import_module("3dgs_study_b200.synth")
vp = importlib.import_module("3dgs_study_b200.view_parallel")
import diff_gaussian_rasterization as dgr
C = dgr._C
dev = torch.device("cuda:0")
which = sys.argv[1] if len(sys.argv) > 1 else "all"
SIZES = tuple(int(x) for x in sys.argv[2].split(",")) if len(sys.argv) > 2 and sys.argv[2] != "all" else (
    100_000, 200_000, 500_000, 1_000_000, 2_000_000, 3_000_000, 4_000_000, 6_000_000)
PRESET = sys.argv[3] if len(sys.argv) > 3 else "opaque"

def sweep():
    """configs[2]: every size of the densification sweep, a different camera of an 8-view orbit every step."""
    W, H, D = 1600, 1060, 3
    for P in SIZES:
        sc = synth.make_scene(P, D, seed=0, preset=PRESET).to(dev)
        cams = [synth.make_camera(W, H, yaw=2.0 * np.pi * k / 8).to(dev) for k in range(8)]
        dL = synth.make_grad_image(W, H).to(dev)
        st = vp.ViewParallelStep(vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations), D, torch.zeros(3, device=dev),
                                 exchange="packed")
        stages, stats = st.stage_profile(cams[0], dL, iters=5)
        C.reset_capacity_policy()
        g = vp.GraphedStep(st, [cams[0]], [dL])          # the step bench.py times: packed rows, one graph replay per view
        for k in range(8):                               # every view of the orbit once: the capacity policy sees the largest R
            g.step([cams[k % 8]], [dL])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(24):
            g.step([cams[k % 8]], [dL])
        e1.record(); torch.cuda.synchronize()
        del g
        binning = sum(stages.get(k, 0.0) for k in ("scan", "duplicate", "sort", "ranges"))
        print(json.dumps({"config": "densification sweep 1600x1060 SH3 fwd+bwd, 8-camera orbit", "preset": PRESET, "step": "packed rows, CUDA-graph replay", "gaussians": P,
                          "visible": stats["visible"], "R": stats["R"], "touched": stats.get("touched"),
                          "ms_per_view": round(e0.elapsed_time(e1) / 24, 4),
                          "stages_ms": {"preprocess": round(stages.get("preprocess", 0.0), 4), "key_sort(scan+duplicate+sort+ranges)": round(binning, 4),
                                        "render_fwd": round(stages.get("render", 0.0), 4), "render_bwd": round(stages.get("render_bwd", 0.0), 4),
                                        "per_gaussian_bwd": round(stages.get("preprocess_bwd", 0.0), 4)}}), flush=True)
        del st, sc
        torch.cuda.empty_cache()

def render4k():
    W, H, D, P = 3840, 2160, 3, 6_000_000
    sc = synth.make_scene(P, D, seed=0).to(dev)
    cams = [synth.make_camera(W, H, yaw=2.0 * np.pi * k / 8).to(dev) for k in range(8)]
    bg = torch.zeros(3, device=dev)
    def fwd(cam):
        return C.rasterize_gaussians(bg, sc.means3D, None, sc.opacities, sc.scales, sc.rotations, 1.0, None, cam.world_view_transform,
                                     cam.full_proj_transform, cam.tanfovx, cam.tanfovy, H, W, sc.shs, D, cam.camera_center, False, False,
                                     antialiasing=True, want_invdepth=True)
    with torch.no_grad():
        for k in range(5):
            r = fwd(cams[k % 8])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(40):
            r = fwd(cams[k % 8])
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 40
        st = C.StageTimes(); r = C.rasterize_gaussians(bg, sc.means3D, None, sc.opacities, sc.scales, sc.rotations, 1.0, None, cams[0].world_view_transform,
                                     cams[0].full_proj_transform, cams[0].tanfovx, cams[0].tanfovy, H, W, sc.shs, D, cams[0].camera_center, False, False,
                                     antialiasing=True, want_invdepth=True, stage_times=st)
    print(json.dumps({"config": "render-only 3840x2160, 6M Gaussians SH3, antialiasing + inverse depth, forward", "fps": round(1000.0 / ms, 2),
                      "ms_per_frame": round(ms, 4), "R": C.forward_status(r[5], H, W)[0], "visible": int((r[2] > 0).sum()),
                      "stages_ms": {k: round(v, 4) for k, v in st.as_dict().items() if "bwd" not in k}}), flush=True)

if which in ("all", "sweep"):
    sweep()
if which in ("all", "4k"):
    render4k()
