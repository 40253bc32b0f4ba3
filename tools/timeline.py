#!/usr/bin/env python
"""GPU timeline of a few bench steps (torch.profiler / CUPTI): every kernel and memcpy in time order with
the idle gap before it.  Diagnostic only; numbers taken under the profiler are never bench values."""
import importlib, json, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from torch.profiler import profile, ProfilerActivity

def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 3_000_000
    pkg = importlib.import_module("3dgs_study_b200"); pkg.ensure_on_path()
    synth = importlib.import_module("3dgs_study_b200.synth")
    vp = importlib.import_module("3dgs_study_b200.view_parallel")
    dev = torch.device("cuda:0")
    W, H, D = 1920, 1080, 3
    sc = synth.make_scene(P, D, seed=0).to(dev)
    cams = [synth.make_camera(W, H, yaw=2.0 * np.pi * k / 8).to(dev) for k in range(8)]
    dL = synth.make_grad_image(W, H).to(dev)
    st = vp.ViewParallelStep(vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations), D, torch.zeros(3, device=dev))
    for i in range(5):
        st.step([cams[i % 8]], [dL])
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for i in range(steps):
            st.step([cams[i % 8]], [dL])
        torch.cuda.synchronize()
    out = os.path.join(ROOT, "gpurun_out", "trace.json")
    prof.export_chrome_trace(out)
    ev = [e for e in json.load(open(out))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    ev.sort(key=lambda e: e["ts"])
    t0 = ev[0]["ts"]; prev_end = t0; busy = 0.0
    lines = []
    for e in ev:
        gap = e["ts"] - prev_end
        lines.append(f"{e['ts'] - t0:10.1f} us  gap {gap:8.1f}  dur {e['dur']:8.1f}  {e['name'][:90]}")
        busy += e["dur"]; prev_end = max(prev_end, e["ts"] + e["dur"])
    total = prev_end - t0
    lines.append(f"total {total:.1f} us over {steps} steps ({total / steps:.1f} us/step), busy {busy:.1f} us ({100 * busy / total:.1f}%)")
    open(os.path.join(ROOT, "gpurun_out", "timeline.txt"), "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[-1:]))
    os.remove(out)

if __name__ == "__main__":
    main()
