#!/usr/bin/env python
"""bench.py -- fwd+bwd views/s of the differentiable Gaussian-splat rasterizer (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" = one pass of the hot path (forward + backward) over one synthetic camera view per rank of the
Garden-scale scene BASELINE.json configs[1] names: 3 M Gaussians, SH degree 3, 1920x1080.  With N > 1
(torchrun, one rank per GPU) the Gaussians are replicated, every rank renders its own view each step
(weak scaling) and the per-Gaussian parameter gradients are sum-allreduced with NCCL (configs[3]).

One JSON line on stdout (rank 0).  `value` = whole-job views/s with inputs resident in HBM;
`e2e` = the same through the drop-in operator with the per-view inputs (camera matrices, upstream image
gradient) coming from pinned HOST memory and the step's loss read back to the host inside the timed region
(flavours with a gradient checksum and with the whole rendered image are reported beside it); `roofline` = dominant kernel against the measured HBM peak;
`cpu_baseline` = the C oracle (all host threads) on a bounded sample of the same view.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch

METRIC = "fwd+bwd views/s at 3M Gaussians SH3 1080p"
UNIT = "views/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--sh-degree", type=int, default=3)
    ap.add_argument("--preset", default="opaque", choices=["opaque", "translucent"],
                    help="opacity distribution of the synthetic scene (3dgs_study_b200/synth.py PRESETS); the headline is 'opaque'")
    ap.add_argument("--exchange", default=None, choices=["allreduce", "factorized", "packed", "auto"],
                    help="N>1 gradient exchange (default auto: packed rows over NVLink peer memory unless the views touch too many "
                         "Gaussians, then the factorized NCCL exchange); allreduce = plain NCCL all-reduce of 59 f32/Gaussian")
    ap.add_argument("--peer-memory", action="store_true", help="factorized exchange through NVLink peer mappings (fused rebuild kernel + multimem all-reduce) instead of NCCL")
    ap.add_argument("--no-graph", action="store_true", help="queue every kernel of every step from Python instead of replaying the captured step (CUDA graph)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train-iter", action="store_true", help="skip the full-training-iteration measurement (N=1 only)")
    ap.add_argument("--stages", action="store_true", help="print the per-stage table to stderr")
    return ap.parse_args()


def config_of(a):
    """The `config` object, identical in both arms (the driver compares them key by key)."""
    return {"workload": workload_name(a), "gaussians": a.gaussians, "sh_degree": a.sh_degree, "width": a.width, "height": a.height,
            "views_per_rank_per_step": 1,
            "l2": "no flush needed: the inputs of a step (0.7 GB of Gaussians + 0.2 GB of per-view keys) exceed the 126 MB L2, and every step renders a different view of the orbit"}


def workload_name(a):
    extra = "" if a.preset == "opaque" else f", opacity preset '{a.preset}'"
    return f"synthetic scene seed 0: {a.gaussians} Gaussians SH{a.sh_degree}, {a.width}x{a.height}, one view per rank per step, fwd+bwd{extra}"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler:
    """Samples SM clock / throttle reasons DURING the timed region (B200_PROFILING.md "clocks line").
    In-process NVML (nvidia_ml_py) on a 25 ms period: spawning nvidia-smi next to a 3 ms step perturbs the
    step itself (driver lock + a competing process); nvidia-smi is only the fallback when NVML is missing."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self._stop, self._th = index, threading.Event(), None
        self.sm, self.mx, self.reasons, self.source = [], [], set(), None
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self._nv, self._h, self.source = pynvml, pynvml.nvmlDeviceGetHandleByIndex(phys), "nvml"
        except Exception:
            self.source = "nvidia-smi"

    def _sample_nvml(self):
        nv, h = self._nv, self._h
        self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
        self.mx.append(float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)))
        r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h)) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
            else int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
        for name, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40)):
            if r & bit:
                self.reasons.add(name)

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout.strip()
        f = [x.strip() for x in out.split(",")]
        self.sm.append(float(f[0])); self.mx.append(float(f[1]))
        for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[3:7]):
            if v.lower().startswith("active"):
                self.reasons.add(name)

    def _run(self):
        period = 0.025 if self._h is not None else 0.5
        while not self._stop.is_set():
            try:
                self._sample_nvml() if self._h is not None else self._sample_smi()
            except Exception:
                pass
            self._stop.wait(period)

    def __enter__(self):
        if os.environ.get("BENCH_NO_CLOCKS"):
            return self
        self._th = threading.Thread(target=self._run, daemon=True)
        self._th.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._th is not None:
            self._th.join(timeout=10)

    def summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": self.source}


def dist_setup(n):
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        torch.cuda.set_device(local)
        # NCCL on a HIGH-PRIORITY stream: the exchange's all-gather is launched while the per-Gaussian backward still
        # fills every SM; at equal priority its CTAs would only be scheduled once that kernel has drained
        opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local), pg_options=opts)
    return rank, world, local


# -----------------------------------------------------------------------------------------------------
# our arm
# -----------------------------------------------------------------------------------------------------
def run_ours(a):
    import torch.distributed as dist
    rank, world, local = dist_setup(a.gpus)
    pkg = importlib.import_module("3dgs_study_b200")
    pkg.ensure_on_path()
    synth = importlib.import_module("3dgs_study_b200.synth")
    vp = importlib.import_module("3dgs_study_b200.view_parallel")
    import diff_gaussian_rasterization as dgr
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    P, W, H, D = a.gaussians, a.width, a.height, a.sh_degree
    scene_cpu = synth.make_scene(P, D, seed=0, preset=a.preset)
    scene = scene_cpu.to(dev)
    M = scene.shs.shape[1]
    n_views = 8
    cams_cpu = [synth.make_camera(W, H, yaw=2.0 * np.pi * k / n_views) for k in range(n_views)]
    cams = [c.to(dev) for c in cams_cpu]
    bg = torch.zeros(3, device=dev)
    dL_cpu = synth.make_grad_image(W, H).pin_memory()
    dL = dL_cpu.to(dev)
    params = vp.GaussianParams(scene.means3D, scene.shs, scene.opacities, scene.scales, scene.rotations)
    exchange = a.exchange
    if exchange is None and world == 1 and not a.no_graph:
        exchange = "packed"          # the row-producing backward + local accumulate: the same step the ranks run at N > 1
    stepper = vp.ViewParallelStep(params, D, bg, world_size=world, exchange=exchange, peer_memory="try" if a.peer_memory else False)

    def view_of(step):   # rank r renders view (step * world + r) of the orbit
        return (step * world + rank) % n_views

    # The step is captured once in a CUDA graph and replayed (view_parallel.GraphedStep): one launch per step instead
    # of ~35 kernels + a dozen tensor ops queued from Python.  Only the packed-row step is capturable (no collective, no
    # host wait); "auto" first settles on a mode with two eager steps.
    gstep = None
    if not a.no_graph:
        for i in range(2):
            stepper.step([cams[view_of(i)]], [dL])
        torch.cuda.synchronize()
        if stepper.exchange == "packed":
            gstep = vp.GraphedStep(stepper, [cams[view_of(0)]], [dL])

    # ---------------- value: inputs resident in HBM -----------------------------------------------
    def step_resident(i):
        if gstep is not None:
            gstep.step([cams[view_of(i)]])
        else:
            stepper.step([cams[view_of(i)]], [dL])

    for i in range(a.warmup):
        step_resident(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    step_evs = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps + 1)] if os.environ.get("BENCH_STEP_TIMES") else None
    with ClockSampler(local) as cs:
        torch.cuda.synchronize()
        ev0.record()
        for i in range(a.steps):
            if step_evs:
                step_evs[i].record()
            step_resident(a.warmup + i)
        ev1.record()
        if step_evs:
            step_evs[a.steps].record()
        torch.cuda.synchronize()
    ms_total = ev0.elapsed_time(ev1)
    if step_evs and rank == 0:      # diagnostic: device time of every timed step
        print("step ms: " + " ".join(f"{step_evs[i].elapsed_time(step_evs[i + 1]):.2f}" for i in range(a.steps)), file=sys.stderr)
    if world > 1:
        t = torch.tensor([ms_total], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_per_step = ms_total / a.steps
    value = world * 1000.0 / ms_per_step
    clocks = cs.summary()

    # ---------------- e2e: per-view inputs from pinned host memory, results read back ---------------
    cam_host = [torch.cat([c.world_view_transform.reshape(-1), c.full_proj_transform.reshape(-1), c.camera_center.reshape(-1)]).pin_memory()
                for c in cams_cpu]
    img_hosts = [torch.empty(3, H, W).pin_memory() for _ in range(2)]
    chk_hosts = [torch.empty(1).pin_memory() for _ in range(2)]
    h2d = cam_host[0].numel() * 4 + dL_cpu.numel() * 4

    # Software pipeline over two streams: the copy stream brings step i+1's inputs from pinned host memory while
    # step i computes.  Every byte of every step still crosses PCIe inside the timed region.  The host consumes step
    # i's result right after it has queued step i+1, as a training loop that logs its loss one step late does, so the
    # device never waits for the host to queue work.  The result read back every step is the LOSS: <image, dL/dimage>,
    # the scalar whose gradient the step back-propagates -- what the reference's loop reads per iteration (loss.item(),
    # /root/reference/train.py:113).  scalar="grad_checksum" reads a checksum of the parameter gradients instead (the
    # definition of rounds 1 and 2 so far: two strided reductions over 0.3 GB that no training loop performs);
    # `readback_image=True` additionally takes the whole rendered image to the host every step.  Both are reported
    # beside the headline (e2e_with_gradient_checksum, e2e_with_image_readback).
    copy_stream = torch.cuda.Stream(device=dev)
    main_stream = torch.cuda.current_stream(dev)
    slots = [{"cm": torch.empty(35, device=dev), "g": torch.empty(3, H, W, device=dev), "ready": torch.cuda.Event()} for _ in range(2)]
    img_done, fwd_done = torch.cuda.Event(), torch.cuda.Event()
    res_main = [torch.cuda.Event(), torch.cuda.Event()]    # step's checksum is in host memory
    res_copy = [torch.cuda.Event(), torch.cuda.Event()]    # step's image is in host memory
    consumed = [0.0]

    def prefetch(i):
        sl = slots[i % 2]
        with torch.cuda.stream(copy_stream):
            sl["cm"].copy_(cam_host[view_of(i)], non_blocking=True)
            sl["g"].copy_(dL_cpu, non_blocking=True)
            sl["ready"].record(copy_stream)

    def run_e2e(readback_image, scalar="loss"):
        def step_e2e(i, prime=False, consume_prev=True):
            if prime:
                prefetch(i)
            sl = slots[i % 2]
            main_stream.wait_event(sl["ready"])
            copy_stream.wait_event(img_done)              # slot (i+1)%2 was last read by step i-1, which has completed
            prefetch(i + 1)
            cm = sl["cm"]
            cam = cams_cpu[view_of(i)]._replace(world_view_transform=cm[:16].view(4, 4), full_proj_transform=cm[16:32].view(4, 4),
                                                camera_center=cm[32:35])

            def on_forward(img):
                if not readback_image:
                    return
                fwd_done.record(main_stream)
                with torch.cuda.stream(copy_stream):
                    copy_stream.wait_event(fwd_done)
                    img.record_stream(copy_stream)
                    img_hosts[i % 2].copy_(img, non_blocking=True)
                    res_copy[i % 2].record(copy_stream)

            if gstep is not None and not readback_image:   # inputs move from the arrival slot into the graph's static inputs, then one launch
                color = gstep.step([cam], [sl["g"]])
            else:                     # (the image-readback flavour starts its copy between forward and backward: queued from Python)
                color = stepper.step([cam], [sl["g"]], on_forward=on_forward)
            res = torch.dot(color.reshape(-1), sl["g"].reshape(-1)) if scalar == "loss" else stepper.grad_checksum()
            chk_hosts[i % 2].copy_(res.reshape(1), non_blocking=True)
            img_done.record(main_stream)
            res_main[i % 2].record(main_stream)
            if consume_prev:
                consume(i - 1)

        def consume(i):                                    # the host reads step i's results
            res_main[i % 2].synchronize()
            consumed[0] += float(chk_hosts[i % 2][0])
            if readback_image:
                res_copy[i % 2].synchronize()
                consumed[0] += float(img_hosts[i % 2][0, 0, 0])

        for i in range(max(3, a.warmup // 2)):
            step_e2e(i, prime=(i == 0), consume_prev=(i > 0))
        e2e_base = max(3, a.warmup // 2)
        consume(e2e_base - 1)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ev0.record()
        for i in range(a.steps):
            step_e2e(e2e_base + i, consume_prev=(i > 0))
        consume(e2e_base + a.steps - 1)
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    ms_e2e = run_e2e(False)
    e2e_value = world * 1000.0 / (ms_e2e / a.steps)
    d2h = 4
    ms_e2e_chk = run_e2e(False, scalar="grad_checksum")
    e2e_chk = {"value": round(world * 1000.0 / (ms_e2e_chk / a.steps), 3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": 4, "ms_per_step": round(ms_e2e_chk / a.steps, 4),
               "note": "as e2e, but the scalar read back is a checksum of the parameter gradients (two strided reductions over the dense gradients): the e2e definition of the earlier records"}
    ms_e2e_img = run_e2e(True)
    e2e_img = {"value": round(world * 1000.0 / (ms_e2e_img / a.steps), 3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(img_hosts[0].numel() * 4 + 4), "ms_per_step": round(ms_e2e_img / a.steps, 4),
               "note": "as e2e, plus the whole rendered image copied to pinned host memory every step (round 1's definition)"}

    # ---------------- optional: GPU timeline of 3 steps (diagnostic; BENCH_TRACE=1) ----------------
    if os.environ.get("BENCH_TRACE"):
        from torch.profiler import profile, ProfilerActivity
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for i in range(3):
                step_resident(i)
            torch.cuda.synchronize()
        if rank in (0, world - 1):
            tr = os.path.join(ROOT, "gpurun_out", f"trace_n{world}_r{rank}.json")
            prof.export_chrome_trace(tr)
            ev = [e for e in json.load(open(tr))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
            ev.sort(key=lambda e: e["ts"])
            t0 = ev[0]["ts"]
            with open(os.path.join(ROOT, "gpurun_out", f"timeline_n{world}_r{rank}.txt"), "w") as f:
                for e in ev:
                    if e["dur"] >= float(os.environ.get("BENCH_TRACE_MIN_US", "8")):
                        f.write(f"{e['ts'] - t0:10.1f} us  dur {e['dur']:8.1f}  stream {e['args'].get('stream')}  {e['name'][:80]}\n")
            os.remove(tr)
        if world > 1:
            dist.barrier()

    # ---------------- per-stage times + roofline (rank 0, CUDA events inside the library) ----------
    stages, roof, launches, stats = {}, None, None, {}
    if rank == 0:
        stages, stats = stepper.stage_profile(cams[0], dL, iters=max(3, min(a.steps, 10)))
        roof = roofline_from_stages(stages, stats, P, M, W, H)
        launches = stepper.launches_per_view()
    if world > 1:
        dist.barrier()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return None
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": config_of(a),
        "workload_detail": {"visible": stats.get("visible"), "tile_pairs_R": stats.get("R"), "touched_gaussians": stats.get("touched"),
                            "parallelism": f"view-parallel x{world}, Gaussians replicated" + ((", exchange=" + stepper.exchange + ({"factorized": ": NCCL all-gather of [P,3] masked colour grads + all-reduce of 11 f32/Gaussian, dL/dsh rebuilt locally",
                                                                           "packed": ": one 64-byte gradient row per Gaussian a view touched, read from every peer over NVLink by the kernel that adds them into the dense gradients (no collective on the data path)"}
                                                                          .get(stepper.exchange, ": NCCL sum-allreduce of 59 f32/Gaussian grads"))) if world > 1 else "")},
        "e2e": {"value": round(e2e_value, 3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": round(ms_e2e / a.steps, 4),
                "note": "Gaussians are model state resident in HBM (as in the reference); camera + upstream image gradient from pinned host memory and the step's loss <image, dL/dimage> (the scalar a training loop reads, loss.item()) back to the host, every step, inside the timed region; the host reads step i's result after queueing step i+1"},
        "e2e_with_gradient_checksum": e2e_chk,
        "e2e_with_image_readback": e2e_img,
        "gpu_launches": int(launches * a.steps) if launches else None,
        "step_submission": ("one CUDA-graph replay per step (the whole step captured once: view_parallel.GraphedStep)" if gstep is not None
                            else "every kernel queued from Python"),
        "clocks": clocks, "stages_ms": {k: round(v, 4) for k, v in stages.items()},
        "roofline": roof,
    }
    if world > 1:
        dist.destroy_process_group()
    return line


def csrc_sha16():
    """Hash of the kernel sources the loaded library was built from: ncu-derived constants in profiles/ are only
    attached to the bench line when they were captured on the same sources."""
    import glob
    import hashlib
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "3dgs_study_b200", "csrc", "*.cu")) + glob.glob(os.path.join(ROOT, "3dgs_study_b200", "csrc", "*.cuh"))):
        with open(f, "rb") as fh:
            h.update(os.path.basename(f).encode() + b"\0" + fh.read())
    return h.hexdigest()[:16]


def fp32_peaks(peaks):
    """FP32 / MUFU / atomic denominators: measured by tools/peaks.cu on a B200 of this pool (profiles/r02_peaks.json);
    else the nominal figure 148 SMs x 128 lanes x 2 flop x SM clock."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_peaks.json")) as f:
            pk = json.load(f)
        return {"fp32_tflops": max(pk["ffma_tflops"], pk["ffma2_tflops"]), "ffma_tflops": pk["ffma_tflops"], "ffma2_tflops": pk["ffma2_tflops"],
                "mufu_ex2_gops": pk["mufu_ex2_gops"], "red_f32_l2_gops": pk["red_f32_l2"]["g_f32_adds_s"], "source": "measured (tools/peaks.cu, profiles/r02_peaks.json)"}
    except Exception:
        mhz = peaks.get("sm_max_mhz", 1965.0) or 1965.0
        return {"fp32_tflops": round(148 * 128 * 2 * mhz * 1e6 / 1e12, 1), "source": "nominal 148 x 128 x 2 x SM clock (no measured figure found)"}


def roofline_from_stages(stages, stats, P, M, W, H):
    """Every stage against the roofline that binds it.

    HBM-bound stages (preprocess fwd/bwd, depth order, emit, tile sort, ranges): algorithmic bytes / live stage time
    vs the measured copy bandwidth.  Render forward / backward: FP32-issue bound -- SURVEY 8(d)'s flop counts
    (fwd 12 E + 8 C, bwd 25 E' + 45 C; E pixel evaluations, C blends, both counted on this view by
    b200splat_render_stats) / live stage time vs the measured FP32 FFMA rate; their HBM figure stays as a secondary
    row.  V visible Gaussians, Nt Gaussians the backward touched, R (tile, Gaussian) pairs, Rv pairs the render
    kernels visit, T tiles, N pixels."""
    peaks, which = measured_peaks()
    fpk = fp32_peaks(peaks)
    V, R = stats.get("visible", 0), stats.get("R", 0)
    Rv = stats.get("pairs_visited", R)
    Rc = stats.get("pairs_contributing", 0)
    Nt = stats.get("touched", V)
    E, C = stats.get("pixel_evals", 0), stats.get("contributions", 0)
    T, N = ((W + 15) // 16) * ((H + 15) // 16), W * H
    kb = 2 if T <= 65536 else 4
    algo = {
        "preprocess": 12 * P + V * (32 + 12 * M) + 20 * P + 48 * V,       # means | scale rot opacity SH | radii rect depth-key | record
        "scan": stats.get("depth_sort_bytes", 4 * 16 * P + 4 * P + 16 * P),  # depth order of the visible Gaussians + scan of tiles touched
        "duplicate": 16 * V + (kb + 4) * R,                                 # order rect offsets | tile key + id per pair
        "sort": stats.get("tile_sort_bytes", kb * R + 2 * 2 * (kb + 4) * R),
        "ranges": kb * R + 8 * T + 4 * T,
        "render": 52 * Rv + 20 * N,                                         # id + 48-byte record per visited pair | colour, final T, n_contrib
        "render_bwd": 52 * Rv + 32 * N + 36 * Rc,                           # ... | final T, n_contrib, dL/dpixel | 9 red.f32 per contributing (strip, Gaussian)
        # per-Gaussian backward visits the TOUCHED Gaussians only: inputs + render sums + SH row in, gradient rows out;
        # the zero-fill of the dense outputs (P rows) runs on a side stream during the render backward and is not this stage
        "preprocess_bwd": Nt * (44 + 48 + 12 * M) + Nt * (52 + 12 * M) + 5 * P,
    }
    flops = {"render": 12 * E + 8 * C, "render_bwd": 25 * E + 45 * C}
    rows = {}
    for k, ms in stages.items():
        if ms > 0 and k in algo:
            gbs = algo[k] / (ms * 1e-3) / 1e9
            rows[k] = {"ms": round(ms, 4), "bound": "hbm", "algo_bytes": int(algo[k]), "gbs": round(gbs, 1),
                       "frac_hbm": round(gbs / peaks["hbm_gbs"], 4)}
            if k in flops and flops[k] > 0:
                tf = flops[k] / (ms * 1e-3) / 1e12
                rows[k].update({"bound": "fp32_issue", "algo_flops": int(flops[k]), "tflops": round(tf, 2),
                                "frac_fp32": round(tf / fpk["fp32_tflops"], 4)})
    if not rows:
        return None
    top = max(rows, key=lambda k: rows[k]["ms"])
    r = rows[top]
    if r["bound"] == "fp32_issue":
        out = {"kernel": top, "bound": "fp32_issue", "achieved": r["tflops"], "peak": fpk["fp32_tflops"], "unit": "TFLOP/s",
               "frac": r["frac_fp32"], "traffic": None, "peak_source": fpk["source"],
               "flops_per_launch": r["algo_flops"], "flops_formula": "SURVEY 8(d): fwd 12 E + 8 C, bwd 25 E' + 45 C (E, C counted on this view)",
               "hbm": {"achieved": r["gbs"], "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": r["frac_hbm"], "peak_source": which,
                       "note": "secondary: the render kernels move far less than HBM could deliver"}}
    else:
        out = {"kernel": top, "bound": "hbm", "achieved": r["gbs"], "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": r["frac_hbm"],
               "traffic": None, "peak_source": which}
    out["peaks"] = fpk
    out["per_stage"] = rows
    # ncu-derived figures (one --set full capture per build): attached only when captured on THESE kernel sources
    try:
        with open(os.path.join(ROOT, "profiles", "r02_ncu_kernels.json")) as f:
            prof = json.load(f)
        key = {"render_bwd": "render_bwd_kernel", "render": "render_fwd_kernel", "preprocess": "preprocess_fwd_kernel",
               "preprocess_bwd": "preprocess_bwd_kernel"}.get(top)
        if prof.get("csrc_sha16") != csrc_sha16():
            out["ncu"] = {"dropped": "profiles/r02_ncu_kernels.json was captured on other kernel sources", "profile_sha": prof.get("csrc_sha16"),
                          "build_sha": csrc_sha16()}
        elif key and key in prof:
            out["traffic"] = prof[key]["dram_bytes"]
            out["issue"] = {"warp_inst_per_launch": prof[key]["warp_inst"], "issue_slots_busy_pct_ncu": prof[key]["issue_active_pct"],
                            "achieved_ginst_s": round(prof[key]["warp_inst"] / (rows[top]["ms"] * 1e-3) / 1e9, 1),
                            "peak_ginst_s": round(148 * 4 * (peaks.get("sm_max_mhz", 1965.0) or 1965.0) * 1e6 / 1e9, 1),
                            "source": prof.get("source", "profiles/r02_ncu_kernels.json")}
            out["issue"]["frac"] = round(out["issue"]["achieved_ginst_s"] / out["issue"]["peak_ginst_s"], 4)
    except Exception:
        pass
    if C:
        out["work"] = {"pairs_visited": Rv, "pixel_evals": E, "blends": C, "lane_efficiency": round(C / max(E, 1), 4),
                       "strip_pairs_contributing": Rc, "touched_gaussians": Nt,
                       "alpha_test_disagreements_fwd_vs_bwd": stats.get("alpha_test_disagreements"),
                       "gblends_per_s_fwd": round(C / (rows["render"]["ms"] * 1e-3) / 1e9, 1) if "render" in rows else None,
                       "gblends_per_s_bwd": round(C / (rows["render_bwd"]["ms"] * 1e-3) / 1e9, 1) if "render_bwd" in rows else None}
    return out


# -----------------------------------------------------------------------------------------------------
# CPU legs (the only places bench.py executes oracle/): cpu_baseline of our arm and --impl reference
# -----------------------------------------------------------------------------------------------------
def host_threads():
    """All host threads this process may use (torchrun exports OMP_NUM_THREADS=1 to its workers: the CPU legs set their
    own thread count instead of inheriting that)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return os.cpu_count() or 1


class CpuOracleView:
    """C oracle (oracle/splat_oracle.c, OpenMP over all host threads) on one view of the same workload.  stride = 1
    (the default for every reported number): the WHOLE view, every tile -- a measured wall time, nothing
    extrapolated.  stride > 1 (warm-up only): the tile render forward + backward runs on every stride-th tile."""

    def __init__(self, a):
        from oracle import c_oracle as co
        synth = importlib.import_module("3dgs_study_b200.synth")
        self.co, self.a = co, a
        self.cores = host_threads()
        os.environ["OMP_NUM_THREADS"] = str(self.cores)
        co.lib().orc_set_threads(self.cores)
        torch.set_num_threads(self.cores)
        P, W, H, D = a.gaussians, a.width, a.height, a.sh_degree
        sc = synth.make_scene(P, D, seed=0, preset=a.preset)
        self.M = sc.shs.shape[1]
        self.arr = {k: getattr(sc, k).numpy() for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        self.cams = [synth.make_camera(W, H, yaw=2.0 * np.pi * k / 8) for k in range(8)]
        self.bg = np.zeros(3, np.float32)
        self.dL = synth.make_grad_image(W, H).numpy()

    def step(self, i, stride=1):
        """-> wall seconds of one fwd+bwd view (all stages inside the timed region)."""
        co, a, A = self.co, self.a, self.arr
        P, W, H, D, M, s = a.gaussians, a.width, a.height, a.sh_degree, self.M, max(1, stride)
        cam = self.cams[i % 8]
        v, p, c = cam.world_view_transform.numpy(), cam.full_proj_transform.numpy(), cam.camera_center.numpy()
        t0 = time.perf_counter()
        g = co.preprocess(P, D, M, W, H, A["means3D"], A["scales"], 1.0, A["rotations"], A["opacities"], A["shs"], None, None,
                          v, p, c, cam.tanfovx, cam.tanfovy)
        b = co.bin_tiles(P, W, H, g)
        color, _, fT, nc = co.render(W, H, g, b, self.bg, tile_stride=s, tile_offset=i % s)
        rg = co.render_backward(P, W, H, g, b, self.bg, fT, nc, self.dL, tile_stride=s, tile_offset=i % s)
        co.preprocess_backward(P, D, M, W, H, A["means3D"], A["scales"], 1.0, A["rotations"], A["opacities"], A["shs"], None,
                               v, p, c, cam.tanfovx, cam.tanfovy, False, g, rg)
        return time.perf_counter() - t0

    def sample_desc(self, n):
        a = self.a
        return (f"C oracle port (OpenMP, {self.cores} threads), {n} whole views measured: all {a.gaussians} Gaussians preprocessed, binned, "
                f"sorted; every tile rendered forward and backward; per-Gaussian backward -- wall time per view, nothing extrapolated")


def cpu_baseline_quick(a, budget_s=20.0):
    """cpu_baseline of our arm: 1 cheap warm-up (1/16 of the tiles) + as many WHOLE views as fit in ~budget_s (>= 2)."""
    ov = CpuOracleView(a)
    ov.step(0, stride=16)
    walls, spent, i = [], 0.0, 1
    while (spent < budget_s or len(walls) < 2) and i <= 8:
        w = ov.step(i)
        walls.append(w); spent += w; i += 1
    w = float(np.median(walls))
    return {"value": round(1.0 / w, 5), "unit": UNIT, "cores": ov.cores, "kind": "port",
            "sample": ov.sample_desc(len(walls)), "ms_per_view": round(w * 1e3, 1)}


def torch_oracle_config1():
    """BASELINE configs[0]: 10k Gaussians SH0, 256x256, fwd+bwd through the pure-PyTorch CPU oracle."""
    from oracle import torch_ref as tr
    synth = importlib.import_module("3dgs_study_b200.synth")
    torch.set_num_threads(os.cpu_count() or 1)
    P, W, H = 10000, 256, 256
    sc, cam, dL = synth.make_scene(P, 0, seed=0), synth.make_camera(W, H), synth.make_grad_image(W, H)
    times = []
    for it in range(3):
        lv = {k: getattr(sc, k).clone().requires_grad_() for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        t0 = time.perf_counter()
        color, radii, _, _ = tr.rasterize_gaussians(image_height=H, image_width=W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
                                                    bg=torch.zeros(3), scale_modifier=1.0, viewmatrix=cam.world_view_transform,
                                                    projmatrix=cam.full_proj_transform, sh_degree=0, campos=cam.camera_center,
                                                    means2D=torch.zeros(P, 3, requires_grad=True), **lv)
        torch.autograd.backward([color], [dL])
        times.append(time.perf_counter() - t0)
    return {"workload": "BASELINE configs[0]: 10k Gaussians SH0 256x256 fwd+bwd, pure-PyTorch CPU oracle",
            "views_per_s": round(1.0 / float(np.median(times[1:])), 4), "cores": os.cpu_count()}


def run_reference(a):
    """--impl reference: the reference has no CPU rasterizer and its CUDA rasterizer source is absent from the
    snapshot (unbuildable), so this arm times the oracle port on the host cores (kind "port").  Every step is one
    WHOLE view (all tiles): `ms_per_step` x steps is wall time actually spent.  Runs on rank 0 only and sets its own
    thread count, so the figure does not depend on how it was launched (python or torchrun)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    ov = CpuOracleView(a)
    for i in range(a.warmup):
        ov.step(i, stride=1 if i == a.warmup - 1 else 16)     # warm-up: cheap samples, then one whole view
    t0 = time.perf_counter()
    for i in range(a.steps):
        ov.step(a.warmup + i)
    ms = (time.perf_counter() - t0) * 1e3 / max(a.steps, 1)
    val = 1000.0 / ms
    return {"impl": "reference", "metric": METRIC, "value": round(val, 5), "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": round(ms, 2), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config_of(a),
            "cpu_baseline": {"value": round(val, 5), "unit": UNIT, "cores": ov.cores, "kind": "port", "sample": ov.sample_desc(a.steps)},
            "e2e": {"value": round(val, 5), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference CUDA rasterizer: unavailable on this box (submodule source absent from /root/reference)"}


def main():
    a = parse_args()
    if a.impl == "reference":
        line = run_reference(a)
        if line is not None:
            print(json.dumps(line), flush=True)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl ours) needs a CUDA GPU: the product has no CPU fallback")
    line = run_ours(a)
    if line is None:
        return
    if int(os.environ.get("WORLD_SIZE", "1")) == 1 and not a.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_quick(a)
        try:
            line["cpu_baseline_torch_config1"] = torch_oracle_config1()
        except Exception as e:   # reported, never fatal
            line["cpu_baseline_torch_config1"] = {"error": str(e)}
    if int(os.environ.get("WORLD_SIZE", "1")) == 1 and a.impl == "ours" and a.sh_degree == 3 and not a.no_train_iter:
        # beside the headline metric: one full training iteration around the operator, fused
        # pieces of this repo vs the reference's own torch code around the same operator; reported, never fatal
        try:
            import importlib.util
            spec = importlib.util.spec_from_file_location("train_iter_bench", os.path.join(ROOT, "tools", "train_iter_bench.py"))
            tib = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(tib)
            torch.cuda.empty_cache()
            line["training_iteration"] = tib.measure(a.gaussians, a.width, a.height, iters=24, warmup=4)
        except Exception as e:
            line["training_iteration"] = {"error": repr(e)}
    if a.stages:
        for k, v in (line.get("roofline") or {}).get("per_stage", {}).items():
            print(f"{k:16s} {v['ms']:8.3f} ms  {v['gbs']:8.1f} GB/s  {100 * v['frac_hbm']:5.1f}% of HBM peak", file=sys.stderr)
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
