#!/usr/bin/env python
"""CPU only.  How much does the one deliberate deviation of the canonical arithmetic from the written hypothesis of
the upstream rasterizer (ndc -> pixel in fp32, `((ndc + 1) * S - 1) * 0.5`, three roundings)
change?  The canonical path (oracle and CUDA alike) evaluates the same expression in double and rounds once.
Both variants of the C oracle are run on the parity scenes and on BASELINE configs[1]; counted: Gaussians whose
pixel centre differs in any bit, whose tile rectangle differs, and the change in R and in sorted-key order.

  python tools/count_pixel_formula.py [--full]     -> JSON lines (committed as profiles/r02_pixel_formula_counts.jsonl)
"""
import argparse
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
importlib.import_module("3dgs_study_b200").ensure_on_path()
synth = importlib.import_module("3dgs_study_b200.synth")
from oracle import c_oracle as co  # noqa: E402


def both(P, W, H, D, seed, s_med, aa=False, yaw=0.0):
    sc = synth.make_scene(P, D, seed=seed, s_med=s_med)
    cam = synth.make_camera(W, H, yaw=yaw)
    M = sc.shs.shape[1]
    A = {k: getattr(sc, k).numpy() for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    v, p, c = cam.world_view_transform.numpy(), cam.full_proj_transform.numpy(), cam.camera_center.numpy()
    out = []
    for mode in (0, 1):
        co.lib().orc_set_pixel_fp32(mode)
        g = co.preprocess(P, D, M, W, H, A["means3D"], A["scales"], 1.0, A["rotations"], A["opacities"], A["shs"], None, None,
                          v, p, c, cam.tanfovx, cam.tanfovy, aa)
        out.append(g)
    co.lib().orc_set_pixel_fp32(0)
    g0, g1 = out
    vis = (g0.radii > 0) | (g1.radii > 0)
    px_diff = int(((g0.means2D != g1.means2D).any(axis=1) & vis).sum())
    max_px = float(np.abs(g0.means2D.astype(np.float64) - g1.means2D)[vis].max()) if vis.any() else 0.0
    tiles_diff = int((g0.tiles_touched != g1.tiles_touched).sum())
    radii_diff = int((g0.radii != g1.radii).sum())
    R0, R1 = int(g0.tiles_touched.astype(np.int64).sum()), int(g1.tiles_touched.astype(np.int64).sum())
    return {"P": P, "W": W, "H": H, "visible": int(vis.sum()), "pixel_centre_bits_differ": px_diff, "max_abs_pixel_diff": max_px,
            "radii_differ": radii_diff, "tiles_touched_differ": tiles_diff, "R_canonical": R0, "R_fp32_order": R1}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--full", action="store_true", help="also BASELINE configs[1] (3 M Gaussians, 1920x1080)")
    a = ap.parse_args()
    import test_gpu_parity as t
    for name, kw in t.CASES.items():
        if kw.get("precomp_cov") or kw.get("precomp_color"):
            continue
        r = both(kw["P"], kw["W"], kw["H"], kw.get("D", 3) if kw.get("M") is None else 3, kw.get("seed", 0), kw.get("s_med"),
                 kw.get("antialiasing", False), kw.get("yaw", 0.0))
        print(json.dumps(dict(case=name, **r)), flush=True)
    if a.full:
        for k in (0, 3):
            r = both(3_000_000, 1920, 1080, 3, 0, None, False, 2.0 * np.pi * k / 8)
            print(json.dumps(dict(case=f"config2_view{k}", **r)), flush=True)


if __name__ == "__main__":
    main()
