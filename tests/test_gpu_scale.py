"""GPU parity at BASELINE.json's FULL sizes (configs[1] and configs[4]): the scenes every headline number is
measured on, not toy cases.

The C oracle preprocesses and bins ALL Gaussians of the view (a few seconds on the host), so the integer gates
cover every Gaussian and every (tile, Gaussian) pair: radii, tiles touched, all R sorted 64-bit keys and values,
all tile ranges -- bit-exact.  The per-pixel stages of the oracle are too slow for 2 M pixels x thousands of
splats, so they run on a SAMPLE of the tiles (every `stride`-th tile index, the same bounded sample bench.py's
cpu_baseline leg uses): image / n_contrib / final_T are compared on those tiles, and the gradients are compared
with the upstream image gradient zeroed outside them (a pixel with zero upstream gradient contributes exactly
nothing to any sum, on both sides).
"""
import importlib

import numpy as np
import pytest
import torch

from _util import n, rel_err

pytestmark = pytest.mark.gpu

synth = importlib.import_module("3dgs_study_b200.synth")

IMG_TOL = 1e-5
GRAD_TOL = 1e-3


def _tile_mask(W, H, stride, offset):
    gx, gy = (W + 15) // 16, (H + 15) // 16
    t = np.arange(gx * gy).reshape(gy, gx)
    sel = (t % stride) == offset
    return np.repeat(np.repeat(sel, 16, axis=0), 16, axis=1)[:H, :W]


def _run(P, W, H, D, stride, offset, antialiasing, want_invdepth, backward, yaw=0.0, exact=True):
    import diff_gaussian_rasterization as dgr
    from oracle import c_oracle as co
    dev = torch.device("cuda:0")
    sc = synth.make_scene(P, D, seed=0)                     # the bench scene (bench.py: seed 0, default s_med)
    cam = synth.make_camera(W, H, yaw=yaw)
    M = sc.shs.shape[1]
    bg = np.array([0.1, 0.2, 0.3], np.float32)
    A = {k: getattr(sc, k).numpy() for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    v, p, c = cam.world_view_transform.numpy(), cam.full_proj_transform.numpy(), cam.camera_center.numpy()

    # ---- oracle: every Gaussian, every pair; pixels of the sampled tiles
    g = co.preprocess(P, D, M, W, H, A["means3D"], A["scales"], 1.0, A["rotations"], A["opacities"], A["shs"], None, None,
                      v, p, c, cam.tanfovx, cam.tanfovy, antialiasing)
    b = co.bin_tiles(P, W, H, g)
    color_o, invd_o, fT_o, nc_o = co.render(W, H, g, b, bg, want_invdepth=want_invdepth, tile_stride=stride, tile_offset=offset)
    mask = _tile_mask(W, H, stride, offset)

    # ---- CUDA path through the C ABI
    d, cd = sc.to(dev), cam.to(dev)
    res = dgr._C.rasterize_gaussians(torch.from_numpy(bg).to(dev), d.means3D, None, d.opacities, d.scales, d.rotations, 1.0, None,
                                     cd.world_view_transform, cd.full_proj_transform, cam.tanfovx, cam.tanfovy, H, W, d.shs, D,
                                     cd.camera_center, False, False, antialiasing=antialiasing, want_invdepth=want_invdepth,
                                     exact=exact)
    R, color, radii, geom, binning, img, invd = res          # R: the pair capacity the backward is handed
    st = dgr._C.export_state(P, W, H, R, radii, geom, binning, img)
    torch.cuda.synchronize()
    assert st["R"] == b.keys.size and st["R"] <= R and (exact is None or st["R"] == R)

    assert np.array_equal(n(radii), g.radii)
    assert np.array_equal(n(st["tiles_touched"]).astype(np.uint32), g.tiles_touched)
    assert np.array_equal(n(st["keys_sorted"]).astype(np.uint64), b.keys)
    assert np.array_equal(n(st["vals_sorted"]).astype(np.uint32), b.vals)
    assert np.array_equal(n(st["ranges"]), b.ranges)
    vis = g.radii > 0
    assert np.array_equal(n(st["means2D"])[vis], g.means2D[vis])
    assert np.array_equal(n(st["conic_opacity"])[vis], g.conic_opacity[vis])
    assert np.array_equal(n(st["rgb"])[vis], g.rgb[vis])
    assert np.array_equal(n(st["n_contrib"]).astype(np.uint32)[mask], nc_o[mask])
    assert float(np.abs(n(st["final_T"]) - fT_o)[mask].max()) <= IMG_TOL
    assert float(np.abs(n(color) - color_o)[:, mask].max()) <= IMG_TOL
    if want_invdepth:
        assert float(np.abs(n(invd) - invd_o)[:, mask].max()) <= IMG_TOL
    # the forward is bit-exact (canonical exp): anything less is unstable at this size
    diff = {"color": int((n(color) != color_o)[:, mask].sum()), "final_T": int((n(st["final_T"]) != fT_o)[mask].sum())}
    if want_invdepth:
        diff["invdepth"] = int((n(invd) != invd_o)[:, mask].sum())
    assert not any(diff.values()), f"forward outputs differ from the oracle in some bit: {diff}"
    out = {"R": st["R"], "visible": int(vis.sum()), "sampled_pixels": int(mask.sum())}
    if not backward:
        return out

    dL = synth.make_grad_image(W, H).numpy() * mask[None].astype(np.float32)
    rg = co.render_backward(P, W, H, g, b, bg, fT_o, nc_o, dL, tile_stride=stride, tile_offset=offset)
    pg = co.preprocess_backward(P, D, M, W, H, A["means3D"], A["scales"], 1.0, A["rotations"], A["opacities"], A["shs"], None,
                                v, p, c, cam.tanfovx, cam.tanfovy, antialiasing, g, rg)
    grads = dgr._C.rasterize_gaussians_backward(
        torch.from_numpy(bg).to(dev), d.means3D, radii, None, d.scales, d.rotations, 1.0, None, cd.world_view_transform,
        cd.full_proj_transform, cam.tanfovx, cam.tanfovy, torch.from_numpy(dL).to(dev), d.shs, D, cd.camera_center, geom, R,
        binning, img, False, opacities=d.opacities, antialiasing=antialiasing, image_height=H, image_width=W)
    torch.cuda.synchronize()
    g2d, _, gop, g3d, _, gsh, gsc, grot = grads
    errs = {"means2D": rel_err(n(g2d)[:, :2], pg.dL_dmean2D), "opacities": rel_err(n(gop).reshape(-1), pg.dL_dopacity),
            "means3D": rel_err(n(g3d), pg.dL_dmeans3D), "shs": rel_err(n(gsh), pg.dL_dsh),
            "scales": rel_err(n(gsc), pg.dL_dscales), "rotations": rel_err(n(grot), pg.dL_drot)}
    bad = {k: e for k, e in errs.items() if not e <= GRAD_TOL}
    where = ""
    if bad:                                                  # name the worst Gaussian of the worst tensor
        pairs = {"means3D": (n(g3d), pg.dL_dmeans3D), "scales": (n(gsc), pg.dL_dscales), "rotations": (n(grot), pg.dL_drot),
                 "shs": (n(gsh).reshape(P, -1), pg.dL_dsh.reshape(P, -1)), "opacities": (n(gop).reshape(P, 1), pg.dL_dopacity.reshape(P, 1)),
                 "means2D": (n(g2d)[:, :2], pg.dL_dmean2D)}
        k = max(bad, key=bad.get)
        a_, b_ = pairs[k]
        i = int(np.abs(a_ - b_).max(axis=1).argmax())
        where = (f"; worst {k}[{i}]: cuda {a_[i]} oracle {b_[i]} scale {A['scales'][i]} radius {g.radii[i]} "
                 f"conic {g.conic_opacity[i]} max|oracle| {np.abs(b_).max():.3e}")
    assert not bad, f"gradient relative errors above {GRAD_TOL}: {bad} (all: {errs}){where}"
    assert float(np.abs(pg.dL_dmeans3D).max()) > 0.0
    out["grad_errs"] = errs
    return out


def test_config2_3M_1080p_full_size():
    """BASELINE configs[1]: 3 M Gaussians SH3, 1920x1080, forward + backward."""
    out = _run(3_000_000, 1920, 1080, 3, stride=16, offset=5, antialiasing=False, want_invdepth=False, backward=True)
    assert out["R"] > 10_000_000 and out["visible"] > 1_500_000


def test_config2_other_view():
    """Same scene from another camera of the bench orbit (the bench renders a different view every step), through
    the no-host-wait path: the binning buffer is sized by the capacity policy, R stays on the device."""
    import diff_gaussian_rasterization as dgr
    dgr._C.reset_capacity_policy()
    _run(3_000_000, 1920, 1080, 3, stride=64, offset=9, antialiasing=False, want_invdepth=False, backward=False, yaw=2.0 * np.pi * 2 / 8,
         exact=None)                                         # first view of the configuration: exact, sets the capacity
    _run(3_000_000, 1920, 1080, 3, stride=32, offset=7, antialiasing=False, want_invdepth=False, backward=True, yaw=2.0 * np.pi * 3 / 8,
         exact=None)
    dgr._C.reset_capacity_policy()


def test_config5_6M_4k_antialiasing_invdepth():
    """BASELINE configs[4]: 6 M Gaussians SH3, 3840x2160, antialiasing + inverse-depth output, forward."""
    out = _run(6_000_000, 3840, 2160, 3, stride=61, offset=11, antialiasing=True, want_invdepth=True, backward=False)
    assert out["R"] > 20_000_000


def test_config3_sweep_point_1600x1060():
    """BASELINE configs[2]: one point of the densification sweep (1 M Gaussians, 1600x1060; ragged tile rows), fwd+bwd."""
    _run(1_000_000, 1600, 1060, 3, stride=16, offset=3, antialiasing=False, want_invdepth=True, backward=True)
