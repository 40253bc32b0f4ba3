"""GPU parity tests proper: the CUDA path (through the drop-in operator -> ctypes -> C ABI) against the
C oracle on the same seeded inputs.

Gates (BASELINE.json north_star): radii / tiles_touched / sorted keys+values / tile ranges bit-exact;
image <= 1e-5 max-abs; gradients <= 1e-3 relative (||a-b||_inf / ||b||_inf per tensor).
"""
import numpy as np
import pytest
import torch

from _util import Case, n, rel_err

pytestmark = pytest.mark.gpu

IMG_TOL = 1e-5     # max-abs, north_star
GRAD_TOL = 1e-3    # relative, north_star (float atomics)

CASES = {
    "sh3_small": dict(P=3000, W=200, H=136, D=3, seed=1, s_med=0.04, bg=(0.1, 0.2, 0.3)),
    "sh0_config1": dict(P=10000, W=256, H=256, D=0, seed=0, bg=(0.0, 0.0, 0.0)),            # BASELINE configs[0]
    "sh3_ragged_edges": dict(P=5000, W=331, H=173, D=3, seed=2, s_med=0.03, bg=(1.0, 1.0, 1.0)),
    "sh1_of_sh3_yaw": dict(P=4000, W=256, H=144, D=1, M=16, seed=3, s_med=0.03, yaw=0.8),
    "dense_opaque": dict(P=6000, W=128, H=96, D=2, seed=4, s_med=0.25, bg=(0.3, 0.3, 0.3)),  # hits the 1e-4 exit / 0.99 clamp
    "precomp": dict(P=3000, W=160, H=120, D=0, seed=5, s_med=0.05, precomp_color=True, precomp_cov=True),
    "scale_mod": dict(P=3000, W=160, H=120, D=3, seed=6, s_med=0.04, mod=1.7),
    # antialiasing (mip filter): opacity scaled by h = sqrt(max(0.000025, det0 / det)), BASELINE configs[4]
    "sh3_aa": dict(P=3000, W=200, H=136, D=3, seed=1, s_med=0.04, bg=(0.1, 0.2, 0.3), antialiasing=True),
    "precomp_aa": dict(P=3000, W=160, H=120, D=0, seed=5, s_med=0.05, precomp_color=True, precomp_cov=True, antialiasing=True),
    "dense_aa": dict(P=6000, W=128, H=96, D=2, seed=4, s_med=0.25, bg=(0.3, 0.3, 0.3), antialiasing=True),
    # sub-pixel splats: det0 / det <= 0.000025 for ~950 of the 2100 visible Gaussians (the clamp branch, whose
    # d(ratio) terms vanish), 69 of which are blended; the rest take the unclamped branch
    "aa_clamp": dict(P=3000, W=200, H=136, D=3, seed=21, s_med=0.001, bg=(0.1, 0.2, 0.3), antialiasing=True),
    # splats with sigma of hundreds of pixels: the conservative {alpha >= 1/255} box of the render kernels must hold
    "huge_splats": dict(P=400, W=640, H=480, D=1, seed=8, s_med=1.2, bg=(0.2, 0.2, 0.2)),
    # 257 x 256 = 65792 tiles > 65536: the 32-bit tile-key instantiations of the binning kernels
    "u32_tile_keys": dict(P=1500, W=4112, H=4096, D=0, seed=12, s_med=0.025),
    # depth keys spanning 28 significant bits (1 .. 1e5 scene units) instead of the 26 of every other case: the depth order
    # sizes its digits from the key range the preprocess kernel saw (kmin / kmax in the control block)
    "deep_range": dict(P=3000, W=200, H=136, D=1, seed=31, s_med=0.04, bg=(0.1, 0.1, 0.1), far=400),
}


def _depth_key_bits(depths, vis):
    k = depths[vis].view(np.uint32)
    return int(k.max() - k.min()).bit_length()


def _integer_gates(case, fr, R, radii, st):
    assert R == fr.binning.keys.size
    assert np.array_equal(n(radii), fr.geom.radii)
    assert np.array_equal(n(st["tiles_touched"]).astype(np.uint32), fr.geom.tiles_touched)
    assert np.array_equal(n(st["keys_sorted"]).astype(np.uint64), fr.binning.keys)
    assert np.array_equal(n(st["vals_sorted"]).astype(np.uint32), fr.binning.vals)
    assert np.array_equal(n(st["ranges"]), fr.binning.ranges)
    assert np.array_equal(n(st["n_contrib"]).astype(np.uint32), fr.n_contrib)
    vis = fr.geom.radii > 0
    assert np.array_equal(n(st["means2D"])[vis], fr.geom.means2D[vis])      # canonical arithmetic: bit-exact
    assert np.array_equal(n(st["depths"])[vis], fr.geom.depths[vis])
    assert np.array_equal(n(st["conic_opacity"])[vis], fr.geom.conic_opacity[vis])
    assert np.array_equal(n(st["rgb"])[vis], fr.geom.rgb[vis])


@pytest.mark.parametrize("name", list(CASES))
def test_forward_state_bit_exact_and_image(name):
    case = Case(**CASES[name])
    fr = case.oracle_forward(want_invdepth=True)
    R, color, radii, invd, st = case.cuda_state(want_invdepth=True)
    _integer_gates(case, fr, R, radii, st)
    if name == "deep_range":
        assert _depth_key_bits(fr.geom.depths, fr.geom.radii > 0) > 27
    assert float(np.abs(n(color) - fr.color).max()) <= IMG_TOL
    assert float(np.abs(n(invd) - fr.invdepth).max()) <= IMG_TOL
    assert float(np.abs(n(st["final_T"]) - fr.final_T).max()) <= IMG_TOL
    # stronger than the 1e-5 gate: with exp part of the canonical arithmetic the whole forward is bit-exact
    # (a flipped alpha >= 1/255 decision would move a pixel by up to 4e-3, so "nearly equal" is not a stable state)
    _bit_exact_forward(color, invd, st, fr)


def _bit_exact_forward(color, invd, st, fr):
    diff = {"color": int((n(color) != fr.color).sum()), "final_T": int((n(st["final_T"]) != fr.final_T).sum())}
    if invd is not None and fr.invdepth is not None:
        diff["invdepth"] = int((n(invd) != fr.invdepth).sum())
    assert not any(diff.values()), f"forward outputs differ from the oracle in some bit: {diff}"


def test_capacity_modes_match_exact_mode_and_report_overflow(monkeypatch):
    """b200splat.h "Pair capacity": a caller-supplied capacity (no host wait) gives the same state as exact mode; a
    capacity that is too small is flagged, writes nothing out of bounds, and the Python policy raises one call late."""
    import diff_gaussian_rasterization as dgr
    C = dgr._C
    case = Case(**CASES["sh3_ragged_edges"])
    fr = case.oracle_forward(want_invdepth=True)
    R0, color0, radii0, invd0, st0 = case.cuda_state(want_invdepth=True)              # exact
    R1, color1, radii1, invd1, st1 = case.cuda_state(want_invdepth=True, exact=False, bin_capacity=R0 + 12345)
    assert R0 == R1 == fr.binning.keys.size
    _integer_gates(case, fr, R1, radii1, st1)
    assert torch.equal(color0, color1) and torch.equal(invd0, invd1)
    dev = torch.device("cuda:0")
    s, c = case.scene.to(dev), case.cam.to(dev)

    def fwd(mod=1.0, **kw):
        return C.rasterize_gaussians(case.bg.to(dev), s.means3D, None, s.opacities, s.scales, s.rotations, mod, None,
                                     c.world_view_transform, c.full_proj_transform, c.tanfovx, c.tanfovy, case.H, case.W, s.shs,
                                     case.D, c.camera_center, False, False, **kw)
    res = fwd(bin_capacity=R0 // 3)                                                  # too small: flagged, in bounds
    torch.cuda.synchronize()
    assert C.forward_status(res[5], case.H, case.W) == (R0, R0 // 3, True)
    assert C.forward_status(fwd(exact=True)[5], case.H, case.W) == (R0, R0, False)
    # the policy: exact first view, then capacities from the largest R seen (head room 1.5x; the fixed slack of the
    # policy, meant for full-size views, is switched off for this small case)
    monkeypatch.setattr(C, "_CAP_SLACK", 0)
    monkeypatch.setattr(C, "_CAP_QUANTUM", 1)
    C.reset_capacity_policy()
    a = fwd()
    assert a[0] == R0                                                                # exact: num_rendered = R
    b = fwd()
    assert b[0] == int(1.5 * R0) and torch.equal(a[1], b[1]) and torch.equal(a[1], color0)
    torch.cuda.synchronize()
    assert C.forward_status(b[5], case.H, case.W)[2] is False
    big = fwd(mod=4.0)                                                               # splats 4x larger: far more pairs
    torch.cuda.synchronize()
    Rbig, capbig, over = C.forward_status(big[5], case.H, case.W)
    assert over and Rbig > capbig
    with pytest.raises(RuntimeError, match="PREVIOUS forward"):
        fwd(mod=4.0)
    ok = fwd(mod=4.0)                                                                # capacity raised: fine now
    torch.cuda.synchronize()
    assert C.forward_status(ok[5], case.H, case.W) == (Rbig, ok[0], False)
    ex = fwd(mod=4.0, exact=True)
    assert torch.equal(ok[1], ex[1])
    C.reset_capacity_policy()
    C.lib().b200splat_shutdown()


@pytest.mark.parametrize("name", list(CASES))
def test_backward_gradients(name):
    case = Case(**CASES[name])
    fr = case.oracle_forward()
    pg = case.oracle_backward(fr)
    color, radii, _, g = case.cuda_run()
    assert float(np.abs(n(color) - fr.color).max()) <= IMG_TOL
    errs = {"means3D": rel_err(n(g["means3D"]), pg.dL_dmeans3D),
            "opacities": rel_err(n(g["opacities"]).reshape(-1), pg.dL_dopacity),
            "means2D": rel_err(n(g["means2D"])[:, :2], pg.dL_dmean2D)}
    assert float(np.abs(n(g["means2D"])[:, 2]).max()) == 0.0
    if "shs" in g:
        errs["shs"] = rel_err(n(g["shs"]), pg.dL_dsh)
    else:
        errs["colors"] = rel_err(n(g["colors_precomp"]), pg.dL_dcolor)
    if "scales" in g:
        errs["scales"] = rel_err(n(g["scales"]), pg.dL_dscales)
        errs["rotations"] = rel_err(n(g["rotations"]), pg.dL_drot)
    else:
        errs["cov3D"] = rel_err(n(g["cov3D_precomp"]), pg.dL_dcov3D)
    bad = {k: v for k, v in errs.items() if not v <= GRAD_TOL}
    assert not bad, f"gradient relative errors above {GRAD_TOL}: {bad} (all: {errs})"
    # culled Gaussians receive exactly zero
    culled = fr.geom.radii == 0
    if culled.any():
        assert float(np.abs(n(g["means3D"])[culled]).max()) == 0.0


@pytest.mark.parametrize("name", ["sh3_small", "dense_opaque", "precomp", "sh3_aa", "aa_clamp"])
def test_backward_with_inverse_depth_gradient(name):
    """The inverse-depth image is differentiable (newer upstream API): a loss on colour AND inverse depth."""
    case = Case(**CASES[name])
    fr = case.oracle_forward(want_invdepth=True)
    pg = case.oracle_backward(fr, invdepth_grad=True)
    pg0 = case.oracle_backward(fr)
    assert rel_err(pg.dL_dmeans3D, pg0.dL_dmeans3D) > 1e-3      # the depth term is not negligible in this case
    color, radii, invd, g = case.cuda_run(invdepth_grad=True)
    errs = {"means3D": rel_err(n(g["means3D"]), pg.dL_dmeans3D),
            "opacities": rel_err(n(g["opacities"]).reshape(-1), pg.dL_dopacity),
            "means2D": rel_err(n(g["means2D"])[:, :2], pg.dL_dmean2D)}
    if "scales" in g:
        errs["scales"] = rel_err(n(g["scales"]), pg.dL_dscales)
        errs["rotations"] = rel_err(n(g["rotations"]), pg.dL_drot)
    else:
        errs["cov3D"] = rel_err(n(g["cov3D_precomp"]), pg.dL_dcov3D)
    bad = {k: v for k, v in errs.items() if not v <= GRAD_TOL}
    assert not bad, f"gradient relative errors above {GRAD_TOL}: {bad} (all: {errs})"


def test_empty_and_all_culled():
    import diff_gaussian_rasterization as dgr
    dev = torch.device("cuda:0")
    case = Case(P=64, W=64, H=48, D=0, seed=9)
    s, c = case.scene.to(dev), case.cam.to(dev)
    behind = s.means3D.clone()
    behind[:, 2] = -50.0                         # everything behind the camera -> all culled
    rs = dgr.GaussianRasterizationSettings(48, 64, c.tanfovx, c.tanfovy, torch.tensor([0.25, 0.5, 0.75], device=dev), 1.0,
                                           c.world_view_transform, c.full_proj_transform, 0, c.camera_center, False, False)
    m = behind.requires_grad_()
    color, radii, _ = dgr.GaussianRasterizer(rs)(means3D=m, means2D=torch.zeros_like(m), opacities=s.opacities, shs=s.shs,
                                                scales=s.scales, rotations=s.rotations)
    assert int(radii.abs().sum()) == 0
    assert torch.equal(color[0], torch.full_like(color[0], 0.25)) and torch.equal(color[2], torch.full_like(color[2], 0.75))
    color.sum().backward()
    assert float(m.grad.abs().max()) == 0.0
    # P = 0
    z = torch.zeros(0, 3, device=dev)
    color0, radii0, _ = dgr.GaussianRasterizer(rs)(means3D=z, means2D=z, opacities=torch.zeros(0, 1, device=dev),
                                                  shs=torch.zeros(0, 1, 3, device=dev), scales=z, rotations=torch.zeros(0, 4, device=dev))
    assert radii0.numel() == 0 and torch.equal(color0[1], torch.full_like(color0[1], 0.5))


def test_mark_visible_and_errors():
    import diff_gaussian_rasterization as dgr
    dev = torch.device("cuda:0")
    case = Case(P=500, W=64, H=48, D=0, seed=11)
    s, c = case.scene.to(dev), case.cam.to(dev)
    rs = dgr.GaussianRasterizationSettings(48, 64, c.tanfovx, c.tanfovy, case.bg.to(dev), 1.0, c.world_view_transform,
                                           c.full_proj_transform, 0, c.camera_center, False, False)
    r = dgr.GaussianRasterizer(rs)
    vis = r.markVisible(s.means3D)
    # the kernel computes view-space z with the canonical fma chain, torch.matmul in its own order: the two may
    # only disagree for points whose z is within rounding of the 0.2 near plane
    p4 = torch.cat([s.means3D, torch.ones(500, 1, device=dev)], 1).double() @ c.world_view_transform.double()
    far_from_plane = (p4[:, 2] - 0.2).abs() > 1e-5
    assert torch.equal(vis[far_from_plane], (p4[:, 2] > 0.2)[far_from_plane])
    assert int(vis.sum()) > 0 and int((~vis).sum()) >= 0
    with pytest.raises(Exception):
        r(means3D=s.means3D, means2D=None, opacities=s.opacities, shs=s.shs, colors_precomp=torch.zeros(500, 3, device=dev),
          scales=s.scales, rotations=s.rotations)
    with pytest.raises(Exception):
        r(means3D=s.means3D, means2D=None, opacities=s.opacities, shs=s.shs)
    with pytest.raises(RuntimeError):   # CPU tensors: no fallback
        r(means3D=s.means3D.cpu(), means2D=None, opacities=s.opacities.cpu(), shs=s.shs.cpu(), scales=s.scales.cpu(),
          rotations=s.rotations.cpu())


def test_factorized_exchange_pieces_match_direct_path():
    """Single GPU: masked colour gradient + b200splat_sh_grad_from_views == the direct dL/dsh, and the factorized
    view-parallel step (2 local views) == the autograd step."""
    import importlib
    vp = importlib.import_module("3dgs_study_b200.view_parallel")
    synth_ = importlib.import_module("3dgs_study_b200.synth")
    dev = torch.device("cuda:0")
    for P, Dg in ((5000, 3), (1234, 1), (777, 0)):
        W, H = 200, 120
        sc = synth_.make_scene(P, Dg, seed=3, s_med=0.04).to(dev)
        cams = [synth_.make_camera(W, H, yaw=0.4 * k).to(dev) for k in range(2)]
        dLs = [synth_.make_grad_image(W, H, seed=40 + k).to(dev) for k in range(2)]
        params = vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)
        bg = torch.tensor([0.2, 0.1, 0.3], device=dev)
        a = vp.ViewParallelStep(params, Dg, bg, exchange="allreduce")
        ca = a.step(cams, dLs)
        for mode in ("factorized", "packed"):      # packed: rows of touched Gaussians + b200splat_accumulate_packed
            b = vp.ViewParallelStep(params, Dg, bg, exchange=mode)
            for _ in range(2):                     # twice: the second step reuses buffers of the first
                b.reset_densification_stats()      # (the statistics accumulate over steps until reset)
                cb = b.step(cams, dLs)
            torch.cuda.synchronize()
            assert torch.equal(ca, cb)
            for name, x, y in zip(vp.GaussianParams._fields, a.grads(), b.grads()):
                err = float((x - y).abs().max()) / max(float(x.abs().max()), 1e-20)
                assert err <= 5e-4, f"{mode} P={P} D={Dg} {name}: {err}"   # float-atomic order differs run to run
            assert torch.allclose(a.xyz_gradient_accum, b.xyz_gradient_accum, rtol=1e-5, atol=1e-12)
            assert torch.equal(a.denom, b.denom) and torch.equal(a.max_radii2D, b.max_radii2D)
            if mode == "packed":
                # b200splat_backward_packed_clean: accumulator rows, touched flags and their count are all zero again
                clean = P * 48 + (P + 255) // 256 * 256 + 256
                assert b._clean_scratch is not None and int(b._clean_scratch[:clean].count_nonzero()) == 0


def test_packed_step_leaves_nothing_of_the_previous_step():
    """The packed step keeps its dense gradients from step to step and clears only the rows the last step wrote
    (b200splat_clear_accumulated): after a step on view A, a step on view B must give what a fresh stepper gives for B --
    also when B touches far fewer Gaussians than A -- and untouched Gaussians must hold exact zeros."""
    import importlib
    vp = importlib.import_module("3dgs_study_b200.view_parallel")
    synth_ = importlib.import_module("3dgs_study_b200.synth")
    dev = torch.device("cuda:0")
    P, Dg, W, H = 8000, 2, 240, 136
    sc = synth_.make_scene(P, Dg, seed=5, s_med=0.04).to(dev)
    cams = [synth_.make_camera(W, H, yaw=y).to(dev) for y in (0.0, 2.5, 1.2)]
    dLs = [synth_.make_grad_image(W, H, seed=60 + k).to(dev) for k in range(3)]
    dLs[1][:, :, W // 4:] = 0.0                      # view B: only a quarter of the image carries gradient
    params = vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)
    bg = torch.tensor([0.2, 0.1, 0.3], device=dev)
    run = vp.ViewParallelStep(params, Dg, bg, exchange="packed")
    for k in (0, 1, 2, 1):
        run.step([cams[k]], [dLs[k]])
        fresh = vp.ViewParallelStep(params, Dg, bg, exchange="allreduce")
        fresh.step([cams[k]], [dLs[k]])
        torch.cuda.synchronize()
        for name, x, y in zip(vp.GaussianParams._fields, run.grads(), fresh.grads()):
            err = float((x - y).abs().max()) / max(float(y.abs().max()), 1e-20)
            assert err <= 5e-4, f"view {k} {name}: {err}"
            untouched = (y.reshape(P, -1) == 0).all(dim=1)
            assert int(untouched.sum()) > P // 4 and bool((x.reshape(P, -1)[untouched] == 0).all()), f"view {k} {name}: stale rows"
    # two views per step: the second one's rows are added onto the first one's
    run.step([cams[0], cams[2]], [dLs[0], dLs[2]])
    fresh = vp.ViewParallelStep(params, Dg, bg, exchange="allreduce")
    fresh.step([cams[0], cams[2]], [dLs[0], dLs[2]])
    torch.cuda.synchronize()
    for name, x, y in zip(vp.GaussianParams._fields, run.grads(), fresh.grads()):
        assert float((x - y).abs().max()) / max(float(y.abs().max()), 1e-20) <= 5e-4, name


def test_graphed_step_matches_eager_step():
    """view_parallel.GraphedStep: the whole packed-row step captured in a CUDA graph and replayed gives what the eagerly
    queued step gives, view after view, and the densification statistics keep accumulating across replays."""
    import importlib
    vp = importlib.import_module("3dgs_study_b200.view_parallel")
    synth_ = importlib.import_module("3dgs_study_b200.synth")
    import diff_gaussian_rasterization as dgr
    dgr._C.reset_capacity_policy()
    dev = torch.device("cuda:0")
    P, Dg, W, H = 6000, 3, 240, 136
    sc = synth_.make_scene(P, Dg, seed=3, s_med=0.04).to(dev)
    cams = [synth_.make_camera(W, H, yaw=0.3 * k).to(dev) for k in range(4)]
    dLs = [synth_.make_grad_image(W, H, seed=50 + k).to(dev) for k in range(4)]
    params = vp.GaussianParams(sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)
    bg = torch.tensor([0.2, 0.1, 0.3], device=dev)
    eager = vp.ViewParallelStep(params, Dg, bg, exchange="packed")
    graphed = vp.ViewParallelStep(params, Dg, bg, exchange="packed")
    g = vp.GraphedStep(graphed, [cams[0]], [dLs[0]])
    graphed.reset_densification_stats()                    # the capture's warm-up steps counted too
    for k in (1, 2, 3, 0, 2):
        ce = eager.step([cams[k]], [dLs[k]])
        cg = g.step([cams[k]], [dLs[k]])
        torch.cuda.synchronize()
        assert torch.equal(ce, cg)
        for name, x, y in zip(vp.GaussianParams._fields, eager.grads(), graphed.grads()):
            err = float((x - y).abs().max()) / max(float(x.abs().max()), 1e-20)
            assert err <= 5e-4, f"view {k} {name}: {err}"
    assert torch.allclose(eager.xyz_gradient_accum, graphed.xyz_gradient_accum, rtol=1e-5, atol=1e-12)
    assert torch.equal(eager.denom, graphed.denom) and float(eager.denom.max()) >= 4
    dgr._C.reset_capacity_policy()
