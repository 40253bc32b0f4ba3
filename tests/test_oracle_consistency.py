"""CPU: the two oracles check each other.  The C oracle (hand-derived backward, canonical fp32 order) and the
PyTorch oracle (autograd backward) must agree: integers bit-exact, image 1e-5, gradients 1e-3 relative
(measured ~1e-6)."""
import numpy as np
import pytest
import torch

from _util import Case, n, rel_err
from oracle import torch_ref as tr

CASES = {
    "sh3": dict(P=1500, W=160, H=112, D=3, seed=3, s_med=0.05, bg=(0.2, 0.5, 0.9)),
    "config1_sh0": dict(P=10000, W=256, H=256, D=0, seed=0),
    "precomp_aa": dict(P=1200, W=96, H=80, D=0, seed=5, s_med=0.06, precomp_color=True, precomp_cov=True, antialiasing=True),
    "dense": dict(P=2000, W=64, H=48, D=1, seed=4, s_med=0.3, bg=(1, 1, 1), mod=1.3),
}


def _torch_run(case, invdepth_grad=False):
    s, c = case.scene, case.cam
    lv = {"means3D": s.means3D.clone().requires_grad_(), "opacities": s.opacities.clone().requires_grad_()}
    kw = dict(shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None)
    if case.colors is not None:
        lv["colors_precomp"] = case.colors.clone().requires_grad_()
    else:
        lv["shs"] = s.shs.clone().requires_grad_()
    if case.cov3D is not None:
        lv["cov3D_precomp"] = case.cov3D.clone().requires_grad_()
    else:
        lv["scales"] = s.scales.clone().requires_grad_(); lv["rotations"] = s.rotations.clone().requires_grad_()
    kw.update({k: v for k, v in lv.items() if k in kw})
    m2d = torch.zeros(case.P, 3, requires_grad=True)
    color, radii, invd, aux = tr.rasterize_gaussians(
        image_height=case.H, image_width=case.W, tanfovx=c.tanfovx, tanfovy=c.tanfovy, bg=case.bg, scale_modifier=case.mod,
        viewmatrix=c.world_view_transform, projmatrix=c.full_proj_transform, sh_degree=case.D, campos=c.camera_center,
        means3D=lv["means3D"], means2D=m2d, opacities=lv["opacities"], antialiasing=case.aa, **kw)
    loss = (color * case.dL).sum()
    if invdepth_grad:
        loss = loss + (invd * case.depth_grad_image()).sum()
    loss.backward()
    return color, radii, invd, aux, lv, m2d


@pytest.mark.parametrize("name", list(CASES))
def test_c_oracle_vs_torch_oracle(name):
    case = Case(**CASES[name])
    fr = case.oracle_forward(want_invdepth=True)
    pg = case.oracle_backward(fr)
    color, radii, invd, aux, lv, m2d = _torch_run(case)
    # integers: counted, expected zero mismatches on these seeds
    assert int((fr.geom.radii != n(radii)).sum()) == 0
    assert int((fr.geom.tiles_touched.astype(np.int64) != n(aux.tiles_touched)).sum()) == 0
    assert np.array_equal(fr.binning.keys.astype(np.int64), n(aux.keys))
    assert np.array_equal(fr.binning.vals.astype(np.int64), n(aux.vals))
    assert np.array_equal(fr.binning.ranges.astype(np.int64), n(aux.ranges))
    assert np.array_equal(fr.n_contrib.astype(np.int64), n(aux.n_contrib))
    # The PyTorch oracle cannot follow the canonical fused-multiply-add order, so its `power` differs from the
    # C oracle's by ~1 ulp; a pair whose alpha lands within 1 ulp of 1/255 then flips the cutoff and moves that
    # pixel by up to T*alpha*c <= 1/255.  Count such pixels instead of hiding them: <= 0.01 % of
    # values, each below 1/255.  (The GPU-vs-C-oracle gate in test_gpu_parity.py has no such allowance.)
    for a, b in ((fr.color, n(color)), (fr.invdepth, n(invd))):
        d = np.abs(a - b)
        assert int((d > 1e-5).sum()) <= max(3, int(1e-4 * d.size)), int((d > 1e-5).sum())
        assert d.max() <= 1.0 / 255.0
    tol = 1e-3
    assert rel_err(pg.dL_dmeans3D, n(lv["means3D"].grad)) <= tol
    assert rel_err(pg.dL_dopacity, n(lv["opacities"].grad).reshape(-1)) <= tol
    assert rel_err(pg.dL_dmean2D, n(m2d.grad)[:, :2]) <= tol
    if "shs" in lv:
        assert rel_err(pg.dL_dsh, n(lv["shs"].grad)) <= tol
    else:
        assert rel_err(pg.dL_dcolor, n(lv["colors_precomp"].grad)) <= tol
    if "scales" in lv:
        assert rel_err(pg.dL_dscales, n(lv["scales"].grad)) <= tol
        assert rel_err(pg.dL_drot, n(lv["rotations"].grad)) <= tol
    else:
        assert rel_err(pg.dL_dcov3D, n(lv["cov3D_precomp"].grad)) <= tol


@pytest.mark.parametrize("name", ["sh3", "dense"])
def test_c_oracle_inverse_depth_gradient_vs_autograd(name):
    """Loss on colour AND the inverse-depth image: the C oracle's hand-derived depth terms against autograd."""
    case = Case(**CASES[name])
    fr = case.oracle_forward(want_invdepth=True)
    pg = case.oracle_backward(fr, invdepth_grad=True)
    pg0 = case.oracle_backward(fr)
    assert rel_err(pg.dL_dmeans3D, pg0.dL_dmeans3D) > 1e-2          # the depth terms matter in these cases
    color, radii, invd, aux, lv, m2d = _torch_run(case, invdepth_grad=True)
    tol = 1e-3
    assert rel_err(pg.dL_dmeans3D, n(lv["means3D"].grad)) <= tol
    assert rel_err(pg.dL_dopacity, n(lv["opacities"].grad).reshape(-1)) <= tol
    assert rel_err(pg.dL_dmean2D, n(m2d.grad)[:, :2]) <= tol
    assert rel_err(pg.dL_dscales, n(lv["scales"].grad)) <= tol
    assert rel_err(pg.dL_drot, n(lv["rotations"].grad)) <= tol


def test_sorted_order_and_ranges_properties():
    """Size-independent properties of the binning stage: keys non-decreasing, ties in Gaussian-index order,
    ranges partition the list, R == sum(tiles_touched)."""
    case = Case(P=20000, W=320, H=200, D=0, seed=8)
    fr = case.oracle_forward()
    k, v, r = fr.binning.keys, fr.binning.vals.astype(np.int64), fr.binning.ranges
    assert k.size == int(fr.geom.tiles_touched.sum())
    assert (np.diff(k.astype(np.int64)) >= 0).all()
    same = np.diff(k.astype(np.int64)) == 0
    assert (np.diff(v)[same] > 0).all()
    nz = r[:, 1] > r[:, 0]
    assert int((r[nz, 1] - r[nz, 0]).sum()) == k.size
    tiles = (k >> np.uint64(32)).astype(np.int64)
    for t in np.nonzero(nz)[0][:50]:
        assert (tiles[r[t, 0]:r[t, 1]] == t).all()
