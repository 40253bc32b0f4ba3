"""CPU tests: pin the oracle (oracle/torch_ref.py restatements and oracle/splat_oracle.c) against
(a) golden vectors produced by the reference's own Python math (tests/golden/make_golden.py) and
(b) closed-form known-answer cases of the published blending rule."""
import math
import os

import numpy as np
import torch

from oracle import c_oracle as co
from oracle import torch_ref as tr

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_math.npz"))


def test_eval_sh_matches_reference():
    sh, dirs = torch.tensor(G["sh_coeffs"]), torch.tensor(G["sh_dirs"])
    for d in range(4):
        assert np.array_equal(tr.eval_sh(d, sh, dirs).numpy(), G[f"sh_eval_deg{d}"])


def test_rotation_covariance_match_reference():
    q, s, mod = torch.tensor(G["quat"]), torch.tensor(G["scale"]), float(G["scale_mod"])
    assert np.array_equal(tr.build_rotation(q).numpy(), G["rotation"])
    assert np.allclose(tr.build_scaling_rotation(mod * s, q).numpy(), G["L"], rtol=0, atol=1e-7)
    assert np.allclose(tr.build_covariance(s, mod, q).numpy(), G["cov6"], rtol=1e-6, atol=1e-7)


def test_camera_matrices_match_reference():
    wvt, full, center = tr.camera_matrices(G["cam_R"], G["cam_T"], float(G["cam_fov"][0]), float(G["cam_fov"][1]))
    assert np.array_equal(wvt.numpy(), G["world_view_transform"])
    assert np.array_equal(full.numpy(), G["full_proj_transform"])
    assert np.array_equal(center.numpy(), G["camera_center"])
    # the product-side harness builds cameras with the same convention
    import importlib
    synth = importlib.import_module("3dgs_study_b200.synth")
    cam = synth.make_camera(640, 360, yaw=0.3)
    Ry = np.array([[math.cos(0.3), 0, math.sin(0.3)], [0, 1, 0], [-math.sin(0.3), 0, math.cos(0.3)]])
    C = Ry @ np.array([0.0, -0.6, -4.0])
    w2, f2, c2 = tr.camera_matrices(Ry, -Ry.T @ C, cam.FoVx, cam.FoVy)
    assert np.allclose(cam.world_view_transform.numpy(), w2.numpy(), atol=1e-6)
    assert np.allclose(cam.full_proj_transform.numpy(), f2.numpy(), atol=1e-6)
    assert np.allclose(cam.camera_center.numpy(), C, atol=1e-5)


def _identity_cam(W, H, tanfov):
    view = np.eye(4, dtype=np.float32)
    proj = tr.getProjectionMatrix(0.01, 100.0, 2 * math.atan(tanfov), 2 * math.atan(tanfov * H / W)).transpose(0, 1).numpy()
    return view, (view @ proj).astype(np.float32), np.zeros(3, np.float32), tanfov, tanfov * H / W


def test_c_oracle_sh_colour_and_cov3d_vs_reference():
    """orc_preprocess's SH colour / cov3D against the reference's eval_sh (+0.5, clamp) and covariance."""
    sh = np.ascontiguousarray(np.transpose(G["sh_coeffs"], (0, 2, 1)))       # [N,16,3]
    dirs = G["sh_dirs"]
    N = dirs.shape[0]
    means = (dirs * 5.0).astype(np.float32)
    qn = G["quat"] / np.linalg.norm(G["quat"], axis=1, keepdims=True)
    W = H = 512
    view, proj, campos, tfx, tfy = _identity_cam(W, H, 8.0)
    for d in range(4):
        g = co.preprocess(N, d, 16, W, H, means, G["scale"] * 0.05, float(G["scale_mod"]), qn.astype(np.float32),
                          np.full(N, 0.5, np.float32), sh, None, None, view, proj, campos, tfx, tfy)
        vis = g.radii > 0
        assert vis.sum() >= 10
        assert np.allclose(g.rgb[vis], G[f"sh_color_deg{d}"][vis], rtol=0, atol=2e-6)
        ref_cov = G["cov6"] * 0.05 ** 2
        assert np.allclose(g.cov3D[vis], ref_cov[vis], rtol=2e-5, atol=1e-9)
        assert (g.depths[vis] > 0.2).all() and not vis[means[:, 2] <= 0.2].any()   # near cull


def _render_single(o, sigma_world=0.02, z=4.0, W=64, H=64, bg=(0.1, 0.2, 0.3), n=1, colors=None):
    view, proj, campos, tfx, tfy = _identity_cam(W, H, 0.5)
    # pixel centre (31.5, 31.5) is the principal point; shift the mean so it lands exactly on pixel (32, 32)
    fx = W / (2 * tfx)
    means = np.tile(np.array([[0.5 * z / fx, 0.5 * z / fx, z]], np.float32), (n, 1))
    means[:, 2] += np.arange(n) * 1e-3
    scales = np.full((n, 3), sigma_world, np.float32)
    rot = np.tile(np.array([[1, 0, 0, 0]], np.float32), (n, 1))
    cols = np.tile(np.array([[0.9, 0.5, 0.2]], np.float32), (n, 1)) if colors is None else colors
    fr = co.forward(n, 0, 0, W, H, np.array(bg, np.float32), means, None, cols, np.full(n, o, np.float32), scales, 1.0, rot,
                    None, view, proj, campos, tfx, tfy)
    return fr, fx, cols


def test_known_answer_single_isotropic_gaussian():
    o, sw, z = 0.8, 0.02, 4.0
    fr, fx, cols = _render_single(o, sw, z)
    assert fr.geom.radii[0] > 0 and abs(fr.geom.means2D[0, 0] - 32.0) < 1e-3
    var = (sw * fx / z) ** 2 + 0.3                      # EWA variance + 0.3 px^2 low-pass
    ys, xs = np.mgrid[0:64, 0:64]
    r2 = (xs - fr.geom.means2D[0, 0]) ** 2 + (ys - fr.geom.means2D[0, 1]) ** 2
    alpha = np.minimum(0.99, o * np.exp(-0.5 * r2 / var))
    alpha[alpha < 1 / 255] = 0.0
    bg = np.array([0.1, 0.2, 0.3])
    want = alpha[None] * cols[0][:, None, None] + (1 - alpha)[None] * bg[:, None, None]
    assert np.abs(fr.color - want).max() < 2e-5
    assert fr.n_contrib[32, 32] == 1 and fr.n_contrib[0, 0] == 0
    # isotropic: mid^2 - det = 0 is floored at 0.1, so lambda_max = var + sqrt(0.1)
    assert fr.geom.radii[0] == math.ceil(3 * math.sqrt(var + math.sqrt(0.1)))


def test_known_answer_alpha_clamp_and_transmittance_exit():
    fr, _, _ = _render_single(1.0)                      # alpha clamps to 0.99 at the centre
    assert abs(fr.final_T[32, 32] - 0.01) < 1e-6
    n = 6
    fr, _, cols = _render_single(0.95, n=n)             # T: .05, .0025, 1.25e-4, then 6.25e-6 < 1e-4 -> stop
    assert fr.n_contrib[32, 32] == 3
    assert abs(fr.final_T[32, 32] - 1.25e-4) < 1e-8
    # ordering: nearer Gaussian first, two colours
    c2 = np.array([[1, 0, 0], [0, 1, 0]], np.float32)
    fr, _, _ = _render_single(0.5, n=2, colors=c2, bg=(0, 0, 1))
    px = fr.color[:, 32, 32]
    assert np.allclose(px, [0.5, 0.25, 0.25], atol=1e-5)
