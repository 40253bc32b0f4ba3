"""GPU parity of the steps either side of the rasterizer: fused L1+SSIM loss, one-launch Adam,
3-NN mean squared distance -- CUDA path (through the host mirrors -> ctypes -> C ABI of
include/b200splat_next.h) against oracle/next_oracle.c and the committed reference outputs in
tests/golden/next_steps.npz; at full size through properties and a plain torch fp32 restatement.
Tolerances are written next to each check."""
import ctypes
import json
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import next_oracle as no

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = np.load(os.path.join(ROOT, "tests", "golden", "next_steps.npz"))
LOSS_CASES = ["chw", "tiny", "batch", "win7", "wide"]
ADAM_NAMES = ["xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
DEV = "cuda:0"


def _relmax(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def _note(key, value):
    """Observed error levels, kept next to the run's other outputs (gpurun_out/ when present)."""
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "widen_observed.jsonl"), "a") as f:
            f.write(json.dumps({key: value}) + "\n")


def _torch_ssim(img1, img2, taps):
    """Plain torch fp32 restatement of _ssim (/root/reference/utils/loss_utils.py:76-108) for full-size checks:
    2-D window = outer product of the taps, grouped conv2d with zero padding."""
    C = img1.shape[-3]
    w2 = (taps[:, None] * taps[None, :]).expand(C, 1, 11, 11).contiguous()
    x, y = (img1[None], img2[None]) if img1.dim() == 3 else (img1, img2)
    mu1, mu2 = F.conv2d(x, w2, padding=5, groups=C), F.conv2d(y, w2, padding=5, groups=C)
    s1 = F.conv2d(x * x, w2, padding=5, groups=C) - mu1 * mu1
    s2 = F.conv2d(y * y, w2, padding=5, groups=C) - mu2 * mu2
    s12 = F.conv2d(x * y, w2, padding=5, groups=C) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))).mean()


@pytest.mark.parametrize("name", LOSS_CASES)
def test_loss_matches_reference_and_oracle(name):
    import gaussian_training as gtr
    img_h, gt_h, ws = G[f"loss_{name}_img"], G[f"loss_{name}_gt"], int(G[f"loss_{name}_ws"])
    gt = torch.from_numpy(gt_h).to(DEV)
    # separate calls, like train.py:103-104 makes them
    a = torch.from_numpy(img_h).to(DEV).requires_grad_()
    l1 = gtr.l1_loss(a, gt)
    (g_l1,) = torch.autograd.grad(l1, a)
    a2 = torch.from_numpy(img_h).to(DEV).requires_grad_()
    s = gtr.ssim(a2, gt, window_size=ws)
    (g_s,) = torch.autograd.grad(s, a2)
    # values vs the reference's own output: 1e-6 absolute
    assert abs(l1.item() - float(G[f"loss_{name}_l1"])) <= 1e-6
    assert abs(s.item() - float(G[f"loss_{name}_ssim"])) <= 1e-6
    assert np.array_equal(g_l1.cpu().numpy(), G[f"loss_{name}_g_l1"])          # sign(a-b)/n is exact
    e_ref = _relmax(g_s.cpu().numpy(), G[f"loss_{name}_g_ssim"])
    assert e_ref <= 1e-4                                                       # 1e-4 of the largest gradient entry
    # the fused training loss (one forward + one backward kernel)
    a3 = torch.from_numpy(img_h).to(DEV).requires_grad_()
    loss, Ll1 = gtr.training_loss(a3, gt, 0.2) if ws == 11 else (None, None)
    if loss is not None:
        loss.backward()
        assert abs(loss.item() - float(G[f"loss_{name}_loss"])) <= 1e-6
        assert _relmax(a3.grad.cpu().numpy(), G[f"loss_{name}_g_loss"]) <= 1e-4
    # vs the C oracle (same separable order): 1e-6 absolute on the values, 1e-5 relative on the gradient
    o_loss, o_l1, o_ssim, _ = no.loss_and_grad(img_h, gt_h, 0.2, ws)
    assert abs(l1.item() - o_l1) <= 1e-6 and abs(s.item() - o_ssim) <= 1e-6
    _, _, maps = no.l1_ssim_forward(img_h, gt_h, ws)
    n = img_h.size
    planes = n // (img_h.shape[-1] * img_h.shape[-2])
    og = no.l1_ssim_backward(img_h, gt_h, maps, np.tile(np.array([0.0, 1.0 / n], np.float32), (planes, 1)), ws)
    e_or = _relmax(g_s.cpu().numpy(), og)
    assert e_or <= 1e-5
    _note(f"loss_{name}", {"grad_vs_reference": e_ref, "grad_vs_oracle": e_or})


def test_loss_per_batch_means_and_no_grad_path():
    import gaussian_training as gtr
    img = torch.from_numpy(G["loss_batch_img"]).to(DEV).requires_grad_()
    gt = torch.from_numpy(G["loss_batch_gt"]).to(DEV)
    sb = gtr.ssim(img, gt, size_average=False)
    assert np.abs(sb.detach().cpu().numpy() - G["loss_batch_ssim_per_batch"]).max() <= 1e-6
    w = torch.from_numpy(G["loss_batch_batch_w"]).to(DEV)
    (g,) = torch.autograd.grad((sb * w).sum(), img)
    assert _relmax(g.cpu().numpy(), G["loss_batch_g_ssim_per_batch_w"]) <= 1e-4
    with torch.no_grad():                                                      # metrics.py:72 evaluates ssim without grad
        s = gtr.ssim(img.detach(), gt)
    assert abs(s.item() - float(G["loss_batch_ssim"])) <= 1e-6


def test_loss_full_size_properties_and_torch_reference():
    """1920x1080x3 (BASELINE configs[1] image size): ssim(x,x) = 1, l1(x,x) = 0 with zero gradient; value and
    gradient against the plain torch fp32 restatement on the same device; determinism."""
    import gaussian_training as gtr
    g = torch.Generator().manual_seed(7)
    gt = torch.rand(3, 1080, 1920, generator=g).to(DEV)
    img = (gt + 0.1 * torch.randn(3, 1080, 1920, generator=g).to(DEV)).clamp(0, 1)
    l1, s = gtr.l1_and_ssim(gt, gt)
    assert l1.item() == 0.0 and abs(s.item() - 1.0) <= 1e-6
    a = img.clone().requires_grad_()
    loss, _ = gtr.training_loss(a, gt, 0.2)
    loss.backward()
    b = img.clone().requires_grad_()
    taps = torch.tensor(list(gtr.loss_utils.window_taps(11)), device=DEV)
    ref = 0.8 * (b - gt).abs().mean() + 0.2 * (1.0 - _torch_ssim(b, gt, taps))
    ref.backward()
    assert abs(loss.item() - ref.item()) <= 2e-6                               # fp32 means over 6.2 M values
    e = _relmax(a.grad.cpu().numpy(), b.grad.cpu().numpy())
    assert e <= 1e-4
    a2 = img.clone().requires_grad_()
    loss2, _ = gtr.training_loss(a2, gt, 0.2)
    loss2.backward()
    assert loss2.item() == loss.item() and torch.equal(a2.grad, a.grad)        # fixed-order reductions
    _note("loss_1080p", {"loss_vs_torch": abs(loss.item() - ref.item()), "grad_vs_torch": e})


def test_loss_errors():
    import gaussian_training as gtr
    with pytest.raises(RuntimeError, match="CUDA"):
        gtr.l1_loss(torch.zeros(3, 8, 8), torch.zeros(3, 8, 8))
    with pytest.raises(RuntimeError, match="equal shape"):
        gtr.ssim(torch.zeros(3, 8, 8, device=DEV), torch.zeros(3, 8, 9, device=DEV))
    with pytest.raises(ValueError):
        gtr.ssim(torch.zeros(3, 8, 8, device=DEV), torch.zeros(3, 8, 8, device=DEV), window_size=13)


def _adam_groups(dev):
    params = {n: torch.nn.Parameter(torch.from_numpy(G[f"adam_init_{n}"].copy()).to(dev)) for n in ADAM_NAMES}
    groups = [{"params": [params[n]], "lr": 0.0, "name": n} for n in ADAM_NAMES]   # gaussian_model.py:185-192
    return params, groups


def test_adam_matches_reference_optimizer_and_oracle():
    import gaussian_training as gtr
    params, groups = _adam_groups(DEV)
    opt = gtr.FusedAdam(groups, lr=0.0, eps=1e-15)                                # gaussian_model.py:197
    iters, lrs = G["adam_iters"], G["adam_lrs"]
    sched = gtr.expon_lr(0.00016 * 2.5, 0.0000016 * 2.5, lr_delay_mult=0.01, max_steps=30000)
    orc = {n: (G[f"adam_init_{n}"].copy(), np.zeros_like(G[f"adam_init_{n}"]), np.zeros_like(G[f"adam_init_{n}"])) for n in ADAM_NAMES}
    mism = 0
    for k, it in enumerate(iters):
        for gi, grp in enumerate(opt.param_groups):
            grp["lr"] = sched(int(it)) if grp["name"] == "xyz" else float(lrs[k, gi])   # update_learning_rate, :207-214
        assert abs(opt.param_groups[0]["lr"] - lrs[k, 0]) <= 1e-12 * lrs[k, 0]
        for n in ADAM_NAMES:
            params[n].grad = torch.from_numpy(G[f"adam_grad_{n}_{k}"]).to(DEV)
        opt.step()
        opt.zero_grad(set_to_none=True)
        for gi, n in enumerate(ADAM_NAMES):
            got = params[n].detach().cpu().numpy()
            ref = G[f"adam_param_{n}_{k}"]
            # vs torch.optim.Adam as the reference drives it: 2e-7 absolute + 1e-6 relative (a few ulp)
            assert np.all(np.abs(got - ref) <= 2e-7 + 1e-6 * np.abs(ref)), (n, k, float(np.abs(got - ref).max()))
            p, m, v = orc[n]
            no.adam_step(p, np.ascontiguousarray(G[f"adam_grad_{n}_{k}"]), m, v, float(opt.param_groups[gi]["lr"]), k + 1)
            # vs the C oracle (same rounding points, IEEE division and square root on both sides): bit-exact
            mism += int((got != p).sum())
            assert np.array_equal(got, p), (n, k, int((got != p).sum()), float(np.abs(got - p).max()))
    for n in ADAM_NAMES:
        st = opt.state[params[n]]
        assert int(st["step"].item()) == len(iters)
        assert _relmax(st["exp_avg"].cpu().numpy(), G[f"adam_m_{n}"]) <= 1e-6
        assert _relmax(st["exp_avg_sq"].cpu().numpy(), G[f"adam_v_{n}"]) <= 1e-6
    _note("adam", {"elements_differing_from_oracle": mism})


def test_adam_large_unaligned_and_missing_grad():
    """Tail handling (n not a multiple of 4 / of the block), a misaligned view, a parameter without .grad,
    and the state surgery the reference's densification performs (state entry replaced by a longer one)."""
    import gaussian_training as gtr
    g = torch.Generator().manual_seed(3)
    big = torch.randn(1_000_003, generator=g)
    base = torch.randn(4099, generator=g).to(DEV)
    p_big = torch.nn.Parameter(big.clone().to(DEV))
    p_odd = torch.nn.Parameter(base[1:4098])                  # a view 4 bytes off a 16-byte boundary: scalar path
    p_skip = torch.nn.Parameter(torch.ones(5, device=DEV))
    opt = gtr.FusedAdam([{"params": [p_big], "lr": 1e-2, "name": "a"}, {"params": [p_odd], "lr": 3e-3, "name": "b"},
                         {"params": [p_skip], "lr": 1.0, "name": "c"}], lr=0.0, eps=1e-15)
    ref_p = [big.clone().numpy(), base[1:4098].cpu().numpy().copy()]
    ref_m = [np.zeros_like(x) for x in ref_p]
    ref_v = [np.zeros_like(x) for x in ref_p]
    for step in (1, 2, 3):
        gs = [torch.randn(1_000_003, generator=g) * 1e-3, torch.randn(4097, generator=g)]
        p_big.grad, p_odd.grad = gs[0].to(DEV), gs[1].to(DEV)
        opt.step()
        for i, lr in enumerate((1e-2, 3e-3)):
            no.adam_step(ref_p[i], gs[i].numpy(), ref_m[i], ref_v[i], lr, step)
    assert _relmax(p_big.detach().cpu().numpy(), ref_p[0]) <= 1e-6
    assert _relmax(p_odd.detach().cpu().numpy(), ref_p[1]) <= 1e-6
    assert torch.equal(p_skip.detach(), torch.ones(5, device=DEV)) and len(opt.state[p_skip]) == 0
    # cat_tensors_to_optimizer-style surgery (scene/gaussian_model.py:387-409): longer tensor, state extended by zeros
    st = opt.state.pop(p_odd)
    new_p = torch.nn.Parameter(torch.cat([p_odd.detach(), torch.zeros(10, device=DEV)]))
    st["exp_avg"] = torch.cat([st["exp_avg"], torch.zeros(10, device=DEV)])
    st["exp_avg_sq"] = torch.cat([st["exp_avg_sq"], torch.zeros(10, device=DEV)])
    opt.param_groups[1]["params"][0] = new_p
    opt.state[new_p] = st
    new_p.grad = torch.ones_like(new_p)
    p_big.grad = None
    opt.step()
    assert int(opt.state[new_p]["step"].item()) == 4 and torch.isfinite(new_p).all()
    assert new_p[-1].item() < 0.0                                                  # moved against its gradient


def test_adam_row_split_learning_rates():
    """One [P, M, 3] SH parameter trained with the f_dc / f_rest learning rates (row_width 48, row_split 3) equals two
    separate Adam groups on the column blocks (the reference's layout, gaussian_model.py:187-188)."""
    import gaussian_training as gtr
    g = torch.Generator().manual_seed(21)
    P, M = 1001, 16
    feats = torch.randn(P, M, 3, generator=g)
    p = torch.nn.Parameter(feats.clone().to(DEV))
    opt = gtr.FusedAdam([{"params": [p], "lr": 0.0025, "lr_tail": 0.0025 / 20.0, "row_width": 3 * M, "row_split": 3,
                          "name": "features"}], lr=0.0, eps=1e-15)
    rp, rm, rv = feats.numpy().copy(), np.zeros((P, M, 3), np.float32), np.zeros((P, M, 3), np.float32)
    for step in (1, 2, 3):
        gr = torch.randn(P, M, 3, generator=g) * 1e-2
        p.grad = gr.to(DEV)
        opt.step()
        no.adam_step_rows(rp, gr.numpy(), rm, rv, 0.0025, 0.0025 / 20.0, 3 * M, 3, step)
    got = p.detach().cpu().numpy()
    assert np.array_equal(got, rp), int((got != rp).sum())                      # bit-exact vs the C oracle
    _note("adam_row_split", {"elements_differing_from_oracle": int((got != rp).sum())})
    # the two column blocks really moved at different rates
    d = np.abs(got - feats.numpy())
    assert d[:, 0, :].mean() > 5 * d[:, 1:, :].mean()


def test_activations_match_reference_and_oracle():
    import gaussian_training as gtr
    raw = {k: torch.from_numpy(G[f"act_raw_{k}"]).to(DEV).requires_grad_() for k in ("scaling", "rotation", "opacity")}
    s, r, o = gtr.activate(raw["scaling"], raw["rotation"], raw["opacity"])
    ups = {k: torch.from_numpy(G[f"act_up_{k}"]).to(DEV) for k in raw}
    ((s * ups["scaling"]).sum() + (r * ups["rotation"]).sum() + (o * ups["opacity"]).sum()).backward()
    os_, or_, oo = no.activate_forward(*(G[f"act_raw_{k}"] for k in ("scaling", "rotation", "opacity")))
    ods, odr, odo = no.activate_backward(G["act_raw_rotation"], os_, oo, G["act_up_scaling"], G["act_up_rotation"], G["act_up_opacity"])
    for got, k, orc in ((s, "scaling", os_), (r, "rotation", or_), (o, "opacity", oo)):
        got = got.detach().cpu().numpy()
        ref = G[f"act_out_{k}"]
        assert np.all(np.abs(got - ref) <= 1e-7 + 2e-6 * np.abs(ref)), k              # vs the reference's model: a few ulp
        assert np.all(np.abs(got - orc) <= 1e-7 + 2e-6 * np.abs(orc)), k              # vs the C oracle
    for k, orc in (("scaling", ods), ("rotation", odr), ("opacity", odo)):
        got = raw[k].grad.cpu().numpy()
        ref = G[f"act_grad_{k}"]
        assert np.all(np.abs(got - ref) <= 1e-6 * np.abs(ref).max() + 2e-5 * np.abs(ref)), (k, float(np.abs(got - ref).max()))
        assert np.all(np.abs(got - orc) <= 1e-6 * np.abs(orc).max() + 2e-5 * np.abs(orc)), k
    # only some outputs used: the others get no gradient kernel input (NULL) and come back as zeros from autograd
    raw2 = {k: torch.from_numpy(G[f"act_raw_{k}"]).to(DEV).requires_grad_() for k in raw}
    s2, _, _ = gtr.activate(raw2["scaling"], raw2["rotation"], raw2["opacity"])
    s2.sum().backward()
    assert torch.allclose(raw2["scaling"].grad, s2.detach())
    with pytest.raises(RuntimeError, match="CUDA"):
        gtr.activate(torch.zeros(2, 3), torch.zeros(2, 4), torch.zeros(2, 1))


def test_knn_bit_exact_against_exhaustive_search():
    from simple_knn._C import distCUDA2
    g = torch.Generator().manual_seed(11)
    for P in (1, 3, 4, 5, 63, 64, 65, 1000, 4097, 20011):
        pts = torch.randn(P, 3, generator=g)
        if P > 100:
            pts[7] = pts[8]                                                        # duplicates: distance 0 counts
            pts[P // 2:] *= 40.0                                                   # two density regimes
        got = distCUDA2(pts.to(DEV)).cpu().numpy()
        ref = no.dist2_knn3(pts.numpy())
        same = (got == ref) | (np.isinf(got) & np.isinf(ref))
        assert same.all(), (P, int((~same).sum()), float(np.abs(got - ref)[np.isfinite(ref)].max()))


def test_knn_against_reference_build_when_present():
    """oracle/_ref/libsimple_knn_ref.so = the reference's own simple_knn.cu compiled where it lies
    (oracle/Makefile `ref`).  Same device, same points: results must agree bit for bit."""
    from simple_knn._C import distCUDA2
    path = os.path.join(ROOT, "oracle", "_ref", "libsimple_knn_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built")
    ref_lib = ctypes.CDLL(path)
    ref_lib.ref_dist_cuda2.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    g = torch.Generator().manual_seed(12)
    for P in (5000, 300_000):
        pts = (torch.randn(P, 3, generator=g) * torch.tensor([1.2, 0.45, 1.2])).to(DEV).contiguous()
        ref = torch.zeros(P, device=DEV)
        torch.cuda.synchronize()
        assert ref_lib.ref_dist_cuda2(P, pts.data_ptr(), ref.data_ptr()) == 0
        got = distCUDA2(pts)
        torch.cuda.synchronize()
        nd = int((got != ref).sum().item())
        _note(f"knn_vs_reference_P{P}", {"differing": nd})
        assert nd == 0, (P, nd, float((got - ref).abs().max().item()))


def test_knn_errors():
    from simple_knn._C import distCUDA2
    with pytest.raises(RuntimeError, match="CUDA"):
        distCUDA2(torch.zeros(4, 3))
    with pytest.raises(RuntimeError, match="num_points"):
        distCUDA2(torch.zeros(4, 2, device=DEV))
    assert distCUDA2(torch.zeros(0, 3, device=DEV)).shape == (0,)
