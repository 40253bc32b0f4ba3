"""CPU: the reference's OWN optimizer handling, unmodified, on top of FusedAdam.

GaussianModel.training_setup builds its optimizer with `torch.optim.Adam(l, lr=0.0, eps=1e-15)`
(/root/reference/scene/gaussian_model.py:197); the one-line change a maintainer makes is to construct
gaussian_training.FusedAdam there instead.  Everything else the reference does to that
optimizer -- update_learning_rate (:207-214), replace_tensor_to_optimizer (:322-335, used by reset_opacity),
_prune_optimizer / prune_points (:337-375), cat_tensors_to_optimizer / densification_postfix (:377-444),
capture / state_dict (:63-77) -- must run unchanged, because FusedAdam keeps torch's state layout.  No kernel is
launched here (no GPU): the optimizer state is filled in by hand as if three steps had been taken."""
import os
import sys
import types

import pytest
import torch

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "scene")), reason="reference not present")

NAMES = ["xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"]


@pytest.fixture()
def model(monkeypatch):
    import gaussian_training as gtr
    monkeypatch.syspath_prepend(REF)
    for name in ("plyfile", "simple_knn", "simple_knn._C"):          # absent third-party modules: import stubs only
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.PlyData = m.PlyElement = m.distCUDA2 = None
            monkeypatch.setitem(sys.modules, name, m)
    zeros = torch.zeros

    def zeros_cpu(*a, **k):                                           # the reference hard-codes device="cuda"
        if str(k.get("device", "")).startswith("cuda"):
            k["device"] = "cpu"
        return zeros(*a, **k)

    monkeypatch.setattr(torch, "zeros", zeros_cpu)
    for mod in [m for m in sys.modules if m == "utils" or m.startswith("utils.") or m == "scene" or m.startswith("scene.")
                or m == "arguments"]:
        monkeypatch.delitem(sys.modules, mod, raising=False)
    from argparse import ArgumentParser
    from arguments import OptimizationParams
    from scene.gaussian_model import GaussianModel
    g = torch.Generator().manual_seed(0)
    P, M = 30, 16
    gm = GaussianModel(3)
    gm.spatial_lr_scale = 1.5
    gm._xyz = torch.nn.Parameter(torch.randn(P, 3, generator=g))
    gm._features_dc = torch.nn.Parameter(torch.randn(P, 1, 3, generator=g))
    gm._features_rest = torch.nn.Parameter(torch.randn(P, M - 1, 3, generator=g))
    gm._opacity = torch.nn.Parameter(torch.randn(P, 1, generator=g))
    gm._scaling = torch.nn.Parameter(torch.randn(P, 3, generator=g))
    gm._rotation = torch.nn.Parameter(torch.randn(P, 4, generator=g))
    gm.max_radii2D = torch.zeros(P)
    monkeypatch.setattr(torch.optim, "Adam", gtr.FusedAdam)          # the maintainer's one-line change, :197
    gm.training_setup(OptimizationParams(ArgumentParser()))
    assert type(gm.optimizer) is gtr.FusedAdam
    for grp in gm.optimizer.param_groups:                             # as if three steps had been taken
        p = grp["params"][0]
        gm.optimizer.state[p] = {"step": torch.tensor(3.0), "exp_avg": torch.randn(p.shape, generator=g),
                                 "exp_avg_sq": torch.rand(p.shape, generator=g)}
    return gm


def _check(gm, P):
    assert [g["name"] for g in gm.optimizer.param_groups] == NAMES
    shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}
    for grp in gm.optimizer.param_groups:
        p = grp["params"][0]
        st = gm.optimizer.state[p]
        assert tuple(p.shape) == shapes[grp["name"]] and p.requires_grad
        assert st["exp_avg"].shape == p.shape and st["exp_avg_sq"].shape == p.shape and float(st["step"]) == 3.0
    assert len(gm.optimizer.state) == 6                               # no stale entries left behind


def test_learning_rate_schedule_and_group_layout(model):
    gm = model
    lr = gm.update_learning_rate(100)
    assert gm.optimizer.param_groups[0]["lr"] == lr and 0 < lr < 0.00016 * 1.5
    assert gm.optimizer.defaults["eps"] == 1e-15
    _check(gm, 30)


def test_prune_densify_and_reset_run_unchanged(model):
    gm = model
    keep_m = gm.optimizer.state[gm._xyz]["exp_avg"].clone()
    mask = torch.zeros(30, dtype=torch.bool)
    mask[::3] = True                                                  # prune every third Gaussian
    gm.prune_points(mask)
    _check(gm, 20)
    assert torch.equal(gm.optimizer.state[gm._xyz]["exp_avg"], keep_m[~mask])
    g = torch.Generator().manual_seed(1)
    new = [torch.randn(7, 3, generator=g), torch.randn(7, 1, 3, generator=g), torch.randn(7, 15, 3, generator=g),
           torch.randn(7, 1, generator=g), torch.randn(7, 3, generator=g), torch.randn(7, 4, generator=g)]
    gm.densification_postfix(*new)
    _check(gm, 27)
    st = gm.optimizer.state[gm._features_rest]
    assert torch.all(st["exp_avg"][20:] == 0) and torch.all(st["exp_avg_sq"][20:] == 0)
    assert torch.equal(gm._features_rest[20:], new[2]) and gm.xyz_gradient_accum.shape == (27, 1)
    gm.reset_opacity()                                                # replace_tensor_to_optimizer
    _check(gm, 27)
    assert torch.all(gm.optimizer.state[gm._opacity]["exp_avg"] == 0)
    assert float(torch.sigmoid(gm._opacity.detach()).max()) <= 0.01 + 1e-6


def test_capture_state_dict_round_trip_and_loud_cpu_failure(model):
    import gaussian_training as gtr
    gm = model
    sd = gm.capture()[-2]                                             # optimizer.state_dict(), :63-77
    other = gtr.FusedAdam([{"params": [g["params"][0]], "lr": 0.0, "name": g["name"]} for g in gm.optimizer.param_groups],
                          lr=0.0, eps=1e-15)
    other.load_state_dict(sd)                                         # restore path, :79-103
    for grp in other.param_groups:
        st = other.state[grp["params"][0]]
        assert float(st["step"]) == 3.0 and st["exp_avg"].shape == grp["params"][0].shape
    gm._xyz.grad = torch.zeros_like(gm._xyz)
    with pytest.raises(RuntimeError, match="CUDA"):                   # the step itself needs the GPU: no CPU fallback
        gm.optimizer.step()
