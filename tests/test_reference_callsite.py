"""CPU, only where /root/reference exists (the build container): drive the reference's UNMODIFIED
gaussian_renderer.render() up to the operator call and check the call contract our package must satisfy
(12 settings kwargs, the 8 keyword tensors, a 2-tuple result in legacy mode).  The rasterizer call itself is
replaced by a recorder -- render() hard-codes device="cuda", so the call cannot execute here, and
/root/reference does not exist on the GPU box."""
import os
import sys
import types

import pytest
import torch

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "gaussian_renderer")), reason="reference not present")


def test_unmodified_render_calls_our_operator_surface(monkeypatch):
    import diff_gaussian_rasterization as dgr
    real_cls = dgr.GaussianRasterizer
    monkeypatch.syspath_prepend(REF)
    # stubs for modules the reference imports but this image lacks (plyfile, simple_knn)
    ply = types.ModuleType("plyfile"); ply.PlyData = object; ply.PlyElement = object
    knn = types.ModuleType("simple_knn"); knn_c = types.ModuleType("simple_knn._C"); knn_c.distCUDA2 = lambda x: x
    monkeypatch.setitem(sys.modules, "plyfile", ply)
    monkeypatch.setitem(sys.modules, "simple_knn", knn)
    monkeypatch.setitem(sys.modules, "simple_knn._C", knn_c)
    for m in [k for k in sys.modules if k.split(".")[0] in ("gaussian_renderer", "scene", "utils")]:
        monkeypatch.delitem(sys.modules, m, raising=False)
    seen = {}

    class Recorder(torch.nn.Module):
        def __init__(self, raster_settings):
            super().__init__()
            seen["settings"] = raster_settings

        def forward(self, **kw):
            seen["kwargs"] = kw
            P = kw["means3D"].shape[0]
            return torch.zeros(3, 4, 4), torch.ones(P, dtype=torch.int32)     # the 2-tuple the call site unpacks

    monkeypatch.setattr(dgr, "GaussianRasterizer", Recorder)
    zl = torch.zeros_like
    monkeypatch.setattr(torch, "zeros_like", lambda t, **k: zl(t, **{a: b for a, b in k.items() if a != "device"}))
    import gaussian_renderer as gr           # the reference's module, importing OUR package by name
    assert gr.GaussianRasterizationSettings is dgr.GaussianRasterizationSettings

    P = 5
    pc = types.SimpleNamespace(get_xyz=torch.zeros(P, 3), get_opacity=torch.ones(P, 1), get_scaling=torch.ones(P, 3),
                               get_rotation=torch.ones(P, 4), get_features=torch.zeros(P, 16, 3), active_sh_degree=2,
                               max_sh_degree=3)
    cam = types.SimpleNamespace(FoVx=1.0, FoVy=0.8, image_height=4, image_width=4, world_view_transform=torch.eye(4),
                                full_proj_transform=torch.eye(4), camera_center=torch.zeros(3))
    pipe = types.SimpleNamespace(debug=False, compute_cov3D_python=False, convert_SHs_python=False)
    out = gr.render(cam, pc, pipe, torch.zeros(3))
    assert set(out) == {"render", "viewspace_points", "visibility_filter", "radii"}
    rs = seen["settings"]
    assert isinstance(rs, dgr.GaussianRasterizationSettings) and rs.sh_degree == 2 and rs.antialiasing is False
    assert set(seen["kwargs"]) == {"means3D", "means2D", "shs", "colors_precomp", "opacities", "scales", "rotations",
                                   "cov3D_precomp"}
    assert seen["kwargs"]["colors_precomp"] is None and seen["kwargs"]["cov3D_precomp"] is None
    # our real module accepts exactly that keyword set
    import inspect
    params = set(inspect.signature(real_cls.forward).parameters) - {"self"}
    assert params == set(seen["kwargs"])
