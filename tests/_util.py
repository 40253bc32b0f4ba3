"""Shared helpers for the parity tests: run the same seeded case through the C oracle (CPU) and the
CUDA path (through the drop-in operator, i.e. through the C ABI)."""
from __future__ import annotations

import importlib
import math

import numpy as np
import torch

synth = importlib.import_module("3dgs_study_b200.synth")


def n(t):
    return t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else t


def rel_err(a, b):
    """||a-b||_inf / max(||b||_inf, eps)  (SURVEY.md 8d gradient gate)."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-20)) if a.size else 0.0


class Case:
    def __init__(self, P, W, H, D=3, seed=0, s_med=None, bg=(0.0, 0.0, 0.0), yaw=0.0, mod=1.0,
                 precomp_color=False, precomp_cov=False, antialiasing=False, M=None, far=0):
        self.P, self.W, self.H, self.D, self.mod, self.aa = P, W, H, D, mod, antialiasing
        maxdeg = D if M is None else int(round(M ** 0.5)) - 1
        self.scene = synth.make_scene(P, maxdeg, seed=seed, s_med=s_med)
        self.cam = synth.make_camera(W, H, yaw=yaw)
        if far:
            # `far` Gaussians moved out along the viewing axis, log-uniform in depth up to ~1e5 scene units: the depth keys
            # then span 28 significant bits instead of the usual 26 (the depth order sizes its digits from the key range)
            g = torch.Generator().manual_seed(seed + 99)
            c = self.cam.camera_center
            axis = -c / c.norm()
            d = torch.exp(torch.empty(far).uniform_(math.log(30.0), math.log(1.0e5), generator=g))
            lateral = (torch.rand(far, 3, generator=g) - 0.5) * 0.3 * d[:, None]
            lateral = lateral - (lateral @ axis)[:, None] * axis[None, :]
            self.scene.means3D[:far] = c[None, :] + axis[None, :] * d[:, None] + lateral
            self.scene.scales[:far] = (0.006 * d)[:, None] * (0.5 + torch.rand(far, 3, generator=g))
            self.scene.opacities[:far] = 0.2 + 0.6 * torch.rand(far, 1, generator=g)
        self.bg = torch.tensor(bg, dtype=torch.float32)
        self.dL = synth.make_grad_image(W, H, seed=seed + 1)
        self.M = self.scene.shs.shape[1]
        self.colors = None
        self.cov3D = None
        if precomp_color:
            g = torch.Generator().manual_seed(seed + 7)
            self.colors = torch.rand(P, 3, generator=g)
        if precomp_cov:
            from oracle import torch_ref as tr
            self.cov3D = tr.build_covariance(self.scene.scales, mod, self.scene.rotations).contiguous()

    # ---- oracle -------------------------------------------------------------------------------
    def oracle_forward(self, want_invdepth=False):
        from oracle import c_oracle as co
        s, c = self.scene, self.cam
        return co.forward(self.P, self.D, self.M, self.W, self.H, n(self.bg), n(s.means3D),
                          None if self.colors is not None else n(s.shs), n(self.colors) if self.colors is not None else None,
                          n(s.opacities), None if self.cov3D is not None else n(s.scales), self.mod,
                          None if self.cov3D is not None else n(s.rotations), n(self.cov3D) if self.cov3D is not None else None,
                          n(c.world_view_transform), n(c.full_proj_transform), n(c.camera_center), c.tanfovx, c.tanfovy,
                          antialiasing=self.aa, want_invdepth=want_invdepth)

    def depth_grad_image(self):
        """Upstream gradient of the inverse-depth image (seeded, same on both sides)."""
        g = torch.Generator().manual_seed(1234 + self.W)
        return (torch.rand(1, self.H, self.W, generator=g) - 0.3).contiguous()

    def oracle_backward(self, fr, invdepth_grad=False):
        from oracle import c_oracle as co
        s, c = self.scene, self.cam
        return co.backward(fr, self.P, self.D, self.M, self.W, self.H, n(self.bg), n(s.means3D),
                           None if self.colors is not None else n(s.shs), n(s.opacities),
                           None if self.cov3D is not None else n(s.scales), self.mod,
                           None if self.cov3D is not None else n(s.rotations), n(self.cov3D) if self.cov3D is not None else None,
                           n(c.world_view_transform), n(c.full_proj_transform), n(c.camera_center), c.tanfovx, c.tanfovy,
                           n(self.dL), dL_dinvdepth=n(self.depth_grad_image()) if invdepth_grad else None,
                           antialiasing=self.aa)

    # ---- CUDA path through the operator -------------------------------------------------------------
    def cuda_run(self, device="cuda:0", backward=True, invdepth_grad=False):
        import diff_gaussian_rasterization as dgr
        dev = torch.device(device)
        s, c = self.scene.to(dev), self.cam.to(dev)
        leaves = {"means3D": s.means3D.clone().requires_grad_(backward), "opacities": s.opacities.clone().requires_grad_(backward)}
        kw = dict(shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None)
        if self.colors is not None:
            leaves["colors_precomp"] = self.colors.to(dev).clone().requires_grad_(backward)
        else:
            leaves["shs"] = s.shs.clone().requires_grad_(backward)
        if self.cov3D is not None:
            leaves["cov3D_precomp"] = self.cov3D.to(dev).clone().requires_grad_(backward)
        else:
            leaves["scales"] = s.scales.clone().requires_grad_(backward)
            leaves["rotations"] = s.rotations.clone().requires_grad_(backward)
        for k in kw:
            if k in leaves:
                kw[k] = leaves[k]
        m2d = torch.zeros(self.P, 3, device=dev, requires_grad=backward)
        rs = dgr.GaussianRasterizationSettings(self.H, self.W, c.tanfovx, c.tanfovy, self.bg.to(dev), self.mod,
                                               c.world_view_transform, c.full_proj_transform, self.D, c.camera_center,
                                               False, False, self.aa)
        color, radii, invdepth = dgr.rasterize_gaussians(leaves["means3D"], m2d, kw["shs"], kw["colors_precomp"],
                                                         leaves["opacities"], kw["scales"], kw["rotations"],
                                                         kw["cov3D_precomp"], rs)
        grads = {}
        if backward:
            loss = (color * self.dL.to(dev)).sum()
            if invdepth_grad:
                loss = loss + (invdepth * self.depth_grad_image().to(dev)).sum()
            loss.backward()
            grads = {k: v.grad for k, v in leaves.items()}
            grads["means2D"] = m2d.grad
        torch.cuda.synchronize()
        return color, radii, invdepth, grads

    def cuda_state(self, device="cuda:0", want_invdepth=False, exact=True, bin_capacity=None):
        """Forward through `_C.rasterize_gaussians` and unpack the opaque state for the integer gates."""
        import diff_gaussian_rasterization as dgr
        dev = torch.device(device)
        s, c = self.scene.to(dev), self.cam.to(dev)
        colors = self.colors.to(dev) if self.colors is not None else None
        cov = self.cov3D.to(dev) if self.cov3D is not None else None
        res = dgr._C.rasterize_gaussians(self.bg.to(dev), s.means3D, colors, s.opacities,
                                         None if cov is not None else s.scales, None if cov is not None else s.rotations,
                                         self.mod, cov, c.world_view_transform, c.full_proj_transform, c.tanfovx, c.tanfovy,
                                         self.H, self.W, None if colors is not None else s.shs, self.D, c.camera_center,
                                         False, False, antialiasing=self.aa, want_invdepth=want_invdepth, exact=exact,
                                         bin_capacity=bin_capacity)
        cap, color, radii, geom, binning, img, invd = res
        st = dgr._C.export_state(self.P, self.W, self.H, cap, radii, geom, binning, img)
        torch.cuda.synchronize()
        return st["R"], color, radii, invd, st
