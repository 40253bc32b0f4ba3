"""Drop-in `diff_gaussian_rasterization` operator package backed by libb200splat.so (B200 / sm_100a).

Mirrors the operator API the reference imports and calls
(/root/reference/gaussian_renderer/__init__.py:14 import, :47-60 settings, :62 module, :98-106 call):

    GaussianRasterizationSettings(image_height, image_width, tanfovx, tanfovy, bg, scale_modifier,
                                  viewmatrix, projmatrix, sh_degree, campos, prefiltered, debug
                                  [, antialiasing=False])
    GaussianRasterizer(raster_settings).forward(means3D, means2D, opacities, shs=None,
        colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None)

Return arity: BASELINE.json's north_star asks for (color, radii, depth); the reference's
own call site unpacks two values (:98).  One switch selects: `LEGACY_OUTPUTS` (module attribute, or env
DGR_LEGACY_OUTPUTS=1) -> (color, radii); default -> (color, radii, invdepth).

Autograd: gradients flow to means3D, shs / colors_precomp, opacities, scales, rotations /
cov3D_precomp; `means2D.grad` receives [P,3] with xy = dL/d(NDC xy), z = 0 -- the tensor
/root/reference/scene/gaussian_model.py:576-580 reads for densification.
"""
from __future__ import annotations

import os
from typing import NamedTuple, Optional

import torch
import torch.nn as nn

from . import _C

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "LEGACY_OUTPUTS"]

LEGACY_OUTPUTS = os.environ.get("DGR_LEGACY_OUTPUTS", "0") not in ("0", "", "false", "False")


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    antialiasing: bool = False


def cpu_deep_copy_tuple(input_tuple):
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, want_invdepth: Optional[bool] = None):
    """-> (color, radii, invdepth).  `want_invdepth=False` (the default in LEGACY_OUTPUTS mode, i.e. for the
    reference's own two-output call site) skips the inverse-depth channel: invdepth is then None."""
    if want_invdepth is None:
        want_invdepth = not LEGACY_OUTPUTS
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings, bool(want_invdepth))


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings, want_invdepth=True):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
                rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        aa = bool(getattr(rs, "antialiasing", False))
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)  # copy first: a failing kernel may corrupt them
            try:
                res = _C.rasterize_gaussians(*args, antialiasing=aa, want_invdepth=want_invdepth)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            res = _C.rasterize_gaussians(*args, antialiasing=aa, want_invdepth=want_invdepth)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepth = res
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.has_means2D = means2D is not None
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities,
                              geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)      # invdepth is differentiable (newer upstream API)
        ctx.set_materialize_grads(False)        # an unused output arrives as None, not as an image of zeros
        return color, radii, invdepth

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii, grad_invdepth):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, sh, rs.sh_degree, rs.campos,
                geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer, rs.debug)
        kw = dict(opacities=opacities, antialiasing=bool(getattr(rs, "antialiasing", False)),
                  image_height=rs.image_height, image_width=rs.image_width, dL_dout_invdepth=grad_invdepth)
        if grad_out_color is None:               # only the inverse-depth image fed the loss
            grad_out_color = torch.zeros(3, rs.image_height, rs.image_width, dtype=torch.float32, device=means3D.device)
            args = args[:12] + (grad_out_color,) + args[13:]
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                out = _C.rasterize_gaussians_backward(*args, **kw)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            out = _C.rasterize_gaussians_backward(*args, **kw)
        g_means2D, g_colors, g_opac, g_means3D, g_cov3D, g_sh, g_scales, g_rot = out
        if g_opac is not None and opacities is not None:
            g_opac = g_opac.reshape(opacities.shape)
        if not ctx.has_means2D:
            g_means2D = None
        return (g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rot, g_cov3D, None, None)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings: GaussianRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions: torch.Tensor) -> torch.Tensor:
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        out = rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                  self.raster_settings)
        color, radii, invdepth = out
        if LEGACY_OUTPUTS:
            return color, radii
        return color, radii, invdepth
