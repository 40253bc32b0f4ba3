"""Activations of the raw Gaussian parameters in one kernel each way (b200splat_activate_forward / _backward):
scales = exp(_scaling), rotations = normalize(_rotation), opacities = sigmoid(_opacity) -- what
GaussianModel.get_scaling / get_rotation / get_opacity compute with separate torch ops
(/root/reference/scene/gaussian_model.py:34-42, 106-127).  CUDA fp32 only, no fallback."""
from __future__ import annotations

import torch

from . import _lib


def _c(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (there is no CPU path)")
    return t.detach().float().contiguous()


class _Activate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scaling, rotation, opacity):
        s, r, o = _c(scaling, "scaling"), _c(rotation, "rotation"), _c(opacity, "opacity")
        P = s.shape[0]
        if s.shape != (P, 3) or r.shape != (P, 4) or o.numel() != P:
            raise RuntimeError("expected scaling [P,3], rotation [P,4], opacity [P,1]")
        scales, rotations, opacities = torch.empty_like(s), torch.empty_like(r), torch.empty_like(o)
        dev = s.device
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().b200splat_activate_forward(P, s.data_ptr(), r.data_ptr(), o.data_ptr(), scales.data_ptr(),
                                                             rotations.data_ptr(), opacities.data_ptr(), _lib.stream_ptr(dev)),
                       "activate forward")
        ctx.save_for_backward(r, scales, opacities)
        return scales, rotations, opacities

    @staticmethod
    def backward(ctx, g_s, g_r, g_o):
        r, scales, opacities = ctx.saved_tensors
        P = r.shape[0]
        dev = r.device
        prep = lambda g: None if g is None else g.float().contiguous()
        g_s, g_r, g_o = prep(g_s), prep(g_r), prep(g_o)
        d_s = torch.empty_like(scales) if g_s is not None else None
        d_r = torch.empty_like(r) if g_r is not None else None
        d_o = torch.empty_like(opacities) if g_o is not None else None
        ptr = lambda t: None if t is None else t.data_ptr()
        if P:
            with torch.cuda.device(dev):
                _lib.check(_lib.lib().b200splat_activate_backward(P, r.data_ptr(), scales.data_ptr(), opacities.data_ptr(), ptr(g_s),
                                                                  ptr(g_r), ptr(g_o), ptr(d_s), ptr(d_r), ptr(d_o),
                                                                  _lib.stream_ptr(dev)), "activate backward")
        return d_s, d_r, d_o


def activate(scaling: torch.Tensor, rotation: torch.Tensor, opacity: torch.Tensor):
    """-> (get_scaling, get_rotation, get_opacity) of the reference, differentiable in all three inputs."""
    return _Activate.apply(scaling, rotation, opacity)
