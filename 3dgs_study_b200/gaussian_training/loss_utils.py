"""Fused L1 + SSIM loss: the host-side mirror of /root/reference/utils/loss_utils.py for the two functions the
training step calls (train.py:103-104) -- same names, arguments and results:

    l1_loss(network_output, gt)                                 loss_utils.py:17-24
    ssim(img1, img2, window_size=11, size_average=True)         loss_utils.py:59-108

plus `l1_and_ssim(img, gt)` (both values from ONE forward kernel and ONE backward kernel) and
`training_loss(img, gt, lambda_dssim)` = (1-l)*L1 + l*(1-SSIM), the expression of train.py:104.

All arithmetic runs in libb200splat.so (b200splat_l1_ssim_forward / _backward, include/b200splat_next.h);
PyTorch supplies device memory, the stream and autograd bookkeeping.  CUDA fp32 tensors only, no fallback.
"""
from __future__ import annotations

import ctypes
from math import exp

import torch

from . import _lib

_TAPS = {}


def window_taps(window_size: int = 11, sigma: float = 1.5):
    """The reference's 1-D window (`gaussian`, loss_utils.py:35-45: exp in double, float32 tensor, divided by its
    float32 sum), centred in 11 taps.  Host memory (ctypes float array); built with the same torch ops so the
    taps are bit-identical to the reference's."""
    key = (int(window_size), float(sigma))
    if key not in _TAPS:
        if window_size % 2 != 1 or not (1 <= window_size <= 11):
            raise ValueError("window_size must be odd and <= 11 (the fused kernels hold an 11-tap window)")
        g = torch.Tensor([exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)])
        g = (g / g.sum()).tolist()
        taps = [0.0] * 11
        o = (11 - window_size) // 2
        taps[o:o + window_size] = g
        _TAPS[key] = (ctypes.c_float * 11)(*taps)
    return _TAPS[key]


def _prep(img1: torch.Tensor, img2: torch.Tensor):
    if not (img1.is_cuda and img2.is_cuda):
        raise RuntimeError("fused L1/SSIM loss needs CUDA tensors (there is no CPU path)")
    if img1.shape != img2.shape or img1.dim() not in (3, 4):
        raise RuntimeError(f"expected two [C,H,W] or [B,C,H,W] tensors of equal shape, got {tuple(img1.shape)} and {tuple(img2.shape)}")
    if img1.device != img2.device:
        raise RuntimeError("both images must be on the same device")
    return img1.float().contiguous(), img2.float().contiguous()


class _L1SSIM(torch.autograd.Function):
    """(img, gt) -> per-plane sums [planes, 2] = (sum |img-gt|, sum of the SSIM map).  Differentiable in img."""

    @staticmethod
    def forward(ctx, img, gt, window_size):
        a, b = _prep(img.detach(), gt.detach())
        H, W = a.shape[-2:]
        planes = a.numel() // (H * W)
        dev = a.device
        L = _lib.lib()
        taps = window_taps(window_size)
        need_grad = bool(ctx.needs_input_grad[0])
        maps = torch.empty((3,) + tuple(a.shape), dtype=torch.float32, device=dev) if need_grad else None
        partials = torch.empty((L.b200splat_l1_ssim_partials(planes, H, W), 2), dtype=torch.float32, device=dev)
        out = torch.empty((planes, 2), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(L.b200splat_l1_ssim_forward(planes, H, W, a.data_ptr(), b.data_ptr(),
                                                   ctypes.addressof(taps), maps.data_ptr() if need_grad else None,
                                                   partials.data_ptr(), out.data_ptr(), _lib.stream_ptr(dev)), "l1_ssim forward")
        ctx.save_for_backward(a, b, maps)
        ctx.window_size = window_size
        ctx.in_dtype = img.dtype
        return out

    @staticmethod
    def backward(ctx, g_out):
        a, b, maps = ctx.saved_tensors
        if maps is None:
            return None, None, None
        H, W = a.shape[-2:]
        planes = a.numel() // (H * W)
        dev = a.device
        scale = g_out.to(torch.float32).contiguous()      # upstream gradients of the per-plane sums
        da = torch.empty_like(a)
        taps = window_taps(ctx.window_size)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().b200splat_l1_ssim_backward(planes, H, W, a.data_ptr(), b.data_ptr(),
                                                             ctypes.addressof(taps), maps.data_ptr(), scale.data_ptr(),
                                                             da.data_ptr(), _lib.stream_ptr(dev)), "l1_ssim backward")
        return da.to(ctx.in_dtype), None, None


class _TrainingLoss(torch.autograd.Function):
    """(img, gt) -> [3] = (loss, Ll1, ssim) with loss = (1-l) Ll1 + l (1 - ssim): two kernels forward, one
    backward, no elementwise glue.  Only `loss` (element 0) carries a gradient."""

    @staticmethod
    def forward(ctx, img, gt, lambda_dssim):
        a, b = _prep(img.detach(), gt.detach())
        H, W = a.shape[-2:]
        planes = a.numel() // (H * W)
        dev = a.device
        L = _lib.lib()
        taps = window_taps(11)
        need_grad = bool(ctx.needs_input_grad[0])
        maps = torch.empty((3,) + tuple(a.shape), dtype=torch.float32, device=dev) if need_grad else None
        partials = torch.empty((L.b200splat_l1_ssim_partials(planes, H, W), 2), dtype=torch.float32, device=dev)
        out = torch.empty(3, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(L.b200splat_training_loss_forward(planes, H, W, a.data_ptr(), b.data_ptr(), ctypes.addressof(taps),
                                                         float(lambda_dssim), maps.data_ptr() if need_grad else None,
                                                         partials.data_ptr(), out.data_ptr(), _lib.stream_ptr(dev)), "training_loss forward")
        ctx.save_for_backward(a, b, maps)
        ctx.lambda_dssim = float(lambda_dssim)
        ctx.in_dtype = img.dtype
        return out

    @staticmethod
    def backward(ctx, g_out):
        a, b, maps = ctx.saved_tensors
        if maps is None:
            return None, None, None
        H, W = a.shape[-2:]
        planes = a.numel() // (H * W)
        dev = a.device
        up = g_out.to(torch.float32).contiguous()     # [3]; element 0 = d L / d loss (read on the device)
        da = torch.empty_like(a)
        taps = window_taps(11)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().b200splat_training_loss_backward(planes, H, W, a.data_ptr(), b.data_ptr(), ctypes.addressof(taps),
                                                                   ctx.lambda_dssim, maps.data_ptr(), up.data_ptr(), da.data_ptr(),
                                                                   _lib.stream_ptr(dev)), "training_loss backward")
        return da.to(ctx.in_dtype), None, None


def l1_and_ssim(img: torch.Tensor, gt: torch.Tensor, window_size: int = 11, size_average: bool = True):
    """(l1_loss(img, gt), ssim(img, gt, window_size, size_average)) from one pass over the images."""
    out = _L1SSIM.apply(img, gt, int(window_size))
    n = img.numel()
    l1 = out[:, 0].sum() / n                  # .mean() of the reference: one division by the element count
    if size_average:
        return l1, out[:, 1].sum() / n
    if img.dim() != 4:
        # the reference's ssim_map.mean(1).mean(1).mean(1) needs a batch dimension too (loss_utils.py:108)
        raise RuntimeError("size_average=False needs a [B,C,H,W] input")
    return l1, out[:, 1].view(img.shape[0], img.shape[1]).sum(1) / (n // img.shape[0])


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    return l1_and_ssim(network_output, gt)[0]


def ssim(img1: torch.Tensor, img2: torch.Tensor, window_size: int = 11, size_average: bool = True) -> torch.Tensor:
    return l1_and_ssim(img1, img2, window_size, size_average)[1]


def training_loss(image: torch.Tensor, gt_image: torch.Tensor, lambda_dssim: float = 0.2):
    """-> (loss, Ll1): loss = (1 - lambda) * Ll1 + lambda * (1 - ssim(image, gt)), train.py:103-104
    (lambda_dssim default: arguments/__init__.py:89)."""
    out = _TrainingLoss.apply(image, gt_image, float(lambda_dssim))
    return out[0], out[1].detach()
