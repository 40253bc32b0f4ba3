"""`simple_knn._C.distCUDA2`, the one function the reference uses from its simple-knn submodule
(/root/reference/submodules/simple-knn/spatial.cu:15-26; call: scene/gaussian_model.py:153):
points [P,3] (CUDA) -> [P] mean squared distance to the 3 nearest other points.  Implemented by
b200splat_dist2_knn3 (csrc/next_steps.cu); CUDA fp32 only, no fallback."""
from __future__ import annotations

import torch

from gaussian_training import _lib


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    if not points.is_cuda:
        raise RuntimeError("distCUDA2 needs a CUDA tensor (there is no CPU path)")
    if points.dim() != 2 or points.shape[1] != 3:
        raise RuntimeError("points must have dimensions (num_points, 3)")
    pts = points.detach().float().contiguous()
    P = pts.shape[0]
    out = torch.zeros(P, dtype=torch.float32, device=pts.device)     # torch::full({P}, 0.0), spatial.cu:21
    if P:
        with torch.cuda.device(pts.device):
            _lib.check(_lib.lib().b200splat_dist2_knn3(P, pts.data_ptr(), out.data_ptr(), _lib.stream_ptr(pts.device)), "dist2_knn3")
    return out


KNN_STATIC_MAX_POINTS = 262144      # B200SPLAT_KNN_STATIC_MAX_POINTS (include/b200splat_next.h)


def distCUDA2_static(points: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """Not part of the reference's module: distCUDA2 for callers that cannot allocate (b200splat_dist2_knn3_static).
    `points` must already be a contiguous fp32 CUDA [P,3] tensor and `out` a contiguous fp32 [P] tensor on the same device;
    nothing is allocated here or in the library (the scratch is a 12 MiB array inside the library, shared by all calls on a
    GPU: do not overlap calls).  More than KNN_STATIC_MAX_POINTS points raise."""
    if not (points.is_cuda and out.is_cuda and points.device == out.device):
        raise RuntimeError("distCUDA2_static needs CUDA tensors on one device")
    if points.dtype != torch.float32 or out.dtype != torch.float32 or not points.is_contiguous() or not out.is_contiguous():
        raise RuntimeError("distCUDA2_static needs contiguous fp32 tensors")
    if points.dim() != 2 or points.shape[1] != 3 or out.shape != (points.shape[0],):
        raise RuntimeError("points must have dimensions (num_points, 3) and out (num_points,)")
    with torch.cuda.device(points.device):
        _lib.check(_lib.lib().b200splat_dist2_knn3_static(points.shape[0], points.data_ptr(), out.data_ptr(),
                                                          _lib.stream_ptr(points.device)), "dist2_knn3_static")
    return out
