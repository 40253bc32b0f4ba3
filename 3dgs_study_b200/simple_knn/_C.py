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
