"""3dgs_study_b200 -- B200-native (sm_100a) differentiable Gaussian-splat rasterizer.

Holds only what the hot path needs:
  csrc/                         hand-written CUDA kernels + the C ABI (include/b200splat.h)
  libb200splat.so               built by __graft_entry__.build() / `make -C 3dgs_study_b200/csrc`
  diff_gaussian_rasterization/  the drop-in operator package (same names as the reference's op)
  synth.py                      deterministic synthetic scenes + reference-convention cameras
  view_parallel.py              view-parallel multi-GPU step (torch.distributed, NCCL / gloo)

The directory name starts with a digit, so import it with importlib
(`importlib.import_module("3dgs_study_b200")`) or put it on sys.path and
`import diff_gaussian_rasterization` -- `ensure_on_path()` does the latter.
"""
import os
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))


def ensure_on_path() -> str:
    """Make `import diff_gaussian_rasterization` resolve to this package's drop-in operator."""
    if PKG_DIR not in sys.path:
        sys.path.insert(0, PKG_DIR)
    return PKG_DIR


ensure_on_path()
