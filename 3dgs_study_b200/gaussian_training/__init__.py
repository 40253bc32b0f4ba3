"""Host-side mirrors of the steps either side of the rasterizer, over the C ABI of
include/b200splat_next.h:

  loss_utils   l1_loss / ssim / training_loss      <-> /root/reference/utils/loss_utils.py, train.py:103-104
  activations  activate (exp / normalize / sigmoid)        <-> scene/gaussian_model.py:34-42, 106-127
  adam         FusedAdam, expon_lr                 <-> scene/gaussian_model.py:185-214, train.py:137-141
  ply_io       save_ply / load_ply                 <-> scene/gaussian_model.py:218-259, 267-320
  (simple_knn._C.distCUDA2 is its own drop-in package next to this one)
"""
from .activations import activate  # noqa: F401
from .adam import FusedAdam, expon_lr  # noqa: F401
from .loss_utils import l1_and_ssim, l1_loss, ssim, training_loss  # noqa: F401
from .ply_io import load_ply, save_ply  # noqa: F401
