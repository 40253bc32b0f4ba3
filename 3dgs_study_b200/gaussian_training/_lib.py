"""ctypes signatures of include/b200splat_next.h on the library `_C.lib()` loads.  No fallback: a missing
libb200splat.so raises in `_C.lib()`."""
from __future__ import annotations

import ctypes

from diff_gaussian_rasterization import _C

ADAM_MAX_GROUPS = 8


class AdamGroup(ctypes.Structure):
    _fields_ = [("param", ctypes.c_void_p), ("grad", ctypes.c_void_p), ("exp_avg", ctypes.c_void_p),
                ("exp_avg_sq", ctypes.c_void_p), ("n", ctypes.c_int64), ("lr", ctypes.c_double),
                ("step", ctypes.c_int32), ("row_width", ctypes.c_int32), ("row_split", ctypes.c_int32),
                ("reserved", ctypes.c_int32), ("lr_tail", ctypes.c_double)]


_ready = False


def lib():
    global _ready
    L = _C.lib()
    if not _ready:
        vp, i32, i64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64
        L.b200splat_l1_ssim_partials.restype = ctypes.c_size_t
        L.b200splat_l1_ssim_partials.argtypes = [i32, i32, i32]
        L.b200splat_l1_ssim_forward.restype = ctypes.c_int
        L.b200splat_l1_ssim_forward.argtypes = [i32, i32, i32, vp, vp, vp, vp, vp, vp, vp]
        L.b200splat_l1_ssim_backward.restype = ctypes.c_int
        L.b200splat_l1_ssim_backward.argtypes = [i32, i32, i32, vp, vp, vp, vp, vp, vp, vp]
        L.b200splat_training_loss_forward.restype = ctypes.c_int
        L.b200splat_training_loss_forward.argtypes = [i32, i32, i32, vp, vp, vp, ctypes.c_float, vp, vp, vp, vp]
        L.b200splat_training_loss_backward.restype = ctypes.c_int
        L.b200splat_training_loss_backward.argtypes = [i32, i32, i32, vp, vp, vp, ctypes.c_float, vp, vp, vp, vp]
        L.b200splat_activate_forward.restype = ctypes.c_int
        L.b200splat_activate_forward.argtypes = [i32, vp, vp, vp, vp, vp, vp, vp]
        L.b200splat_activate_backward.restype = ctypes.c_int
        L.b200splat_activate_backward.argtypes = [i32] + [vp] * 10
        L.b200splat_adam_step.restype = ctypes.c_int
        L.b200splat_adam_step.argtypes = [i32, vp, ctypes.c_double, ctypes.c_double, ctypes.c_double, vp]
        L.b200splat_dist2_knn3.restype = ctypes.c_int
        L.b200splat_dist2_knn3.argtypes = [i32, vp, vp, vp]
        L.b200splat_dist2_knn3_static.restype = ctypes.c_int
        L.b200splat_dist2_knn3_static.argtypes = [i32, vp, vp, vp]
        L.b200splat_ply_write.restype = ctypes.c_int
        L.b200splat_ply_write.argtypes = [ctypes.c_char_p, i64, i32, vp, vp, vp, vp, vp, vp]
        L.b200splat_ply_read_header.restype = ctypes.c_int
        L.b200splat_ply_read_header.argtypes = [ctypes.c_char_p, ctypes.POINTER(i64), ctypes.POINTER(i32)]
        L.b200splat_ply_read.restype = ctypes.c_int
        L.b200splat_ply_read.argtypes = [ctypes.c_char_p, i64, i32, vp, vp, vp, vp, vp, vp]
        _ready = True
    return L


check = _C._check
stream_ptr = _C._stream_ptr
