"""ctypes front-end for oracle/next_oracle.c  --  TEST INFRASTRUCTURE ONLY.

CPU restatements of the steps either side of the rasterizer: L1 + SSIM loss
(/root/reference/utils/loss_utils.py:17-24,35-108), the Adam step over the Gaussian parameters
(/root/reference/scene/gaussian_model.py:185-214, train.py:137-141), distCUDA2
(/root/reference/submodules/simple-knn/simple_knn.cu:46-222) and the Gaussian PLY file
(/root/reference/scene/gaussian_model.py:218-259,267-320; numpy, below).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libnext_oracle.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(_HERE, "next_oracle.c")
        if not os.path.exists(_LIB_PATH) or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(_LIB_PATH)):
            subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "libnext_oracle.so"])
        _lib = ctypes.CDLL(_LIB_PATH)
    return _lib


def _p(a):
    assert a.flags["C_CONTIGUOUS"]
    return ctypes.c_void_p(a.ctypes.data)


def window_taps(window_size: int = 11, sigma: float = 1.5) -> np.ndarray:
    """The 1-D window of /root/reference/utils/loss_utils.py:35-45: exp in double, stored as float32
    (torch.Tensor), normalised by its float32 sum; centred in 11 taps (zeros outside)."""
    assert window_size % 2 == 1 and 1 <= window_size <= 11
    import torch  # the reference builds the taps with torch (float32 division by torch's float32 sum)
    t = torch.tensor([math.exp(-(x - window_size // 2) ** 2 / (2 * sigma ** 2)) for x in range(window_size)], dtype=torch.float32)
    g = (t / t.sum()).numpy()
    out = np.zeros(11, np.float32)
    o = (11 - window_size) // 2
    out[o:o + window_size] = g
    return out


def l1_ssim_forward(a: np.ndarray, b: np.ndarray, window_size: int = 11, want_maps: bool = True):
    """a, b: [..., H, W] float32.  Returns (l1_sum[planes], ssim_sum[planes], (dmu1, dE1, dE12) | None)."""
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    H, W = a.shape[-2:]
    planes = a.size // (H * W)
    w = window_taps(window_size)
    l1 = np.zeros(planes, np.float64)
    ss = np.zeros(planes, np.float64)
    maps = tuple(np.zeros_like(a) for _ in range(3)) if want_maps else (None, None, None)
    lib().orc_l1_ssim_forward(planes, H, W, _p(a), _p(b), _p(w), _p(l1), _p(ss),
                              *(ctypes.c_void_p(0) if m is None else _p(m) for m in maps))
    return l1, ss, (maps if want_maps else None)


def l1_ssim_backward(a, b, maps, scale, window_size: int = 11) -> np.ndarray:
    """scale: [planes, 2] float32 = (factor of sign(a-b), factor of the SSIM-map gradient) per plane."""
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    H, W = a.shape[-2:]
    planes = a.size // (H * W)
    scale = np.ascontiguousarray(scale, np.float32).reshape(planes, 2)
    w = window_taps(window_size)
    da = np.zeros_like(a)
    lib().orc_l1_ssim_backward(planes, H, W, _p(a), _p(b), _p(w), _p(maps[0]), _p(maps[1]), _p(maps[2]), _p(scale), _p(da))
    return da


def loss_and_grad(a, b, lambda_dssim: float = 0.2, window_size: int = 11):
    """The training loss of /root/reference/train.py:103-104 and d loss / d a.  Returns (loss, l1, ssim, grad)."""
    l1s, sss, maps = l1_ssim_forward(a, b, window_size)
    n = a.size
    l1 = float(l1s.sum() / n)
    ssim = float(sss.sum() / n)
    planes = l1s.size
    scale = np.tile(np.array([(1.0 - lambda_dssim) / n, -lambda_dssim / n], np.float32), (planes, 1))
    return (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - ssim), l1, ssim, l1_ssim_backward(a, b, maps, scale, window_size)


def adam_step(p, g, m, v, lr, step, beta1=0.9, beta2=0.999, eps=1e-15):
    """In place on float32 arrays p, m, v (torch.optim.Adam, the reference's settings: gaussian_model.py:197)."""
    for x in (p, g, m, v):
        assert x.dtype == np.float32 and x.flags["C_CONTIGUOUS"]
    f = ctypes.c_double
    lib().orc_adam_step(ctypes.c_int64(p.size), _p(p), _p(g), _p(m), _p(v), f(lr), f(beta1), f(beta2), f(eps), ctypes.c_int(step))


def activate_forward(scaling, rotation, opacity):
    s, r, o = (np.ascontiguousarray(x, np.float32) for x in (scaling, rotation, opacity))
    P = s.shape[0]
    out = (np.zeros_like(s), np.zeros_like(r), np.zeros_like(o))
    lib().orc_activate_forward(ctypes.c_int(P), _p(s), _p(r), _p(o), _p(out[0]), _p(out[1]), _p(out[2]))
    return out


def activate_backward(rotation, scales, opacities, g_scales, g_rotations, g_opacities):
    a = [np.ascontiguousarray(x, np.float32) for x in (rotation, scales, opacities, g_scales, g_rotations, g_opacities)]
    P = a[0].shape[0]
    out = (np.zeros_like(a[1]), np.zeros_like(a[0]), np.zeros_like(a[2]))
    lib().orc_activate_backward(ctypes.c_int(P), *(_p(x) for x in a), _p(out[0]), _p(out[1]), _p(out[2]))
    return out


def adam_step_rows(p, g, m, v, lr, lr_tail, row_width, row_split, step, beta1=0.9, beta2=0.999, eps=1e-15):
    """Adam on a [rows, row_width] tensor whose columns >= row_split use lr_tail (FusedAdam's row_width / row_split /
    lr_tail group keys): restated as two plain Adam steps on the two column blocks."""
    P = p.size // row_width
    views = [x.reshape(P, row_width) for x in (p, g, m, v)]
    for lo, hi, rate in ((0, row_split, lr), (row_split, row_width, lr_tail)):
        if hi > lo:
            blk = [np.ascontiguousarray(x[:, lo:hi]) for x in views]
            adam_step(blk[0], blk[1], blk[2], blk[3], rate, step, beta1, beta2, eps)
            for x, b in zip((views[0], views[2], views[3]), (blk[0], blk[2], blk[3])):
                x[:, lo:hi] = b


def dist2_knn3(xyz: np.ndarray) -> np.ndarray:
    xyz = np.ascontiguousarray(xyz, np.float32)
    out = np.zeros(xyz.shape[0], np.float32)
    lib().orc_dist2_knn3(ctypes.c_int(xyz.shape[0]), _p(xyz), _p(out))
    return out


# ---- Gaussian PLY file (numpy restatement) -----------------------------------------------------------
def ply_attribute_names(n_rest: int):
    """construct_list_of_attributes, /root/reference/scene/gaussian_model.py:218-238 (f_dc 3, f_rest
    3*(M-1) channel-major, 3 scales, 4 rotations)."""
    l = ["x", "y", "z", "nx", "ny", "nz"]
    l += [f"f_dc_{i}" for i in range(3)]
    l += [f"f_rest_{i}" for i in range(n_rest)]
    l.append("opacity")
    l += [f"scale_{i}" for i in range(3)]
    l += [f"rot_{i}" for i in range(4)]
    return l


def ply_bytes(xyz, f_dc, f_rest, opacity, scale, rot) -> bytes:
    """save_ply, /root/reference/scene/gaussian_model.py:240-259: f_dc [P,1,3] / f_rest [P,M-1,3] are written
    transposed(1,2) and flattened (channel-major), normals are zeros, all properties f4 little-endian, in
    the binary_little_endian 1.0 layout plyfile (the absent third-party writer) produces for one `vertex`
    element built from a structured array."""
    P = xyz.shape[0]
    fdc = np.ascontiguousarray(np.transpose(f_dc, (0, 2, 1))).reshape(P, 3 * f_dc.shape[1])
    frest = np.ascontiguousarray(np.transpose(f_rest, (0, 2, 1))).reshape(P, 3 * f_rest.shape[1])
    attrs = np.concatenate([xyz, np.zeros_like(xyz), fdc, frest, opacity.reshape(P, 1), scale, rot], axis=1).astype("<f4")
    names = ply_attribute_names(frest.shape[1])
    assert attrs.shape[1] == len(names)
    hdr = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P
    hdr += "".join(f"property float {n}\n" for n in names) + "end_header\n"
    return hdr.encode("ascii") + attrs.tobytes()
