"""Gaussian PLY file: the wire format of GaussianModel.save_ply / load_ply
(/root/reference/scene/gaussian_model.py:218-259, 267-320) without the `plyfile` package (absent from this image).
The codec is native host code in libb200splat.so (csrc/ply_io.cu: b200splat_ply_write / _read); tensors may live
on any device and come back on the CPU (the reference moves them to "cuda" itself, :309-317)."""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

from . import _lib


def _host(t, shape):
    a = t.detach().to("cpu", torch.float32).contiguous().numpy() if torch.is_tensor(t) else np.ascontiguousarray(t, np.float32)
    if a.size != int(np.prod(shape)):
        raise ValueError(f"expected {shape}, got {a.shape}")
    return a.reshape(shape)


def save_ply(path: str, xyz, features_dc, features_rest, opacity, scaling, rotation) -> None:
    """Raw (pre-activation) values, as the reference stores them: xyz [P,3], features_dc [P,1,3],
    features_rest [P,M-1,3], opacity [P,1], scaling [P,3], rotation [P,4].  Creates the directory like
    save_ply's mkdir_p (:241)."""
    P = int(xyz.shape[0])
    K = int(features_rest.shape[1]) if features_rest is not None and features_rest.dim() == 3 else 0
    a = [_host(xyz, (P, 3)), _host(features_dc, (P, 3)),
         _host(features_rest, (P, K * 3)) if K else np.zeros((P, 0), np.float32), _host(opacity, (P,)),
         _host(scaling, (P, 3)), _host(rotation, (P, 4))]
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    ptr = lambda x: x.ctypes.data if x.size else None
    _lib.check(_lib.lib().b200splat_ply_write(os.fsencode(path), P, 3 * K, *(ptr(x) for x in a)), "ply_write")


def load_ply(path: str):
    """-> dict(xyz [P,3], features_dc [P,1,3], features_rest [P,M-1,3], opacity [P,1], scaling [P,3],
    rotation [P,4]) of CPU fp32 tensors, in the layouts load_ply builds (:279-317)."""
    L = _lib.lib()
    P = ctypes.c_int64(0)
    n_rest = ctypes.c_int32(0)
    _lib.check(L.b200splat_ply_read_header(os.fsencode(path), ctypes.byref(P), ctypes.byref(n_rest)), "ply_read_header")
    P, n_rest = int(P.value), int(n_rest.value)
    if n_rest % 3 != 0:
        raise RuntimeError(f"{path}: {n_rest} f_rest properties (not a multiple of 3)")
    K = n_rest // 3
    out = {"xyz": np.empty((P, 3), np.float32), "features_dc": np.empty((P, 1, 3), np.float32),
           "features_rest": np.empty((P, K, 3), np.float32), "opacity": np.empty((P, 1), np.float32),
           "scaling": np.empty((P, 3), np.float32), "rotation": np.empty((P, 4), np.float32)}
    ptr = lambda x: x.ctypes.data if x.size else None
    _lib.check(L.b200splat_ply_read(os.fsencode(path), P, n_rest, *(ptr(out[k]) for k in
                                    ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation"))), "ply_read")
    return {k: torch.from_numpy(v) for k, v in out.items()}


def sh_degree_of(features_rest) -> int:
    """max SH degree from the number of rest coefficients: (D+1)^2 - 1 = K (cf. the assert at :287)."""
    K = int(features_rest.shape[1])
    D = int(round((K + 1) ** 0.5)) - 1
    if (D + 1) ** 2 - 1 != K:
        raise ValueError(f"{K} rest coefficients do not form a full SH band set")
    return D
