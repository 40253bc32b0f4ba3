"""ctypes front-end for oracle/splat_oracle.c  --  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package never does (tests/test_no_oracle_in_product.py checks).

All arrays are numpy, C-contiguous.  Stage functions mirror the stages of the hot path
(SURVEY.md section 8a): preprocess -> bin -> render -> render_backward -> preprocess_backward.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import NamedTuple, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libsplat_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the C oracle if its .so is missing or older than the source."""
    src = os.path.join(_HERE, "splat_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or (
        os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(_LIB_PATH)
    ):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "libsplat_oracle.so"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.orc_bin.restype = ctypes.c_int64
    return _lib


def _p(a: Optional[np.ndarray]):
    if a is None:
        return ctypes.c_void_p(0)
    assert a.flags["C_CONTIGUOUS"], "oracle arrays must be C-contiguous"
    return ctypes.c_void_p(a.ctypes.data)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


class Geom(NamedTuple):
    radii: np.ndarray          # [P] i32
    means2D: np.ndarray        # [P,2] f32 pixel coords
    depths: np.ndarray         # [P] f32
    cov3D: np.ndarray          # [P,6]
    conic_opacity: np.ndarray  # [P,4] (A,B,C,opacity_eff)
    rgb: np.ndarray            # [P,3]
    clamped: np.ndarray        # [P,3] u8
    tiles_touched: np.ndarray  # [P] u32


class Binning(NamedTuple):
    keys: np.ndarray    # [R] u64 sorted
    vals: np.ndarray    # [R] u32 sorted
    ranges: np.ndarray  # [T,2] i32


def preprocess(P, D, M, W, H, means3D, scales, mod, rotations, opacities, shs, colors_precomp,
               cov3D_precomp, view, proj, campos, tanfovx, tanfovy, antialiasing=False) -> Geom:
    means3D, scales, rotations = _f32(means3D), _f32(scales), _f32(rotations)
    opacities, shs, colors_precomp = _f32(opacities), _f32(shs), _f32(colors_precomp)
    cov3D_precomp, view, proj, campos = _f32(cov3D_precomp), _f32(view), _f32(proj), _f32(campos)
    g = Geom(np.zeros(P, np.int32), np.zeros((P, 2), np.float32), np.zeros(P, np.float32),
             np.zeros((P, 6), np.float32), np.zeros((P, 4), np.float32), np.zeros((P, 3), np.float32),
             np.zeros((P, 3), np.uint8), np.zeros(P, np.uint32))
    lib().orc_preprocess(
        ctypes.c_int(P), ctypes.c_int(D), ctypes.c_int(M), ctypes.c_int(W), ctypes.c_int(H),
        _p(means3D), _p(scales), ctypes.c_float(mod), _p(rotations), _p(opacities), _p(shs),
        _p(colors_precomp), _p(cov3D_precomp), _p(view), _p(proj), _p(campos),
        ctypes.c_float(tanfovx), ctypes.c_float(tanfovy), ctypes.c_int(int(antialiasing)),
        _p(g.radii), _p(g.means2D), _p(g.depths), _p(g.cov3D), _p(g.conic_opacity), _p(g.rgb),
        _p(g.clamped), _p(g.tiles_touched))
    return g


def bin_tiles(P, W, H, g: Geom) -> Binning:
    R = int(g.tiles_touched.astype(np.int64).sum())
    T = ((W + 15) // 16) * ((H + 15) // 16)
    keys = np.zeros(max(R, 1), np.uint64)
    vals = np.zeros(max(R, 1), np.uint32)
    ranges = np.zeros((T, 2), np.int32)
    n = lib().orc_bin(ctypes.c_int(P), ctypes.c_int(W), ctypes.c_int(H), _p(g.means2D), _p(g.depths),
                      _p(g.radii), _p(keys), _p(vals), ctypes.c_int64(R), _p(ranges))
    assert n == R, (n, R)
    return Binning(keys[:R], vals[:R], ranges)


def render(W, H, g: Geom, b: Binning, bg, want_invdepth=False, tile_stride=1, tile_offset=0):
    bg = _f32(bg)
    color = np.zeros((3, H, W), np.float32)
    invd = np.zeros((1, H, W), np.float32) if want_invdepth else None
    final_T = np.zeros((H, W), np.float32)
    n_contrib = np.zeros((H, W), np.uint32)
    vals = b.vals if b.vals.size else np.zeros(1, np.uint32)
    lib().orc_render(ctypes.c_int(W), ctypes.c_int(H), _p(b.ranges), _p(vals), _p(g.means2D),
                     _p(g.conic_opacity), _p(g.rgb), _p(g.depths), _p(bg), _p(color), _p(invd),
                     _p(final_T), _p(n_contrib), ctypes.c_int(tile_stride), ctypes.c_int(tile_offset))
    return color, invd, final_T, n_contrib


class RenderGrads(NamedTuple):
    dL_dmean2D: np.ndarray   # [P,2], NDC units
    dL_dconic: np.ndarray    # [P,3], true partials wrt (A,B,C)
    dL_dopacity: np.ndarray  # [P]
    dL_dcolor: np.ndarray    # [P,3]
    dL_dinvd: np.ndarray     # [P]


def render_backward(P, W, H, g: Geom, b: Binning, bg, final_T, n_contrib, dL_dpix, dL_dinvdepth=None,
                    tile_stride=1, tile_offset=0) -> RenderGrads:
    bg, dL_dpix, dL_dinvdepth = _f32(bg), _f32(dL_dpix), _f32(dL_dinvdepth)
    out = RenderGrads(np.zeros((P, 2), np.float32), np.zeros((P, 3), np.float32), np.zeros(P, np.float32),
                      np.zeros((P, 3), np.float32), np.zeros(P, np.float32))
    vals = b.vals if b.vals.size else np.zeros(1, np.uint32)
    lib().orc_render_backward(ctypes.c_int(P), ctypes.c_int(W), ctypes.c_int(H), _p(b.ranges), _p(vals),
                              _p(g.means2D), _p(g.conic_opacity), _p(g.rgb), _p(g.depths), _p(bg),
                              _p(final_T), _p(n_contrib), _p(dL_dpix), _p(dL_dinvdepth),
                              _p(out.dL_dmean2D), _p(out.dL_dconic), _p(out.dL_dopacity), _p(out.dL_dcolor),
                              _p(out.dL_dinvd), ctypes.c_int(tile_stride), ctypes.c_int(tile_offset))
    return out


class ParamGrads(NamedTuple):
    dL_dmeans3D: np.ndarray
    dL_dsh: Optional[np.ndarray]
    dL_dscales: Optional[np.ndarray]
    dL_drot: Optional[np.ndarray]
    dL_dcov3D: np.ndarray
    dL_dopacity: np.ndarray
    dL_dcolor: np.ndarray
    dL_dmean2D: np.ndarray


def preprocess_backward(P, D, M, W, H, means3D, scales, mod, rotations, opacities, shs, cov3D_precomp,
                        view, proj, campos, tanfovx, tanfovy, antialiasing, g: Geom, rg: RenderGrads,
                        use_invdepth_grad=False) -> ParamGrads:
    means3D, scales, rotations = _f32(means3D), _f32(scales), _f32(rotations)
    opacities, shs, cov3D_precomp = _f32(opacities), _f32(shs), _f32(cov3D_precomp)
    view, proj, campos = _f32(view), _f32(proj), _f32(campos)
    dmeans = np.zeros((P, 3), np.float32)
    dsh = np.zeros((P, M, 3), np.float32) if shs is not None else None
    dscales = np.zeros((P, 3), np.float32) if cov3D_precomp is None else None
    drot = np.zeros((P, 4), np.float32) if cov3D_precomp is None else None
    dcov = np.zeros((P, 6), np.float32)
    dop = rg.dL_dopacity.copy()
    lib().orc_preprocess_backward(
        ctypes.c_int(P), ctypes.c_int(D), ctypes.c_int(M), ctypes.c_int(W), ctypes.c_int(H),
        _p(means3D), _p(scales), ctypes.c_float(mod), _p(rotations), _p(opacities), _p(shs),
        _p(cov3D_precomp), _p(view), _p(proj), _p(campos), ctypes.c_float(tanfovx), ctypes.c_float(tanfovy),
        ctypes.c_int(int(antialiasing)), _p(g.radii), _p(g.cov3D), _p(g.clamped),
        _p(rg.dL_dmean2D), _p(rg.dL_dconic), _p(dop), _p(rg.dL_dcolor),
        _p(rg.dL_dinvd if use_invdepth_grad else None),
        _p(dmeans), _p(dsh), _p(dscales), _p(drot), _p(dcov))
    return ParamGrads(dmeans, dsh, dscales, drot, dcov, dop, rg.dL_dcolor, rg.dL_dmean2D)


class ForwardResult(NamedTuple):
    color: np.ndarray
    invdepth: Optional[np.ndarray]
    geom: Geom
    binning: Binning
    final_T: np.ndarray
    n_contrib: np.ndarray


def forward(P, D, M, W, H, bg, means3D, shs, colors_precomp, opacities, scales, mod, rotations,
            cov3D_precomp, view, proj, campos, tanfovx, tanfovy, antialiasing=False, want_invdepth=False) -> ForwardResult:
    g = preprocess(P, D, M, W, H, means3D, scales, mod, rotations, opacities, shs, colors_precomp,
                   cov3D_precomp, view, proj, campos, tanfovx, tanfovy, antialiasing)
    b = bin_tiles(P, W, H, g)
    color, invd, fT, nc = render(W, H, g, b, bg, want_invdepth)
    return ForwardResult(color, invd, g, b, fT, nc)


def backward(fr: ForwardResult, P, D, M, W, H, bg, means3D, shs, opacities, scales, mod, rotations,
             cov3D_precomp, view, proj, campos, tanfovx, tanfovy, dL_dpix, dL_dinvdepth=None,
             antialiasing=False) -> ParamGrads:
    rg = render_backward(P, W, H, fr.geom, fr.binning, bg, fr.final_T, fr.n_contrib, dL_dpix, dL_dinvdepth)
    return preprocess_backward(P, D, M, W, H, means3D, scales, mod, rotations, opacities, shs, cov3D_precomp,
                               view, proj, campos, tanfovx, tanfovy, antialiasing, fr.geom, rg,
                               use_invdepth_grad=dL_dinvdepth is not None)
