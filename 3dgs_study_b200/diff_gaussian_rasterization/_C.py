"""`_C` of the drop-in `diff_gaussian_rasterization` package: a thin ctypes binding over the C ABI of
libb200splat.so (include/b200splat.h).

It exposes the native surface BASELINE.json's north_star names -- `rasterize_gaussians`,
`rasterize_gaussians_backward` (+ `mark_visible`) -- with the positional argument order of the pybind
module it replaces (the submodule pinned at /root/reference/.gitmodules:4-6, absent from the
snapshot; call site /root/reference/gaussian_renderer/__init__.py:98-106).  PyTorch is used for
device memory and the current CUDA stream only; every kernel lives in libb200splat.so.

There is NO fallback: if the library is missing or the tensors are not CUDA tensors this raises.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Tuple

import torch

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# B200SPLAT_LIB selects another build of the same library (kernel-variant experiments); never a fallback
LIB_PATH = os.environ.get("B200SPLAT_LIB") or os.path.join(_PKG_DIR, "libb200splat.so")

_alloc_fn_t = ctypes.CFUNCTYPE(ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)


class _View(ctypes.Structure):
    _fields_ = [
        ("image_height", ctypes.c_int32), ("image_width", ctypes.c_int32),
        ("tanfovx", ctypes.c_float), ("tanfovy", ctypes.c_float), ("scale_modifier", ctypes.c_float),
        ("sh_degree", ctypes.c_int32), ("prefiltered", ctypes.c_int32), ("debug", ctypes.c_int32),
        ("antialiasing", ctypes.c_int32),
        ("bg", ctypes.c_void_p), ("viewmatrix", ctypes.c_void_p), ("projmatrix", ctypes.c_void_p),
        ("campos", ctypes.c_void_p),
    ]


class _Gaussians(ctypes.Structure):
    _fields_ = [
        ("P", ctypes.c_int32), ("M", ctypes.c_int32),
        ("means3D", ctypes.c_void_p), ("shs", ctypes.c_void_p), ("colors_precomp", ctypes.c_void_p),
        ("opacities", ctypes.c_void_p), ("scales", ctypes.c_void_p), ("rotations", ctypes.c_void_p),
        ("cov3D_precomp", ctypes.c_void_p),
    ]


class _ForwardOut(ctypes.Structure):
    _fields_ = [("color", ctypes.c_void_p), ("radii", ctypes.c_void_p), ("invdepth", ctypes.c_void_p)]


class _Allocators(ctypes.Structure):
    _fields_ = [("alloc", _alloc_fn_t), ("geom_ctx", ctypes.c_void_p), ("binning_ctx", ctypes.c_void_p),
                ("image_ctx", ctypes.c_void_p), ("bin_capacity", ctypes.c_int64)]


class StageTimes(ctypes.Structure):
    _fields_ = [(n, ctypes.c_float) for n in (
        "preprocess_ms", "scan_ms", "duplicate_ms", "sort_ms", "ranges_ms", "render_ms",
        "render_bwd_ms", "preprocess_bwd_ms")]

    def as_dict(self):
        return {n: float(getattr(self, n)) for n, _ in self._fields_}


class _Grads(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in (
        "dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
        "dL_drotations")]


_hook_fn_t = ctypes.CFUNCTYPE(None, ctypes.c_void_p)


class _BackwardHooks(ctypes.Structure):
    _fields_ = [("masked_colors", ctypes.c_void_p), ("after_masked_colors", _hook_fn_t), ("ctx", ctypes.c_void_p)]


class _StateOut(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in (
        "means2D", "depths", "conic_opacity", "rgb", "tiles_touched", "point_offsets", "keys_sorted",
        "vals_sorted", "ranges", "final_T", "n_contrib")]


_lib = None


def lib():
    """Load libb200splat.so (once).  Raises if it has not been built: there is no CPU fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `make -C 3dgs_study_b200/csrc`). This package has no CPU or PyTorch fallback.")
        L = ctypes.CDLL(LIB_PATH)
        L.b200splat_last_error.restype = ctypes.c_char_p
        L.b200splat_version.restype = ctypes.c_int
        for f in ("b200splat_geom_bytes", "b200splat_image_bytes", "b200splat_binning_bytes"):
            getattr(L, f).restype = ctypes.c_size_t
        L.b200splat_forward.restype = ctypes.c_int
        L.b200splat_forward.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.b200splat_backward.restype = ctypes.c_int
        L.b200splat_backward.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_void_p, _alloc_fn_t, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_void_p]
        L.b200splat_backward_ex.restype = ctypes.c_int
        L.b200splat_backward_ex.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, _alloc_fn_t, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_void_p]
        L.b200splat_mark_visible.restype = ctypes.c_int
        L.b200splat_mark_visible.argtypes = [ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.c_void_p]
        L.b200splat_sh_grad_from_views.restype = ctypes.c_int
        L.b200splat_sh_grad_from_views.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p,
                                                   ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                                   ctypes.c_void_p]
        L.b200splat_densify_stats.restype = ctypes.c_int
        L.b200splat_densify_stats.argtypes = [ctypes.c_int32] + [ctypes.c_void_p] * 6
        L.b200splat_backward_packed.restype = ctypes.c_int
        L.b200splat_backward_packed.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int32] + [ctypes.c_void_p] * 5 + [
            _alloc_fn_t, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.b200splat_backward_scratch_bytes.restype = ctypes.c_size_t
        L.b200splat_backward_scratch_bytes.argtypes = [ctypes.c_int32]
        L.b200splat_backward_packed_clean.restype = ctypes.c_int
        L.b200splat_backward_packed_clean.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int32] + [ctypes.c_void_p] * 7 + [
            ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
        L.b200splat_densify_stats_visible.restype = ctypes.c_int
        L.b200splat_densify_stats_visible.argtypes = [ctypes.c_int32] + [ctypes.c_void_p] * 4
        L.b200splat_clear_accumulated.restype = ctypes.c_int
        L.b200splat_clear_accumulated.argtypes = [ctypes.c_int32, ctypes.c_int32] + [ctypes.c_void_p] * 4
        L.b200splat_accumulate_packed.restype = ctypes.c_int
        L.b200splat_accumulate_packed.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int32,
                                                  ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                                  ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p]
        L.b200splat_render_stats.restype = ctypes.c_int
        L.b200splat_render_stats.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                             ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                             ctypes.c_void_p]
        L.b200splat_forward_status.restype = ctypes.c_int
        L.b200splat_forward_status.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p,
                                               ctypes.c_void_p, ctypes.c_void_p]
        L.b200splat_image_status_offset.restype = ctypes.c_size_t
        L.b200splat_image_status_offset.argtypes = [ctypes.c_int32, ctypes.c_int32]
        L.b200splat_shutdown.restype = None
        L.b200splat_zero_fill.restype = ctypes.c_int
        L.b200splat_zero_fill.argtypes = [ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.b200splat_export_state.restype = ctypes.c_int
        L.b200splat_export_state.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                             ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.c_void_p]
        _lib = L
    return _lib


def _check(rc: int, what: str):
    if rc != 0:
        msg = lib().b200splat_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"b200splat {what} failed (status {rc}): {msg}")


def _ptr(t: Optional[torch.Tensor]):
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32c(t: Optional[torch.Tensor], name: str, device) -> Optional[torch.Tensor]:
    """Empty / None -> None; otherwise a contiguous fp32 CUDA tensor on `device`."""
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (this rasterizer has no CPU path)")
    if t.device != device:
        raise RuntimeError(f"{name} is on {t.device}, expected {device}")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


class _TensorAllocator:
    """Serves the library's allocator callbacks from torch's caching allocator, keeping the byte
    tensors alive so autograd can save them for the backward pass."""

    def __init__(self, device):
        self.device = device
        self.buffers = {}
        self.error = None

        def _cb(ctx, nbytes):
            try:
                # The binning buffer's size follows the pair count R, which differs from view to view; a new
                # size every call fragments the caching allocator (it kept cudaMalloc-ing: +10 GB reserved over
                # 30 views of the 3M scene).  Sizes are therefore quantised to 1/8 of their power of two, so that
                # consecutive views reuse each other's blocks.
                n = max(int(nbytes), 1)
                if n > (1 << 20):
                    q = 1 << (n.bit_length() - 4)
                    n = (n + q - 1) // q * q
                t = torch.empty(n, dtype=torch.uint8, device=self.device)
                self.buffers[int(ctx) if ctx else 0] = t
                return t.data_ptr()
            except Exception as e:  # surfaced after the C call returns B200SPLAT_ERR_ALLOC
                self.error = e
                return None

        self.cb = _alloc_fn_t(_cb)

    def release(self):
        """Break the self -> callback -> closure -> self cycle: otherwise the scratch tensors (hundreds of MB
        per call) stay alive until Python's cyclic GC happens to run, and the caching allocator keeps growing."""
        bufs, self.buffers, self.cb = self.buffers, {}, None
        return bufs


def _stream_ptr(device) -> int:
    return int(torch.cuda.current_stream(device).cuda_stream)


# ---- pair capacity policy (include/b200splat.h "Pair capacity") ------------------------------------------------
# The first view of a (device, P, W, H) configuration is rendered in exact mode (one host round trip for the pair
# count R).  Every later view gets a binning buffer sized from the largest R seen so far with head room and is
# enqueued WITHOUT waiting for the device; its status words {R, capacity, overflow} are copied to pinned host
# memory behind it and looked at when the next view is about to be enqueued -- by then they have long landed.
# An overflow (R grew by more than the head room from one view to the next) therefore surfaces one call late, as a
# RuntimeError, the way an asynchronous CUDA error does; B200SPLAT_EXACT=1 (or exact=True) turns the policy off.
_CAP_HEADROOM, _CAP_SLACK, _CAP_QUANTUM = 1.5, 1 << 18, 1 << 16
_cap_state = {}


def _grow(R: int) -> int:
    c = int(R * _CAP_HEADROOM) + _CAP_SLACK
    return min((c + _CAP_QUANTUM - 1) // _CAP_QUANTUM * _CAP_QUANTUM, 0x7fffffff)


def reset_capacity_policy():
    """Forget the pair-capacity history (the next view of every configuration is rendered in exact mode again)."""
    _cap_state.clear()


def capacity_status(device, P, image_height, image_width):
    """-> (R, capacity, overflow) as last landed in pinned memory for this configuration (no synchronisation), or None:
    what the capacity policy itself looks at; for callers that replay captured forwards and must check for themselves."""
    dev = torch.device(device)
    key = (dev.index if dev.index is not None else torch.cuda.current_device(), int(P), int(image_width), int(image_height))
    st = _cap_state.get(key)
    if st is None or not st["pending"]:
        return None
    lo, hi, cap, over = (int(x) & 0xffffffff for x in st["pinned"].tolist())
    return lo | (hi << 32), cap, bool(over)


def forward_status(imageBuffer, image_height, image_width):
    """-> (R, capacity, overflow) of the view whose image buffer is given (b200splat_forward_status; synchronises)."""
    dev = imageBuffer.device
    R, cap, over = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int32(0)
    with torch.cuda.device(dev):
        rc = lib().b200splat_forward_status(_ptr(imageBuffer), int(image_width), int(image_height), ctypes.byref(R),
                                            ctypes.byref(cap), ctypes.byref(over), _stream_ptr(dev))
    _check(rc, "forward_status")
    return int(R.value), int(cap.value), bool(over.value)


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, debug, antialiasing=False, want_invdepth=False, stage_times: StageTimes = None,
                        exact: Optional[bool] = None, bin_capacity: Optional[int] = None, status_stream=None):
    """-> (num_rendered, out_color [3,H,W], radii [P] i32, geomBuffer, binningBuffer, imgBuffer, invdepth|None)

    num_rendered is the pair capacity of the binning buffer (what the backward needs): the exact pair count R when
    the view was rendered in exact mode, an upper bound otherwise; forward_status(imgBuffer, H, W) gives R itself.
    exact=True: wait for R (one host round trip).  bin_capacity=n: allocate for n pairs, never wait, the caller
    checks forward_status for overflow.  Neither (default): the capacity policy above.
    status_stream (capacity policy only): a torch stream on which the 16 status bytes travel to pinned memory instead of
    the current one, so that the kernels queued after the forward do not wait for a copy-engine round trip (0.01 ms);
    the caller joins that stream before imgBuffer is released (and before a stream capture ends)."""
    if not means3D.is_cuda:
        raise RuntimeError("means3D must be a CUDA tensor (this rasterizer has no CPU path)")
    dev = means3D.device
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    P = means3D.shape[0]
    H, W = int(image_height), int(image_width)
    means3D = _f32c(means3D, "means3D", dev)
    colors = _f32c(colors, "colors_precomp", dev)
    opacity = _f32c(opacity, "opacities", dev)
    scales = _f32c(scales, "scales", dev)
    rotations = _f32c(rotations, "rotations", dev)
    cov3D_precomp = _f32c(cov3D_precomp, "cov3D_precomp", dev)
    sh = _f32c(sh, "shs", dev)
    background = _f32c(background, "bg", dev)
    viewmatrix = _f32c(viewmatrix, "viewmatrix", dev)
    projmatrix = _f32c(projmatrix, "projmatrix", dev)
    campos = _f32c(campos, "campos", dev)
    M = int(sh.shape[1]) if sh is not None else 0

    # ---- capacity of the binning buffer
    st = None
    if bin_capacity is not None:
        cap = int(bin_capacity)
        if cap <= 0:
            raise RuntimeError("bin_capacity must be positive")
    elif exact or (exact is None and os.environ.get("B200SPLAT_EXACT") == "1") or debug or stage_times is not None or P == 0:
        cap = 0
    else:
        key = (dev.index if dev.index is not None else torch.cuda.current_device(), P, W, H)
        st = _cap_state.get(key)
        if st is None:
            cap = 0
            st = _cap_state[key] = {"hint": 0, "pinned": torch.zeros(4, dtype=torch.int32).pin_memory(), "pending": False}
        else:
            if st["pending"]:
                lo, hi, cap_prev, over = (int(x) & 0xffffffff for x in st["pinned"].tolist())
                R_prev = lo | (hi << 32)
                st["hint"] = max(st["hint"], _grow(R_prev))
                if over:
                    st["pending"] = False
                    raise RuntimeError(
                        f"b200splat: the PREVIOUS forward call on this device produced {R_prev} (tile, Gaussian) pairs, more than the "
                        f"{cap_prev} its binning buffer was sized for; its image and everything derived from it are invalid. "
                        "The capacity has been raised; re-render that view, or set B200SPLAT_EXACT=1.")
            cap = st["hint"]

    out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    invdepth = torch.empty((1, H, W), dtype=torch.float32, device=dev) if want_invdepth else None

    view = _View(H, W, float(tan_fovx), float(tan_fovy), float(scale_modifier), int(degree), int(bool(prefiltered)),
                 int(bool(debug)), int(bool(antialiasing)), _ptr(background), _ptr(viewmatrix), _ptr(projmatrix),
                 _ptr(campos))
    g = _Gaussians(P, M, _ptr(means3D), _ptr(sh), _ptr(colors), _ptr(opacity), _ptr(scales), _ptr(rotations),
                   _ptr(cov3D_precomp))
    out = _ForwardOut(_ptr(out_color), _ptr(radii), _ptr(invdepth))
    ta = _TensorAllocator(dev)
    allocs = _Allocators(ta.cb, 1, 2, 3, cap)
    num_rendered = ctypes.c_int32(0)
    with torch.cuda.device(dev):
        rc = lib().b200splat_forward(ctypes.addressof(view), ctypes.addressof(g), ctypes.addressof(out),
                                     ctypes.addressof(allocs), ctypes.addressof(num_rendered),
                                     ctypes.addressof(stage_times) if stage_times is not None else None,
                                     _stream_ptr(dev))
    bufs = ta.release()
    if ta.error is not None:
        raise ta.error
    _check(rc, "forward")
    empty = torch.empty(0, dtype=torch.uint8, device=dev)
    img = bufs.get(3, empty)
    if st is not None:
        if cap == 0:                   # the exact first view: R is known on the host
            st["hint"] = _grow(int(num_rendered.value))
        else:                          # status words follow the view to pinned memory; read before the next one
            off = lib().b200splat_image_status_offset(W, H)
            if status_stream is not None:
                done = torch.cuda.Event(); done.record(torch.cuda.current_stream(dev))
                status_stream.wait_event(done)
                with torch.cuda.stream(status_stream):
                    st["pinned"].copy_(img[off:off + 16].view(torch.int32), non_blocking=True)
            else:
                st["pinned"].copy_(img[off:off + 16].view(torch.int32), non_blocking=True)
            st["pending"] = True
    return (int(num_rendered.value), out_color, radii, bufs.get(1, empty), bufs.get(2, empty), img, invdepth)


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                 viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, debug, opacities=None, antialiasing=False,
                                 image_height=None, image_width=None, stage_times: StageTimes = None,
                                 masked_colors: bool = False, out: Optional[dict] = None,
                                 early_colors: Optional[torch.Tensor] = None, after_early_colors=None,
                                 dL_dout_invdepth: Optional[torch.Tensor] = None):
    """-> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)

    dL_dout_invdepth ([1,H,W], optional): upstream gradient of the inverse-depth image (newer upstream API).

    masked_colors=True (SH mode only): dL_dsh is not produced (None); dL_dcolors is the colour gradient times the
    SH clamp mask, the quantity the view-parallel exchange ships (see sh_grad_from_views).
    out: optional preallocated fp32 CUDA outputs by name (e.g. views into one flat all-reduce buffer).
    early_colors ([P,3] fp32 CUDA, SH mode): receives the masked colour gradient right after the render backward;
    `after_early_colors()` is then called on the host, before the per-Gaussian backward is enqueued, so the caller
    can start exchanging the buffer on another stream (b200splat_backward_ex).  Implies no dL_dsh / dL_dcolors."""
    dev = means3D.device
    P = means3D.shape[0]
    dL_dout_color = _f32c(dL_dout_color, "dL_dout_color", dev)
    dL_dout_invdepth = _f32c(dL_dout_invdepth, "dL_dout_invdepth", dev)
    H = int(image_height) if image_height is not None else int(dL_dout_color.shape[1])
    W = int(image_width) if image_width is not None else int(dL_dout_color.shape[2])
    means3D = _f32c(means3D, "means3D", dev)
    colors = _f32c(colors, "colors_precomp", dev)
    scales = _f32c(scales, "scales", dev)
    rotations = _f32c(rotations, "rotations", dev)
    cov3D_precomp = _f32c(cov3D_precomp, "cov3D_precomp", dev)
    sh = _f32c(sh, "shs", dev)
    opacities = _f32c(opacities, "opacities", dev)
    if opacities is None:
        raise RuntimeError("opacities are required by the backward pass")
    background = _f32c(background, "bg", dev)
    viewmatrix = _f32c(viewmatrix, "viewmatrix", dev)
    projmatrix = _f32c(projmatrix, "projmatrix", dev)
    campos = _f32c(campos, "campos", dev)
    M = int(sh.shape[1]) if sh is not None else 0
    if dL_dout_invdepth is not None and dL_dout_invdepth.numel() != H * W:
        raise RuntimeError(f"dL_dout_invdepth must have {H}*{W} elements")

    def new(name, *shape):
        if out is not None and name in out:
            t = out[name]
            if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == int(torch.Size(shape).numel())):
                raise RuntimeError(f"out[{name!r}] must be a contiguous fp32 CUDA tensor with {shape} elements")
            return t.view(shape)
        return torch.empty(shape, dtype=torch.float32, device=dev)

    if (masked_colors or early_colors is not None) and sh is None:
        raise RuntimeError("masked_colors / early_colors need shs")
    if early_colors is not None:
        if not (early_colors.is_cuda and early_colors.dtype == torch.float32 and early_colors.is_contiguous()
                and early_colors.numel() == 3 * P):
            raise RuntimeError("early_colors must be a contiguous fp32 CUDA tensor with 3 P elements")
        masked_colors = True
    dL_dmeans2D = new("dL_dmeans2D", P, 3)
    dL_dcolors = new("dL_dcolors", P, 3) if (colors is not None or (masked_colors and early_colors is None)) else None
    dL_dopacity = new("dL_dopacity", P, 1)
    dL_dmeans3D = new("dL_dmeans3D", P, 3)
    dL_dcov3D = new("dL_dcov3D", P, 6) if cov3D_precomp is not None else None
    dL_dsh = new("dL_dsh", P, M, 3) if (sh is not None and not masked_colors) else None
    dL_dscales = new("dL_dscales", P, 3) if scales is not None else None
    dL_drotations = new("dL_drotations", P, 4) if rotations is not None else None

    view = _View(H, W, float(tan_fovx), float(tan_fovy), float(scale_modifier), int(degree), 0, int(bool(debug)),
                 int(bool(antialiasing)), _ptr(background), _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos))
    g = _Gaussians(P, M, _ptr(means3D), _ptr(sh), _ptr(colors), _ptr(opacities), _ptr(scales), _ptr(rotations),
                   _ptr(cov3D_precomp))
    grads = _Grads(_ptr(dL_dmeans2D), _ptr(dL_dcolors), _ptr(dL_dopacity), _ptr(dL_dmeans3D), _ptr(dL_dcov3D),
                   _ptr(dL_dsh), _ptr(dL_dscales), _ptr(dL_drotations))
    ta = _TensorAllocator(dev)
    hooks, hook_err = None, []
    if early_colors is not None:
        def _after(_ctx):
            try:
                if after_early_colors is not None:
                    after_early_colors()
            except Exception as e:   # re-raised after the C call returns
                hook_err.append(e)
        hook_cb = _hook_fn_t(_after)
        hooks = _BackwardHooks(_ptr(early_colors), hook_cb, None)
    with torch.cuda.device(dev):
        rc = lib().b200splat_backward_ex(ctypes.addressof(view), ctypes.addressof(g), _ptr(radii), _ptr(geomBuffer),
                                         _ptr(binningBuffer), _ptr(imageBuffer), int(R), _ptr(dL_dout_color),
                                         _ptr(dL_dout_invdepth),
                                         ctypes.addressof(grads), ctypes.addressof(hooks) if hooks is not None else None,
                                         ta.cb, 1, ctypes.addressof(stage_times) if stage_times is not None else None,
                                         _stream_ptr(dev))
    if hook_err:
        ta.release()
        raise hook_err[0]
    ta.release()
    if ta.error is not None:
        raise ta.error
    _check(rc, "backward")
    # the scratch tensor may be released now: stream-ordered reuse by the caching allocator is safe
    return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)


def mark_visible(means3D, viewmatrix, projmatrix):
    if not means3D.is_cuda:
        raise RuntimeError("means3D must be a CUDA tensor (this rasterizer has no CPU path)")
    dev = means3D.device
    P = means3D.shape[0]
    means3D = _f32c(means3D, "means3D", dev)
    viewmatrix = _f32c(viewmatrix, "viewmatrix", dev)
    projmatrix = _f32c(projmatrix, "projmatrix", dev)
    present = torch.zeros((P,), dtype=torch.bool, device=dev)
    with torch.cuda.device(dev):
        rc = lib().b200splat_mark_visible(P, _ptr(means3D), _ptr(viewmatrix), _ptr(projmatrix), _ptr(present),
                                          _stream_ptr(dev))
    _check(rc, "mark_visible")
    return present


def sh_grad_from_views(means3D, dcols, campos, M: int, degree: int, out: Optional[torch.Tensor] = None):
    """dL_dsh[P,M,3] = sum_v Y(normalize(means3D - campos[v])) (x) dcols[v]  (b200splat_sh_grad_from_views).
    dcols: sequence of [P,3] fp32 CUDA tensors, or raw device addresses (ints) of such arrays -- e.g. peer-GPU
    buffers mapped through symmetric memory.  campos: [n_views,3] fp32 CUDA tensor."""
    dev = means3D.device
    P = means3D.shape[0]
    means3D = _f32c(means3D, "means3D", dev)
    campos = _f32c(campos, "campos", dev)
    n = len(dcols)
    keep = []
    ptrs = (ctypes.c_void_p * n)()
    for v, d in enumerate(dcols):
        if isinstance(d, torch.Tensor):
            d = _f32c(d, "dcol", dev)
            if d is None or d.numel() != 3 * P:
                raise RuntimeError("every dcol must be a [P,3] tensor")
            keep.append(d)
            ptrs[v] = d.data_ptr()
        else:
            ptrs[v] = int(d)
    if campos is None or campos.numel() != 3 * n:
        raise RuntimeError("campos must be [n_views,3]")
    dL_dsh = out if out is not None else torch.empty((P, int(M), 3), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = lib().b200splat_sh_grad_from_views(P, int(M), int(degree), _ptr(means3D), n, ptrs, _ptr(campos), _ptr(dL_dsh),
                                                _stream_ptr(dev))
    _check(rc, "sh_grad_from_views")
    return dL_dsh


PACKED_ROW_WORDS = 16     # f32 words per packed gradient row (include/b200splat.h "packed gradient rows")


def rasterize_gaussians_backward_packed(background, means3D, radii, scales, rotations, scale_modifier, viewmatrix, projmatrix,
                                        tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R, binningBuffer,
                                        imageBuffer, opacities, packed_rows, packed_count, image_height, image_width,
                                        antialiasing=False, dL_dmeans2D=None, dL_dout_invdepth=None,
                                        stage_times: StageTimes = None, clean_scratch: Optional[torch.Tensor] = None,
                                        xyz_gradient_accum: Optional[torch.Tensor] = None):
    """The backward pass of one view, producing packed gradient rows (one 64-byte row per Gaussian the view touched)
    for the sparse view-parallel exchange instead of dense gradients (b200splat_backward_packed).
    packed_rows: fp32 CUDA tensor with room for P rows of 16 words, or the raw device address (int) of such a
    buffer; packed_count: int32/uint32 CUDA tensor of one element, or its address.  dL_dmeans2D ([P,3], optional)
    is still written densely (densification statistics are rank-local).
    clean_scratch: a tensor from `clean_backward_scratch(P, device)`, kept by the caller from call to call
    (b200splat_backward_packed_clean): no allocation and no 150 MB clear per call.  After an exception it must be
    zero-filled (or replaced) before it is used again.
    xyz_gradient_accum ([P] or [P,1] fp32, with clean_scratch only): the backward adds ||dL/dmean2D|| of the Gaussians it
    visits into it (the gradient-norm line of `densify_stats`; `densify_stats_visible` does the other two)."""
    dev = means3D.device
    P = means3D.shape[0]
    H, W = int(image_height), int(image_width)
    dL_dout_color = _f32c(dL_dout_color, "dL_dout_color", dev)
    dL_dout_invdepth = _f32c(dL_dout_invdepth, "dL_dout_invdepth", dev)
    means3D = _f32c(means3D, "means3D", dev); scales = _f32c(scales, "scales", dev)
    rotations = _f32c(rotations, "rotations", dev); sh = _f32c(sh, "shs", dev)
    opacities = _f32c(opacities, "opacities", dev); background = _f32c(background, "bg", dev)
    viewmatrix = _f32c(viewmatrix, "viewmatrix", dev); projmatrix = _f32c(projmatrix, "projmatrix", dev)
    campos = _f32c(campos, "campos", dev)
    if sh is None or scales is None or rotations is None or opacities is None:
        raise RuntimeError("packed rows need shs, scales, rotations and opacities")
    M = int(sh.shape[1])
    if isinstance(packed_rows, torch.Tensor):
        if not (packed_rows.is_cuda and packed_rows.dtype == torch.float32 and packed_rows.is_contiguous()
                and packed_rows.numel() >= PACKED_ROW_WORDS * P):
            raise RuntimeError("packed_rows must be a contiguous fp32 CUDA tensor with room for P rows of 16 words")
    if dL_dmeans2D is not None and not (dL_dmeans2D.is_cuda and dL_dmeans2D.dtype == torch.float32
                                        and dL_dmeans2D.is_contiguous() and dL_dmeans2D.numel() == 3 * P):
        raise RuntimeError("dL_dmeans2D must be a contiguous fp32 CUDA tensor [P,3]")
    rows_ptr = packed_rows.data_ptr() if isinstance(packed_rows, torch.Tensor) else int(packed_rows)
    cnt_ptr = packed_count.data_ptr() if isinstance(packed_count, torch.Tensor) else int(packed_count)
    view = _View(H, W, float(tan_fovx), float(tan_fovy), float(scale_modifier), int(degree), 0, 0,
                 int(bool(antialiasing)), _ptr(background), _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos))
    g = _Gaussians(P, M, _ptr(means3D), _ptr(sh), None, _ptr(opacities), _ptr(scales), _ptr(rotations), None)
    if xyz_gradient_accum is not None and (clean_scratch is None or not (
            xyz_gradient_accum.is_cuda and xyz_gradient_accum.dtype == torch.float32 and xyz_gradient_accum.is_contiguous()
            and xyz_gradient_accum.numel() == P)):
        raise RuntimeError("xyz_gradient_accum must be a contiguous fp32 CUDA tensor of P elements, and needs clean_scratch")
    if clean_scratch is not None:
        if not (clean_scratch.is_cuda and clean_scratch.device == dev and clean_scratch.is_contiguous()):
            raise RuntimeError("clean_scratch must be a contiguous CUDA tensor on the Gaussians' device")
        with torch.cuda.device(dev):
            rc = lib().b200splat_backward_packed_clean(
                ctypes.addressof(view), ctypes.addressof(g), _ptr(radii), _ptr(geomBuffer), _ptr(binningBuffer), _ptr(imageBuffer),
                int(R), _ptr(dL_dout_color), _ptr(dL_dout_invdepth), _ptr(dL_dmeans2D), _ptr(xyz_gradient_accum), rows_ptr, cnt_ptr,
                clean_scratch.data_ptr(), clean_scratch.numel() * clean_scratch.element_size(),
                ctypes.addressof(stage_times) if stage_times is not None else None, _stream_ptr(dev))
        _check(rc, "backward_packed_clean")
        return
    ta = _TensorAllocator(dev)
    with torch.cuda.device(dev):
        rc = lib().b200splat_backward_packed(ctypes.addressof(view), ctypes.addressof(g), _ptr(radii), _ptr(geomBuffer),
                                             _ptr(binningBuffer), _ptr(imageBuffer), int(R), _ptr(dL_dout_color),
                                             _ptr(dL_dout_invdepth), _ptr(dL_dmeans2D), rows_ptr, cnt_ptr, ta.cb, 1,
                                             ctypes.addressof(stage_times) if stage_times is not None else None,
                                             _stream_ptr(dev))
    ta.release()
    if ta.error is not None:
        raise ta.error
    _check(rc, "backward_packed")


def clear_accumulated(dL_dsh: torch.Tensor, dL_dgeom: torch.Tensor, marks: torch.Tensor):
    """Zeroes the rows of dL_dsh [P,M,3] / dL_dgeom [P,12] that `accumulate_packed(..., marks=marks)` wrote, and the marks
    ([P] uint8) by which it finds them (b200splat_clear_accumulated)."""
    P, M = dL_dsh.shape[0], dL_dsh.shape[1]
    for t in (dL_dsh, dL_dgeom):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise RuntimeError("clear_accumulated needs contiguous fp32 CUDA tensors")
    if dL_dgeom.numel() != 12 * P:
        raise RuntimeError("dL_dgeom must be [P,12]")
    if not (marks.is_cuda and marks.dtype == torch.uint8 and marks.is_contiguous() and marks.numel() == P):
        raise RuntimeError("marks must be a contiguous uint8 CUDA tensor [P]")
    with torch.cuda.device(dL_dsh.device):
        _check(lib().b200splat_clear_accumulated(P, M, dL_dsh.data_ptr(), dL_dgeom.data_ptr(), marks.data_ptr(),
                                                 _stream_ptr(dL_dsh.device)), "clear_accumulated")


def clean_backward_scratch(P: int, device) -> torch.Tensor:
    """A zero-filled scratch for `rasterize_gaussians_backward_packed(..., clean_scratch=...)`."""
    return torch.zeros(int(lib().b200splat_backward_scratch_bytes(int(P))), dtype=torch.uint8, device=device)


def accumulate_packed(means3D, rows, counts, campos, M: int, degree: int, dL_dsh, dL_dgeom, store: bool = False,
                      marks: Optional[torch.Tensor] = None):
    """Adds the packed gradient rows of len(rows) views into dense gradients the caller has cleared: dL_dsh [P,M,3],
    rebuilt from each row's colour gradient and its view's camera centre, and dL_dgeom [P,12] = per Gaussian
    (d rotation[4], d mean[3], d scale[3], d opacity, 0) interleaved (b200splat_accumulate_packed).
    rows[v] / counts[v]: tensors or raw device addresses (ints) -- e.g. a peer GPU's buffers mapped through symmetric
    memory: the kernel then reads them over NVLink.  campos: [len(rows),3] fp32 CUDA tensor.
    store=True (one source): written instead of added, for the first source into freshly cleared gradients.
    marks ([P] uint8, optional): set to 1 for every Gaussian a row goes into -- what `clear_accumulated` goes by."""
    dev = means3D.device
    P = means3D.shape[0]
    means3D = _f32c(means3D, "means3D", dev)
    campos = _f32c(campos, "campos", dev)      # None: the camera centres are in the headers counts[v] points at (words 4..6)
    n = len(rows)
    if len(counts) != n or (campos is not None and campos.numel() != 3 * n):
        raise RuntimeError("rows, counts and campos must describe the same number of views")
    rp, cp = (ctypes.c_void_p * n)(), (ctypes.c_void_p * n)()
    for v in range(n):
        rp[v] = rows[v].data_ptr() if isinstance(rows[v], torch.Tensor) else int(rows[v])
        cp[v] = counts[v].data_ptr() if isinstance(counts[v], torch.Tensor) else int(counts[v])
    for name, t, numel in (("dL_dsh", dL_dsh, 3 * int(M) * P), ("dL_dgeom", dL_dgeom, 12 * P)):
        if t is None or not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == numel):
            raise RuntimeError(f"{name} must be a contiguous fp32 CUDA tensor with {numel} elements")
    if marks is not None and not (marks.is_cuda and marks.dtype == torch.uint8 and marks.is_contiguous() and marks.numel() == P):
        raise RuntimeError("marks must be a contiguous uint8 CUDA tensor [P]")
    with torch.cuda.device(dev):
        rc = lib().b200splat_accumulate_packed(P, int(M), int(degree), _ptr(means3D), n, rp, cp, _ptr(campos),
                                               _ptr(dL_dsh), _ptr(dL_dgeom), _ptr(marks), int(bool(store)), _stream_ptr(dev))
    _check(rc, "accumulate_packed")


def zero_fill(tensors):
    """Zero up to 8 contiguous CUDA tensors with the library's background-clear kernel (one block per SM: made to run
    beside a compute-bound kernel on another stream) on the current stream (b200splat_zero_fill)."""
    tensors = [t for t in tensors if t is not None and t.numel()]
    if not tensors:
        return
    dev = tensors[0].device
    n = len(tensors)
    if n > 8 or any((not t.is_cuda) or (not t.is_contiguous()) or t.device != dev for t in tensors):
        raise RuntimeError("zero_fill takes up to 8 contiguous CUDA tensors of one device")
    ptrs, sizes = (ctypes.c_void_p * n)(), (ctypes.c_size_t * n)()
    for i, t in enumerate(tensors):
        ptrs[i], sizes[i] = t.data_ptr(), t.numel() * t.element_size()
    with torch.cuda.device(dev):
        rc = lib().b200splat_zero_fill(n, ptrs, sizes, _stream_ptr(dev))
    _check(rc, "zero_fill")


def densify_stats(dL_dmeans2D, radii, xyz_gradient_accum, denom, max_radii2D):
    """In place, one kernel: accum += ||dL_dmeans2D[:, :2]||, denom += 1 (visible only), max_radii2D = max(., radii)."""
    dev = radii.device
    P = radii.shape[0]
    for t, n in ((dL_dmeans2D, 3 * P), (xyz_gradient_accum, P), (denom, P), (max_radii2D, P)):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == n):
            raise RuntimeError("densify_stats: contiguous fp32 CUDA tensors of matching size required")
    if radii.dtype != torch.int32:
        raise RuntimeError("radii must be int32")
    with torch.cuda.device(dev):
        rc = lib().b200splat_densify_stats(P, _ptr(dL_dmeans2D), _ptr(radii), _ptr(xyz_gradient_accum), _ptr(denom),
                                           _ptr(max_radii2D), _stream_ptr(dev))
    _check(rc, "densify_stats")


def densify_stats_visible(radii, denom, max_radii2D):
    """In place: denom += 1 and max_radii2D = max(., radii) for radii > 0 -- the lines of `densify_stats` that need the
    forward's radii only (b200splat_densify_stats_visible)."""
    dev = radii.device
    P = radii.shape[0]
    for t in (denom, max_radii2D):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == P):
            raise RuntimeError("densify_stats_visible: contiguous fp32 CUDA tensors of P elements required")
    if radii.dtype != torch.int32:
        raise RuntimeError("radii must be int32")
    with torch.cuda.device(dev):
        _check(lib().b200splat_densify_stats_visible(P, _ptr(radii), _ptr(denom), _ptr(max_radii2D), _stream_ptr(dev)),
               "densify_stats_visible")


def render_stats(P, W, H, num_rendered, geomBuffer, binningBuffer, imageBuffer):
    """Inspection: workload counters of the render stage (see b200splat_render_stats in the header)."""
    dev = geomBuffer.device
    out = torch.zeros(16, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        rc = lib().b200splat_render_stats(int(P), int(W), int(H), int(num_rendered), _ptr(geomBuffer), _ptr(binningBuffer),
                                          _ptr(imageBuffer), _ptr(out), _stream_ptr(dev))
    _check(rc, "render_stats")
    v = [int(x) for x in out.cpu().tolist()]
    return {"pairs_by_subblocks": v[0:9], "pairs_visited": v[9], "pixel_evals": v[10], "exp_evals": v[11],
            "contributions": v[12], "pairs_contributing": v[13], "alpha_test_disagreements": v[14]}


def export_state(P, W, H, num_rendered, radii, geomBuffer, binningBuffer, imageBuffer):
    """Parity/inspection helper: unpack the opaque forward state into named tensors.  num_rendered: as returned by
    rasterize_gaussians (the pair capacity); the sorted lists hold the view's true pair count R entries ("R")."""
    dev = radii.device
    T = ((W + 15) // 16) * ((H + 15) // 16)
    R = forward_status(imageBuffer, H, W)[0] if imageBuffer.numel() else 0
    o = {
        "means2D": torch.zeros((P, 2), dtype=torch.float32, device=dev),
        "depths": torch.zeros((P,), dtype=torch.float32, device=dev),
        "conic_opacity": torch.zeros((P, 4), dtype=torch.float32, device=dev),
        "rgb": torch.zeros((P, 3), dtype=torch.float32, device=dev),
        "tiles_touched": torch.zeros((P,), dtype=torch.int32, device=dev),
        "point_offsets": torch.zeros((P,), dtype=torch.int32, device=dev),
        "keys_sorted": torch.zeros((R,), dtype=torch.int64, device=dev),
        "vals_sorted": torch.zeros((R,), dtype=torch.int32, device=dev),
        "ranges": torch.zeros((T, 2), dtype=torch.int32, device=dev),
        "final_T": torch.zeros((H, W), dtype=torch.float32, device=dev),
        "n_contrib": torch.zeros((H, W), dtype=torch.int32, device=dev),
    }
    so = _StateOut(*[_ptr(o[n]) for n, _ in _StateOut._fields_])
    with torch.cuda.device(dev):
        rc = lib().b200splat_export_state(P, W, H, int(num_rendered), _ptr(radii), _ptr(geomBuffer), _ptr(binningBuffer),
                                          _ptr(imageBuffer), ctypes.addressof(so), _stream_ptr(dev))
    _check(rc, "export_state")
    o["R"] = R
    return o
