"""b200splat_dist2_knn3_static (include/b200splat_next.h): the nearest-neighbour search with its scratch inside the
library's device image, for callers that cannot allocate.  No counterpart in the reference; checked against the
existing entry point b200splat_dist2_knn3, which tests/test_gpu_widen.py pins against the oracle and against the
reference's own build.  Same kernels on the same inputs: the results must be bit-identical (the tolerance the knn
tests use), on two calls in a row that share the scratch array, in a fresh process or after every other test.

Collected last on purpose (the file name sorts after every other test file)."""
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu

importlib.import_module("3dgs_study_b200").ensure_on_path()


def _same(a, b):
    return bool(((a == b) | (torch.isinf(a) & torch.isinf(b))).all())


def test_static_scratch_knn_matches_allocating_entry_point_twice_in_a_row():
    from gaussian_training import _lib
    from simple_knn import _C as knn
    dev = torch.device("cuda:0")
    L = _lib.lib()
    g = torch.Generator().manual_seed(21)
    # second call: fewer points than the first (stale scratch beyond its own extent), then the maximum, then tiny
    sizes = (50_000, 20_011, knn.KNN_STATIC_MAX_POINTS, 5, 1)
    pts = []
    for P in sizes:
        p = torch.randn(P, 3, generator=g) * torch.tensor([1.2, 0.45, 1.2])
        if P > 100:
            p[7] = p[8]
            p[P // 2:] *= 40.0
        pts.append(p.to(dev).contiguous())
    want = [knn.distCUDA2(p) for p in pts]                       # existing entry point (stream-ordered allocation)
    outs = [torch.full((p.shape[0],), -1.0, device=dev) for p in pts]
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
    stream = _lib.stream_ptr(dev)
    for p, o in zip(pts, outs):                                  # back to back on one stream, no synchronisation between
        rc = L.b200splat_dist2_knn3_static(p.shape[0], p.data_ptr(), o.data_ptr(), stream)
        _lib.check(rc, "dist2_knn3_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats(dev)["allocation.all.allocated"] == before      # torch's allocator was not involved
    for P, w, o in zip(sizes, want, outs):
        assert _same(o, w), (P, int((o != w).sum()))
    # the same inputs again through the Python wrapper: the array keeps the last call's contents, results do not change
    for p, w in zip(pts[:2], want[:2]):
        o = torch.empty(p.shape[0], device=dev)
        assert _same(knn.distCUDA2_static(p, o), w)


def test_static_scratch_knn_rejects_sizes_beyond_the_maximum():
    from gaussian_training import _lib
    from simple_knn import _C as knn
    dev = torch.device("cuda:0")
    P = knn.KNN_STATIC_MAX_POINTS + 1
    p = torch.zeros(P, 3, device=dev)
    o = torch.zeros(P, device=dev)
    with pytest.raises(RuntimeError, match="exceed the built-in scratch"):
        knn.distCUDA2_static(p, o)
    L = _lib.lib()
    assert L.b200splat_dist2_knn3_static(-1, p.data_ptr(), o.data_ptr(), None) != 0
    assert L.b200splat_dist2_knn3_static(0, None, None, None) == 0
    assert L.b200splat_dist2_knn3_static(10, None, o.data_ptr(), None) != 0 and b"null" in L.b200splat_last_error()
