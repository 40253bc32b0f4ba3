"""CPU: host logic of the steps either side of the rasterizer -- the native PLY codec (host code in
libb200splat.so) against the numpy restatement of save_ply/load_ply, the ABI of include/b200splat_next.h, and
the loud failure of the GPU entry points without CUDA tensors (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from oracle import next_oracle as no

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gaussians(P, K, seed=0):
    g = torch.Generator().manual_seed(seed)
    return {"xyz": torch.randn(P, 3, generator=g), "features_dc": torch.randn(P, 1, 3, generator=g),
            "features_rest": torch.randn(P, K, 3, generator=g), "opacity": torch.randn(P, 1, generator=g),
            "scaling": torch.randn(P, 3, generator=g), "rotation": torch.randn(P, 4, generator=g)}


@pytest.mark.parametrize("P,K", [(0, 15), (1, 15), (7, 0), (1000, 3), (70001, 15)])
def test_ply_round_trip_and_bytes(tmp_path, P, K):
    import gaussian_training as gtr
    d = _gaussians(P, K)
    path = str(tmp_path / "point_cloud" / "iteration_7" / "point_cloud.ply")        # save_ply makes the directories
    gtr.save_ply(path, *d.values())
    assert open(path, "rb").read() == no.ply_bytes(*(v.numpy() for v in d.values()))   # byte-identical to the restatement
    r = gtr.load_ply(path)
    for k in d:
        assert r[k].shape == d[k].shape and torch.equal(r[k], d[k]), k
    if K:
        assert gtr.ply_io.sh_degree_of(r["features_rest"]) == int(round((K + 1) ** 0.5)) - 1


def test_ply_reader_finds_properties_by_name(tmp_path):
    """load_ply reads by property name (scene/gaussian_model.py:269-306): shuffled columns, an extra uchar
    property and a trailing empty element must not matter."""
    import gaussian_training as gtr
    P, K = 33, 15
    d = _gaussians(P, K, seed=4)
    names = no.ply_attribute_names(3 * K)
    cols = np.concatenate([d["xyz"].numpy(), np.zeros((P, 3), np.float32),
                           d["features_dc"].numpy().transpose(0, 2, 1).reshape(P, -1),
                           d["features_rest"].numpy().transpose(0, 2, 1).reshape(P, -1), d["opacity"].numpy(),
                           d["scaling"].numpy(), d["rotation"].numpy()], axis=1)
    perm = np.random.default_rng(0).permutation(len(names))
    dt = np.dtype([("flag", "u1")] + [(names[i], "<f4") for i in perm])
    rows = np.zeros(P, dt)
    for i in perm:
        rows[names[i]] = cols[:, i]
    hdr = "ply\nformat binary_little_endian 1.0\ncomment made by a test\nelement vertex %d\nproperty uchar flag\n" % P
    hdr += "".join(f"property float {names[i]}\n" for i in perm) + "element face 0\nproperty list uchar int vertex_indices\nend_header\n"
    path = tmp_path / "shuffled.ply"
    path.write_bytes(hdr.encode() + rows.tobytes())
    r = gtr.load_ply(str(path))
    for k in d:
        assert torch.equal(r[k], d[k]), k


def test_ply_errors(tmp_path):
    import gaussian_training as gtr
    with pytest.raises(RuntimeError, match="cannot open"):
        gtr.load_ply(str(tmp_path / "missing.ply"))
    bad = tmp_path / "ascii.ply"
    bad.write_bytes(b"ply\nformat ascii 1.0\nelement vertex 1\nproperty float x\nend_header\n0\n")
    with pytest.raises(RuntimeError, match="binary_little_endian"):
        gtr.load_ply(str(bad))
    d = _gaussians(5, 15)
    path = str(tmp_path / "t.ply")
    gtr.save_ply(path, *d.values())
    blob = open(path, "rb").read()
    open(path, "wb").write(blob[:-10])
    with pytest.raises(RuntimeError, match="shorter"):
        gtr.load_ply(path)
    open(path, "wb").write(blob.replace(b"property float opacity\n", b"property double opacity\n"))
    with pytest.raises(RuntimeError, match="opacity"):
        gtr.load_ply(path)


def test_next_header_symbols_exported_and_plain_c(tmp_path):
    hdr = open(os.path.join(ROOT, "include", "b200splat_next.h")).read()
    names = sorted(set(re.findall(r"B200SPLAT_API\s+[\w\s\*]+?\b(b200splat_\w+)\s*\(", hdr)))
    assert names == ["b200splat_activate_backward", "b200splat_activate_forward", "b200splat_adam_step", "b200splat_dist2_knn3",
                     "b200splat_dist2_knn3_static",
                     "b200splat_l1_ssim_backward", "b200splat_l1_ssim_forward", "b200splat_l1_ssim_partials", "b200splat_ply_read",
                     "b200splat_ply_read_header", "b200splat_ply_write", "b200splat_training_loss_backward",
                     "b200splat_training_loss_forward"]
    lib = ctypes.CDLL(os.path.join(ROOT, "3dgs_study_b200", "libb200splat.so"))
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200splat_next.h but not exported"
    lib.b200splat_l1_ssim_partials.restype = ctypes.c_size_t
    assert lib.b200splat_l1_ssim_partials(3, 1080, 1920) == 3 * 34 * 60
    c = tmp_path / "t.c"
    c.write_text('#include "b200splat_next.h"\nint main(void){ b200splat_adam_group g; (void)g; return sizeof(g) == 72 ? 0 : 1; }\n')
    assert os.system(f"/usr/bin/gcc -std=c99 -Wall -Werror -I{ROOT}/include {c} -o {tmp_path}/t && {tmp_path}/t") == 0
    from gaussian_training import _lib
    assert ctypes.sizeof(_lib.AdamGroup) == 72


def test_gpu_entry_points_fail_loudly_on_cpu_tensors():
    import gaussian_training as gtr
    from simple_knn._C import distCUDA2
    with pytest.raises(RuntimeError, match="CUDA"):
        gtr.l1_loss(torch.zeros(3, 8, 8), torch.zeros(3, 8, 8))
    with pytest.raises(RuntimeError, match="CUDA"):
        gtr.ssim(torch.zeros(3, 8, 8), torch.zeros(3, 8, 8))
    with pytest.raises(RuntimeError, match="CUDA"):
        distCUDA2(torch.zeros(10, 3))
    p = torch.nn.Parameter(torch.zeros(4))
    opt = gtr.FusedAdam([{"params": [p], "lr": 0.1, "name": "xyz"}], lr=0.0, eps=1e-15)
    p.grad = torch.ones(4)
    with pytest.raises(RuntimeError, match="CUDA"):
        opt.step()


def test_fused_adam_keeps_torch_optimizer_surface():
    """What the reference does to its optimizer besides step(): named groups, lr writes, state_dict round trip."""
    import gaussian_training as gtr
    p = torch.nn.Parameter(torch.zeros(4))
    opt = gtr.FusedAdam([{"params": [p], "lr": 0.1, "name": "xyz"}], lr=0.0, eps=1e-15)
    assert isinstance(opt, torch.optim.Adam) and opt.param_groups[0]["name"] == "xyz" and opt.defaults["eps"] == 1e-15
    opt.param_groups[0]["lr"] = 0.5
    sd = opt.state_dict()
    opt.load_state_dict(sd)
    assert opt.param_groups[0]["lr"] == 0.5
    opt.zero_grad(set_to_none=True)
    assert opt.step() is None                                  # nothing has a gradient: no launch, no error


def test_ply_written_by_the_reference_save_ply(tmp_path):
    """tests/golden/reference_save_ply_7.ply was produced by the reference's OWN GaussianModel.save_ply
    (/root/reference/scene/gaussian_model.py:218-258) run in the build container with only the absent `plyfile`
    writer restated (tests/golden/make_golden_ply.py).  The codec must read it back to the tensors that went in and
    write the identical bytes."""
    import gaussian_training as gtr
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    path = os.path.join(here, "reference_save_ply_7.ply")
    src = np.load(os.path.join(here, "reference_save_ply_7.npz"))
    r = gtr.load_ply(path)
    for k in ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation"):
        assert r[k].shape == src[k].shape, k
        assert np.array_equal(r[k].numpy(), src[k]), k          # features_rest comes back [P,15,3] from channel-major columns
    out = str(tmp_path / "again.ply")
    gtr.save_ply(out, *(r[k] for k in ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")))
    assert open(out, "rb").read() == open(path, "rb").read()
