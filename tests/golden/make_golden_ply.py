#!/usr/bin/env python
"""Generates tests/golden/reference_save_ply_7.ply (+ .npz of the tensors that went in) by running the REFERENCE'S
OWN `GaussianModel.save_ply` (/root/reference/scene/gaussian_model.py:240-258, with its construct_list_of_attributes,
:218-238) in this container.  Run here only (it imports /root/reference); the outputs are committed.

What is the reference's and what is restated:
  * reference, unmodified: the attribute list and its order (x, y, z, nx, ny, nz, f_dc_*, f_rest_*, opacity, scale_*,
    rot_*), the transpose(1, 2).flatten of the SH tensors (which makes f_rest CHANNEL-major), the zero normals, the
    np.concatenate of the columns, the structured array with dtype 'f4' per attribute.
  * restated (the `plyfile` package is absent from this image and from /root/reference -- requirements name it, no
    wheel is available offline): `PlyElement.describe(elements, 'vertex')` + `PlyData([el]).write(path)`.  The
    stand-in below follows the PLY 1.0 format as plyfile 0.8/1.0 emits it for a little-endian structured array:
        ply\\n format binary_little_endian 1.0\\n element vertex <N>\\n property float <name>\\n ... end_header\\n
    followed by the array's raw bytes (`elements.tofile`).  numpy 'f4' maps to the PLY type name `float`; plyfile
    writes no comment lines unless asked and defaults to the machine's byte order ('<' here).
So the golden file pins everything the codec must agree on with the reference's writer except the five header
keywords of the PLY container itself; DESIGN.md section 7 says so.
"""
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


class _PlyElement:
    def __init__(self, data, name):
        self.data, self.name = data, name

    @staticmethod
    def describe(data, name):
        assert data.dtype.names and all(data.dtype[n] == np.dtype("<f4") for n in data.dtype.names)
        return _PlyElement(data, name)


class _PlyData:
    def __init__(self, elements):
        self.elements = elements

    def write(self, path):
        with open(path, "wb") as f:
            f.write(b"ply\nformat binary_little_endian 1.0\n")
            for el in self.elements:
                f.write(f"element {el.name} {len(el.data)}\n".encode("ascii"))
                for n in el.data.dtype.names:
                    f.write(f"property float {n}\n".encode("ascii"))
            f.write(b"end_header\n")
            for el in self.elements:
                el.data.tofile(f)


def main():
    sys.path.insert(0, REF)
    ply = types.ModuleType("plyfile")
    ply.PlyData, ply.PlyElement = _PlyData, _PlyElement
    sys.modules["plyfile"] = ply
    for name in ("simple_knn", "simple_knn._C"):
        m = types.ModuleType(name)
        m.distCUDA2 = None
        sys.modules[name] = m
    from scene.gaussian_model import GaussianModel
    g = torch.Generator().manual_seed(7)
    P, deg = 7, 3
    M = (deg + 1) ** 2
    gm = GaussianModel(deg)
    gm._xyz = torch.nn.Parameter(torch.randn(P, 3, generator=g))
    gm._features_dc = torch.nn.Parameter(torch.randn(P, 1, 3, generator=g))          # [P, 1, 3] as create_from_pcd leaves it
    gm._features_rest = torch.nn.Parameter(torch.randn(P, M - 1, 3, generator=g))    # [P, 15, 3]
    gm._opacity = torch.nn.Parameter(torch.randn(P, 1, generator=g))
    gm._scaling = torch.nn.Parameter(torch.randn(P, 3, generator=g))
    gm._rotation = torch.nn.Parameter(torch.randn(P, 4, generator=g))
    out = os.path.join(HERE, "reference_save_ply_7.ply")
    gm.save_ply(out)
    np.savez(os.path.join(HERE, "reference_save_ply_7.npz"), xyz=gm._xyz.detach().numpy(), features_dc=gm._features_dc.detach().numpy(),
             features_rest=gm._features_rest.detach().numpy(), opacity=gm._opacity.detach().numpy(),
             scaling=gm._scaling.detach().numpy(), rotation=gm._rotation.detach().numpy())
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
