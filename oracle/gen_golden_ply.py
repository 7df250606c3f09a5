"""ORACLE -- TEST INFRASTRUCTURE ONLY.

Generates tests/golden/ply_p37.ply and ply_p37.npz by running the REFERENCE's own GaussianModel.save_ply and
GaussianModel.load_ply (/root/reference/scene/gaussian_model.py:329-407; dev container only) on seeded parameters.
``plyfile`` is absent from this image: its three calls (PlyElement.describe, PlyData([...]).write, PlyData.read) are
stubbed with oracle/ply_oracle.py's restatement of the container format; every tensor-to-row statement that runs is
the reference's.

Run:  python oracle/gen_golden_ply.py     (needs /root/reference)
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import gen_golden_lbs, ply_oracle  # noqa: E402


class _Prop:
    def __init__(self, name):
        self.name = name


class _Element:
    def __init__(self, data):
        self.data = data
        self.properties = [_Prop(n) for n in data.dtype.names]

    def __getitem__(self, name):
        return self.data[name]


class PlyElement:
    @staticmethod
    def describe(elements, name):
        assert name == 'vertex'
        return _Element(elements)


class PlyData:
    def __init__(self, elements):
        self.elements = elements

    def write(self, path):
        d = self.elements[0].data
        names = list(d.dtype.names)
        rows = np.stack([d[n] for n in names], axis=1)
        ply_oracle.write_ply(path, rows, names)

    @staticmethod
    def read(path):
        rows, names = ply_oracle.read_ply(path)
        arr = np.empty(rows.shape[0], dtype=[(n, 'f4') for n in names])
        for i, n in enumerate(names):
            arr[n] = rows[:, i]
        return PlyData([_Element(arr)])


def main(P=37, seed=0):
    gm, _, _ = gen_golden_lbs.import_reference()
    gm.PlyData, gm.PlyElement = PlyData, PlyElement
    g = torch.Generator().manual_seed(seed)
    model = gm.GaussianModel.__new__(gm.GaussianModel)
    model.max_sh_degree = 3
    mk = lambda *s: torch.nn.Parameter(torch.randn(*s, generator=g))
    model._xyz, model._normal, model._albedo, model._roughness = mk(P, 3), mk(P, 3), mk(P, 3), mk(P, 1)
    model._features_dc, model._features_rest = mk(P, 1, 3), mk(P, 15, 3)
    model._opacity, model._scaling, model._rotation = mk(P, 1), mk(P, 3), mk(P, 4)
    inputs = dict(xyz=model._xyz, normal=model._normal, albedo=model._albedo, roughness=model._roughness,
                  features_dc=model._features_dc, features_rest=model._features_rest, opacity=model._opacity,
                  scaling=model._scaling, rotation=model._rotation)
    inputs = {"in_" + k: v.detach().numpy().copy() for k, v in inputs.items()}
    out_dir = os.path.join(ROOT, "tests", "golden")
    path = os.path.join(out_dir, "ply_p37.ply")
    model.save_ply(path)                      # reference statements :329-348
    loaded = gm.GaussianModel.__new__(gm.GaussianModel)
    loaded.max_sh_degree = 3
    loaded.load_ply(path)                     # reference statements :355-407
    outs = dict(xyz=loaded._xyz, normal=loaded._normal, albedo=loaded._albedo, roughness=loaded._roughness,
                features_dc=loaded._features_dc, features_rest=loaded._features_rest, opacity=loaded._opacity,
                scaling=loaded._scaling, rotation=loaded._rotation)
    outs = {"out_" + k: v.detach().numpy().copy() for k, v in outs.items()}
    np.savez_compressed(os.path.join(out_dir, "ply_p37.npz"), **inputs, **outs)
    print("wrote", path, os.path.getsize(path), "bytes; load_ply(save_ply(x)) == x:",
          all(np.array_equal(inputs["in_" + k[4:]], v) for k, v in outs.items()))


if __name__ == "__main__":
    main()
