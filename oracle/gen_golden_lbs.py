"""ORACLE -- TEST INFRASTRUCTURE ONLY.

Generates tests/golden/lbs_*.npz by running the REFERENCE's own Python LBS code
(/root/reference, dev container only) on the CPU:

* scene/gaussian_model.py  GaussianModel.coarse_deform_c2source (:768-872),
  get_transform_params_torch (:947-980), GaussianModel.get_covariance (:208-209)
* utils/sh_utils.py        eval_sh (:135-190)
* smplx/lbs.py             lbs (:156-252)   (BASELINE.json config 1)

The reference module hard-wires CUDA (``.cuda()``, ``device="cuda"``) and
imports packages that are absent here (knn_cuda, simple_knn, plyfile, trimesh);
those imports are stubbed, ``Tensor.cuda`` is made the identity and
``device="cuda"`` factory kwargs are redirected to the CPU.  No reference
arithmetic is replaced except knn_cuda's nearest-neighbour search (third-party,
not vendored), for which an fp64 brute-force argmin stands in.

Run:  python oracle/gen_golden_lbs.py     (needs /root/reference)
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("GAUHUMAN_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)


def _stub(name, **attrs):
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def import_reference():
    class _KNN:  # knn_cuda.KNN(k, transpose_mode=True)(ref[1,V,3], query[1,P,3]) -> (dist, idx[1,P,k])
        def __init__(self, k, transpose_mode=True):
            self.k = k

        def __call__(self, ref, query):
            d = torch.cdist(query.double(), ref.double())
            dist, idx = torch.topk(d, self.k, dim=-1, largest=False)
            return dist.float(), idx

    _stub("knn_cuda", KNN=_KNN)
    _stub("simple_knn")
    _stub("simple_knn._C", distCUDA2=None)
    _stub("plyfile", PlyData=None, PlyElement=None)
    _stub("trimesh")
    torch.Tensor.cuda = lambda self, *a, **k: self
    for fname in ("zeros", "ones", "eye", "empty", "tensor", "full"):
        orig = getattr(torch, fname)

        def wrap(*a, __orig=orig, **k):
            if str(k.get("device", "")).startswith("cuda"):
                k["device"] = "cpu"
            return __orig(*a, **k)
        setattr(torch, fname, wrap)
    sys.path.insert(0, REF)
    import importlib.util
    import utils.sh_utils as shu
    # load the module by path: scene/__init__.py pulls in the dataset readers (imageio, cv2, ...)
    spec = importlib.util.spec_from_file_location("ref_gaussian_model", os.path.join(REF, "scene", "gaussian_model.py"))
    gm = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gm)
    # smplx/lbs.py does `from .utils import ...`: register a bare package so the relative import resolves
    # without running smplx/__init__.py (body models, licensed assets)
    pkg = _stub("smplx")
    pkg.__path__ = [os.path.join(REF, "smplx")]
    import importlib
    lbs_mod = importlib.import_module("smplx.lbs")
    return gm, shu, lbs_mod


def main():
    from mygauhuman_b200 import synthetic as S
    gm, shu, lbs_mod = import_reference()
    out_dir = os.path.join(ROOT, "tests", "golden")
    smpl = S.make_smpl_model(0)
    model = gm.GaussianModel.__new__(gm.GaussianModel)
    model.setup_functions()
    model.SMPL_NEUTRAL = {k: v.clone() for k, v in smpl.items()}
    model.knn = sys.modules["knn_cuda"].KNN(k=1)

    big = S.big_pose_params()
    # big-pose "world" vertices the cameras carry (dataset_readers.py:596-600): the SMPL surface in the big pose
    A_big, _, _, _ = gm.get_transform_params_torch(model.SMPL_NEUTRAL, big)
    t_vertices = (torch.einsum("vj,jab->vab", smpl["weights"], A_big[0])[:, :3, :3] @ smpl["v_template"][:, :, None])[:, :, 0] \
        + torch.einsum("vj,jab->vab", smpl["weights"], A_big[0])[:, :3, 3]

    for name, P, seed, use_offsets in (("lbs_plain_p1500", 1500, 3, False), ("lbs_offsets_p1200", 1200, 4, True)):
        g = S.make_gaussians(P, t_vertices.numpy(), seed=seed, sh_degree=3)
        params = S.make_pose_params(seed)
        gen = torch.Generator().manual_seed(100 + seed)
        lbs_w = (torch.randn(1, P, 24, generator=gen) * 0.3) if use_offsets else None
        cr = None
        if use_offsets:
            cr = gm.batch_rodrigues_torch(torch.randn(23, 3, generator=gen) * 0.05)[None]
        with torch.no_grad():
            smpl_pts, world_pts, bw, transforms, transl, world_n = model.coarse_deform_c2source(
                g["xyz"][None], params, big, t_vertices[None], lbs_weights=lbs_w, correct_Rs=cr,
                return_transl=use_offsets, normals=g["normals"][None])
            scaling = torch.exp(g["log_scale"])
            model._scaling = g["log_scale"]
            model._rotation = g["rotation"] * 1.7  # un-normalised on purpose: build_rotation normalises
            cov = model.get_covariance(1.0, transforms.squeeze(0))
            campos = torch.tensor([0.3, 1.1, 2.8])
            d = world_pts[0] - campos
            d = d / d.norm(dim=1, keepdim=True)
            cols = []
            for deg in range(4):
                rgb = shu.eval_sh(deg, g["features"].transpose(1, 2), d)
                cols.append(torch.clamp_min(rgb + 0.5, 0.0))
            A_t, _, _, joints_t = gm.get_transform_params_torch(model.SMPL_NEUTRAL, params)
        np.savez_compressed(
            os.path.join(out_dir, name + ".npz"),
            xyz=g["xyz"].numpy(), normals=g["normals"].numpy(), log_scale=g["log_scale"].numpy(),
            rotation=(g["rotation"] * 1.7).numpy(), features=g["features"].numpy(), campos=campos.numpy(),
            poses=params["poses"].numpy(), shapes=params["shapes"].numpy(), R=params["R"].numpy(), Th=params["Th"].numpy(),
            big_poses=big["poses"].numpy(), t_vertices=t_vertices.numpy(),
            lbs_weights=np.zeros(0, np.float32) if lbs_w is None else lbs_w[0].numpy(),
            correct_Rs=np.zeros(0, np.float32) if cr is None else cr[0].numpy(),
            out_smpl_pts=smpl_pts[0].numpy(), out_world_pts=world_pts[0].numpy(), out_bweights=bw[0].numpy(),
            out_transforms=transforms[0].numpy(),
            out_translation=np.zeros(0, np.float32) if transl is None else transl[0].numpy(),
            out_world_normals=world_n[0].numpy(), out_cov3D=cov.numpy(),
            out_colors=torch.stack(cols).numpy(), out_A_big=A_big[0].numpy(), out_A_target=A_t[0].numpy(),
            out_joints_target=joints_t[0].numpy(), seed=seed, smpl_seed=0)
        print("wrote", name, "world_pts range", float(world_pts.min()), float(world_pts.max()))

    # BASELINE.json config 1: smplx/lbs.py lbs() over the 6,890 template vertices
    params = S.make_pose_params(7)
    with torch.no_grad():
        verts, Jt, A, T = lbs_mod.lbs(params["shapes"], params["poses"], smpl["v_template"][None], smpl["shapedirs"],
                                      smpl["posedirs"].t().contiguous(), smpl["J_regressor"],
                                      smpl["kintree_table"][0].clone(), smpl["weights"])
    np.savez_compressed(os.path.join(out_dir, "lbs_smplx_v6890.npz"), poses=params["poses"].numpy(),
                        shapes=params["shapes"].numpy(), out_verts=verts[0].numpy(), out_joints=Jt[0].numpy(),
                        out_A=A[0].numpy(), out_T_sample=T[0, ::53].numpy(), smpl_seed=0, seed=7)
    print("wrote lbs_smplx_v6890")


if __name__ == "__main__":
    main()
