"""Pins oracle/lbs_oracle.py (the CPU restatement of the LBS part of the hot path) against the golden
vectors the reference's own Python code produced (oracle/gen_golden_lbs.py -> tests/golden/lbs_*.npz)."""
import os

import numpy as np
import pytest
import torch

from oracle import lbs_oracle as LO

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _load(name):
    return dict(np.load(os.path.join(GOLD, name + ".npz")))


def _t(a, dt=torch.float32):
    return torch.from_numpy(np.asarray(a)).to(dt)


def lbs_inputs(g, smpl_model, dt=torch.float32):
    smpl = {k: (v.to(dt) if v.is_floating_point() else v) for k, v in smpl_model.items()}
    params = {"poses": _t(g["poses"], dt), "shapes": _t(g["shapes"], dt), "R": _t(g["R"], dt), "Th": _t(g["Th"], dt)}
    big = {"poses": _t(g["big_poses"], dt), "shapes": torch.zeros(1, 10, dtype=dt), "R": torch.eye(3, dtype=dt),
           "Th": torch.zeros(1, 3, dtype=dt)}
    lw = _t(g["lbs_weights"], dt) if g["lbs_weights"].size else None
    cr = _t(g["correct_Rs"], dt) if g["correct_Rs"].size else None
    return smpl, params, big, lw, cr


@pytest.mark.parametrize("name", ["lbs_plain_p1500", "lbs_offsets_p1200"])
@pytest.mark.parametrize("dt,tol", [(torch.float32, 2e-5), (torch.float64, 2e-5)])
def test_coarse_deform_matches_reference(name, dt, tol, smpl_model):
    g = _load(name)
    smpl, params, big, lw, cr = lbs_inputs(g, smpl_model, dt)
    out = LO.coarse_deform_c2source(smpl, _t(g["xyz"], dt), params, big, _t(g["t_vertices"], dt), lbs_weights=lw,
                                    correct_Rs=cr, return_transl=lw is not None, normals=_t(g["normals"], dt))
    smpl_pts, world_pts, bw, transforms, transl, world_n, vid = out
    np.testing.assert_allclose(smpl_pts.numpy(), g["out_smpl_pts"], atol=tol)
    np.testing.assert_allclose(world_pts.numpy(), g["out_world_pts"], atol=tol)
    np.testing.assert_allclose(bw.numpy(), g["out_bweights"], atol=1e-6)
    np.testing.assert_allclose(transforms.numpy(), g["out_transforms"], atol=tol)
    np.testing.assert_allclose(world_n.numpy(), g["out_world_normals"], atol=tol)
    if lw is not None:
        np.testing.assert_allclose(transl.numpy(), g["out_translation"], atol=tol)
    # chain matrices
    A_big, _, _ = LO.transform_params(smpl, big)
    np.testing.assert_allclose(A_big.numpy(), g["out_A_big"], atol=1e-6)
    A_t, _, joints = LO.transform_params(smpl, params)
    np.testing.assert_allclose(A_t.numpy(), g["out_A_target"], atol=2e-6)
    np.testing.assert_allclose(joints.numpy(), g["out_joints_target"], atol=1e-6)


@pytest.mark.parametrize("name", ["lbs_plain_p1500", "lbs_offsets_p1200"])
def test_covariance_and_sh_match_reference(name):
    g = _load(name)
    cov = LO.covariance(torch.exp(_t(g["log_scale"])), 1.0, _t(g["rotation"]), _t(g["out_transforms"]))
    np.testing.assert_allclose(cov.numpy(), g["out_cov3D"], rtol=1e-4, atol=1e-9)
    for deg in range(4):
        col = LO.view_colors(deg, _t(g["features"]), _t(g["out_world_pts"]), _t(g["campos"]))
        np.testing.assert_allclose(col.numpy(), g["out_colors"][deg], atol=2e-6)


def test_smplx_lbs_matches_reference(smpl_model):
    g = _load("lbs_smplx_v6890")
    s = smpl_model
    verts, joints, A, T = LO.smplx_lbs(_t(g["shapes"]).reshape(-1), _t(g["poses"]), s["v_template"], s["shapedirs"],
                                       s["posedirs"].t().contiguous(), s["J_regressor"],
                                       s["kintree_table"][0].tolist(), s["weights"])
    np.testing.assert_allclose(verts.numpy(), g["out_verts"], atol=5e-6)
    np.testing.assert_allclose(joints.numpy(), g["out_joints"], atol=2e-6)
    np.testing.assert_allclose(A.numpy(), g["out_A"], atol=2e-6)
    np.testing.assert_allclose(T[::53].numpy(), g["out_T_sample"], atol=2e-6)


def test_nearest_vertex_is_unambiguous(smpl_model):
    """The synthetic Gaussians are jittered off-vertex, so the knn_cuda stand-in has no distance ties."""
    g = _load("lbs_plain_p1500")
    q = _t(g["xyz"], torch.float64)
    tv = _t(g["t_vertices"], torch.float64)
    d = torch.cdist(q, tv)
    two = torch.topk(d, 2, dim=1, largest=False).values
    assert float((two[:, 1] - two[:, 0]).min()) > 1e-7
