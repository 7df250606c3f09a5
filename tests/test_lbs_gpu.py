"""-m gpu parity tests of the LBS part of the hot path: libghr's CUDA kernels (through
mygauhuman_b200.lbs, i.e. the C ABI) against the golden vectors produced by the reference's own
Python LBS (tests/golden/lbs_*.npz) and against the CPU oracle (oracle/lbs_oracle.py) in fp64.

Tolerances (fp32 arithmetic on both sides; BASELINE.json north_star): positions / transforms / colours
2e-5 abs (means feed the rasteriser whose images must agree to 1e-5), gradients 1e-4 relative to the
tensor's scale; nearest-vertex indices exact.
"""
import os

import numpy as np
import pytest
import torch

from tests.common import rel_err
from tests.test_lbs_oracle_cpu import _load, _t, lbs_inputs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def deformer(smpl_model):
    from mygauhuman_b200.lbs import SMPLDeformer
    return SMPLDeformer(smpl_model, "cuda")


def _cuda_params(g):
    dev = "cuda"
    params = {"poses": _t(g["poses"]).to(dev), "shapes": _t(g["shapes"]).to(dev), "R": _t(g["R"]).to(dev),
              "Th": _t(g["Th"]).to(dev)}
    big = {"poses": _t(g["big_poses"]).to(dev), "shapes": torch.zeros(1, 10, device=dev), "R": torch.eye(3, device=dev),
           "Th": torch.zeros(1, 3, device=dev)}
    return params, big


@pytest.mark.parametrize("name", ["lbs_plain_p1500", "lbs_offsets_p1200"])
def test_pose_chain_matches_reference(name, deformer):
    g = _load(name)
    params, big = _cuda_params(g)
    cr = _t(g["correct_Rs"]).cuda()[None] if g["correct_Rs"].size else None
    A, R, Th, joints = deformer.get_transform_params_torch(params)
    np.testing.assert_allclose(A[0].cpu().numpy(), g["out_A_target"], atol=3e-6)
    np.testing.assert_allclose(joints[0].cpu().numpy(), g["out_joints_target"], atol=2e-6)
    A_big, *_ = deformer.get_transform_params_torch(big)
    np.testing.assert_allclose(A_big[0].cpu().numpy(), g["out_A_big"], atol=3e-6)


@pytest.mark.parametrize("name", ["lbs_plain_p1500", "lbs_offsets_p1200"])
def test_coarse_deform_matches_reference(name, deformer):
    g = _load(name)
    params, big = _cuda_params(g)
    lw = _t(g["lbs_weights"]).cuda()[None] if g["lbs_weights"].size else None
    cr = _t(g["correct_Rs"]).cuda()[None] if g["correct_Rs"].size else None
    out = deformer.coarse_deform_c2source(_t(g["xyz"]).cuda()[None], params, big, _t(g["t_vertices"]).cuda()[None],
                                          lbs_weights=lw, correct_Rs=cr, return_transl=lw is not None,
                                          normals=_t(g["normals"]).cuda()[None])
    smpl_pts, world_pts, bw, transforms, transl, world_n = out
    assert world_pts.shape == (1, g["xyz"].shape[0], 3) and transforms.shape[-2:] == (3, 3)
    tol = 2e-5
    np.testing.assert_allclose(smpl_pts[0].cpu().numpy(), g["out_smpl_pts"], atol=tol)
    np.testing.assert_allclose(world_pts[0].cpu().numpy(), g["out_world_pts"], atol=tol)
    np.testing.assert_allclose(bw[0].cpu().numpy(), g["out_bweights"], atol=2e-6)
    np.testing.assert_allclose(transforms[0].cpu().numpy(), g["out_transforms"], atol=tol)
    np.testing.assert_allclose(world_n[0].cpu().numpy(), g["out_world_normals"], atol=tol)
    if lw is not None:
        np.testing.assert_allclose(transl[0].cpu().numpy(), g["out_translation"], atol=tol)
    else:
        assert transl is None


@pytest.mark.parametrize("name", ["lbs_plain_p1500", "lbs_offsets_p1200"])
def test_fused_cov_and_colour_match_reference(name, deformer):
    g = _load(name)
    params, big = _cuda_params(g)
    lw = _t(g["lbs_weights"]).cuda() if g["lbs_weights"].size else None
    cr = _t(g["correct_Rs"]).cuda()[None] if g["correct_Rs"].size else None
    for deg in range(4):
        d = deformer.deform_gaussians(_t(g["xyz"]).cuda(), params, big, _t(g["t_vertices"]).cuda(),
                                      normals=_t(g["normals"]).cuda(), lbs_weights=lw, correct_Rs=cr,
                                      scaling=torch.exp(_t(g["log_scale"])).cuda(), rotation=_t(g["rotation"]).cuda(),
                                      scaling_modifier=1.0, features=_t(g["features"]).cuda(), sh_degree=deg,
                                      campos=_t(g["campos"]).cuda())
        np.testing.assert_allclose(d["colors"].cpu().numpy(), g["out_colors"][deg], atol=5e-6)
    np.testing.assert_allclose(d["cov3D"].cpu().numpy(), g["out_cov3D"], rtol=2e-4, atol=1e-9)
    np.testing.assert_allclose(d["means3D"].cpu().numpy(), g["out_world_pts"], atol=2e-5)


def test_strided_lbs_weights(deformer):
    """The reference passes lweight_offset_decoder(...).permute(0,2,1): a [1,P,24] view of a [1,24,P] tensor."""
    g = _load("lbs_offsets_p1200")
    params, big = _cuda_params(g)
    lw = _t(g["lbs_weights"]).cuda()           # [P,24]
    lw_t = lw.t().contiguous()[None]            # [1,24,P] as the decoder emits it
    cr = _t(g["correct_Rs"]).cuda()[None]
    out = deformer.coarse_deform_c2source(_t(g["xyz"]).cuda()[None], params, big, _t(g["t_vertices"]).cuda()[None],
                                          lbs_weights=lw_t.permute(0, 2, 1), correct_Rs=cr, normals=_t(g["normals"]).cuda()[None])
    np.testing.assert_allclose(out[1][0].cpu().numpy(), g["out_world_pts"], atol=2e-5)


def test_nearest_vertex_exact(deformer, smpl_model):
    g = _load("lbs_plain_p1500")
    tv = _t(g["t_vertices"])
    gen = torch.Generator().manual_seed(5)
    P = 20000
    q = tv[torch.randint(0, tv.shape[0], (P,), generator=gen)] + torch.randn(P, 3, generator=gen) * 0.02
    q[:500] = torch.randn(500, 3, generator=gen) * 3.0           # far outside the body's bounding box
    q[500:600] = tv[:100]                                          # exactly on vertices
    vid = deformer.nearest_vertex(tv.cuda(), q.cuda()).cpu().numpy()
    d = torch.cdist(q.double(), tv.double())
    ref = d.argmin(1).numpy()
    two = torch.topk(d, 2, dim=1, largest=False).values
    clear = ((two[:, 1] - two[:, 0]) > 1e-6).numpy()              # skip fp32-ambiguous near-ties
    assert clear.mean() > 0.99
    assert np.array_equal(vid[clear], ref[clear])
    assert vid.min() >= 0 and vid.max() < tv.shape[0]


def test_smplx_lbs_matches_reference(deformer):
    g = _load("lbs_smplx_v6890")
    verts, joints, A, T = deformer.lbs(_t(g["shapes"]).cuda(), _t(g["poses"]).cuda())
    np.testing.assert_allclose(verts[0].cpu().numpy(), g["out_verts"], atol=1e-5)
    np.testing.assert_allclose(joints[0].cpu().numpy(), g["out_joints"], atol=3e-6)
    np.testing.assert_allclose(A[0].cpu().numpy(), g["out_A"], atol=3e-6)
    np.testing.assert_allclose(T[0, ::53].cpu().numpy(), g["out_T_sample"], atol=3e-6)


def test_batch_rodrigues(deformer):
    from mygauhuman_b200.lbs import batch_rodrigues
    from oracle import lbs_oracle as LO
    gen = torch.Generator().manual_seed(2)
    r = torch.randn(50, 3, generator=gen)
    r[3] = 0.0
    np.testing.assert_allclose(batch_rodrigues(r.cuda()).cpu().numpy(), LO.rodrigues(r).numpy(), atol=2e-6)


@pytest.mark.parametrize("name", ["lbs_plain_p1500", "lbs_offsets_p1200"])
def test_backward_matches_fp64_autograd(name, deformer, smpl_model):
    """Gradients of a random linear functional of all differentiable outputs vs. torch.autograd through the
    fp64 oracle (same nearest-vertex ids)."""
    from oracle import lbs_oracle as LO
    g = _load(name)
    P = g["xyz"].shape[0]
    use = g["lbs_weights"].size > 0
    gen = torch.Generator().manual_seed(11)
    seeds = {k: torch.randn(*s, generator=gen, dtype=torch.float64) for k, s in
             (("world", (P, 3)), ("T", (P, 3, 3)), ("n", (P, 3)), ("cov", (P, 6)), ("col", (P, 3)))}
    seeds["cov"] *= 1e3  # covariances are ~1e-4: give them weight in the functional

    # ---- ours (fp32, CUDA) ----
    params, big = _cuda_params(g)
    leaf = lambda a: _t(a).cuda().requires_grad_(True)
    xyz, nrm, logs, rot, feat = leaf(g["xyz"]), leaf(g["normals"]), leaf(g["log_scale"]), leaf(g["rotation"]), leaf(g["features"])
    lw = leaf(g["lbs_weights"]) if use else None
    cr = leaf(g["correct_Rs"]) if use else None
    d = deformer.deform_gaussians(xyz, params, big, _t(g["t_vertices"]).cuda(), normals=nrm, lbs_weights=lw,
                                  correct_Rs=None if cr is None else cr[None], scaling=torch.exp(logs), rotation=rot,
                                  scaling_modifier=1.3, features=feat, sh_degree=3, campos=_t(g["campos"]).cuda())
    c = lambda k: seeds[k].float().cuda()
    loss = ((d["means3D"] * c("world")).sum() + (d["transforms"] * c("T")).sum() + (d["world_normals"] * c("n")).sum() +
            (d["cov3D"] * c("cov")).sum() + (d["colors"] * c("col")).sum())
    loss.backward()
    vid = deformer.nearest_vertex(_t(g["t_vertices"]).cuda(), xyz.detach()).cpu().long()

    # ---- oracle (fp64, CPU autograd) ----
    dt = torch.float64
    smpl, oparams, obig, _, _ = lbs_inputs(g, smpl_model, dt)
    oleaf = lambda a: _t(a, dt).requires_grad_(True)
    oxyz, onrm, ologs, orot, ofeat = oleaf(g["xyz"]), oleaf(g["normals"]), oleaf(g["log_scale"]), oleaf(g["rotation"]), oleaf(g["features"])
    olw = oleaf(g["lbs_weights"]) if use else None
    ocr = oleaf(g["correct_Rs"]) if use else None
    o = LO.coarse_deform_c2source(smpl, oxyz, oparams, obig, _t(g["t_vertices"], dt), lbs_weights=olw, correct_Rs=ocr,
                                  normals=onrm, vert_ids=vid)
    _, world, _, T, _, wn, _ = o
    cov = LO.covariance(torch.exp(ologs), 1.3, orot, T)
    col = LO.view_colors(3, ofeat, world, _t(g["campos"], dt))
    oloss = ((world * seeds["world"]).sum() + (T * seeds["T"]).sum() + (wn * seeds["n"]).sum() +
             (cov * seeds["cov"]).sum() + (col * seeds["col"]).sum())
    oloss.backward()
    assert abs(float(loss) - float(oloss)) <= 2e-4 * max(1.0, abs(float(oloss)))

    pairs = [("xyz", xyz, oxyz), ("normals", nrm, onrm), ("log_scale", logs, ologs), ("rotation", rot, orot),
             ("features", feat, ofeat)]
    if use:
        pairs += [("lbs_weights", lw, olw), ("correct_Rs", cr, ocr)]
    for nm, a, b in pairs:
        assert a.grad is not None, nm
        e = rel_err(a.grad.cpu().numpy(), b.grad.numpy())
        assert e <= 1e-4, (nm, e)


def test_fused_activations_match_separate_ops(deformer):
    """log_scaling= / opacity_logit= (exp and sigmoid inside the deform kernels) give what exp / sigmoid applied
    outside give: outputs to 1e-6 relative, gradients to 1e-5 of each tensor's scale."""
    g = _load("lbs_plain_p1500")
    P = g["xyz"].shape[0]
    params, big = _cuda_params(g)
    tv = _t(g["t_vertices"]).cuda()
    gen = torch.Generator().manual_seed(4)
    logit0 = torch.randn(P, 1, generator=gen)
    w_cov, w_op = (torch.randn(P, 6, generator=gen) * 1e3).cuda(), torch.randn(P, 1, generator=gen).cuda()

    def run(fused):
        leaf = lambda a: a.clone().cuda().requires_grad_(True)
        xyz, logs, rot, logit = leaf(_t(g["xyz"])), leaf(_t(g["log_scale"])), leaf(_t(g["rotation"])), leaf(logit0)
        kw = dict(rotation=rot, scaling_modifier=1.1)
        if fused:
            d = deformer.deform_gaussians(xyz, params, big, tv, log_scaling=logs, opacity_logit=logit, **kw)
            op = d["opacity"]
        else:
            d = deformer.deform_gaussians(xyz, params, big, tv, scaling=torch.exp(logs), **kw)
            assert d["opacity"] is None
            op = torch.sigmoid(logit)
        ((d["cov3D"] * w_cov).sum() + (op * w_op).sum()).backward()
        return d["cov3D"].detach(), op.detach(), logs.grad, logit.grad, rot.grad, xyz.grad

    a, b = run(True), run(False)
    assert a[1].shape == (P, 1)
    for x, y, tol in zip(a, b, (1e-6, 1e-6, 1e-5, 1e-5, 1e-5, 1e-5)):
        assert rel_err(x.cpu().numpy(), y.cpu().numpy()) <= tol


@pytest.mark.parametrize("name", ["lbs_plain_p1500", "lbs_offsets_p1200"])
def test_standalone_covariance_and_colours_match_reference_goldens(name):
    """get_covariance / eval_sh_colors on their own against the reference's golden outputs."""
    from mygauhuman_b200.lbs import eval_sh_colors, get_covariance
    g = _load(name)
    cov = get_covariance(torch.exp(_t(g["log_scale"])).cuda(), _t(g["rotation"]).cuda(), 1.0, _t(g["out_transforms"]).cuda())
    np.testing.assert_allclose(cov.cpu().numpy(), g["out_cov3D"], rtol=2e-4, atol=1e-9)
    for deg in range(4):
        col = eval_sh_colors(deg, _t(g["features"]).cuda(), _t(g["out_world_pts"]).cuda(), _t(g["campos"]).cuda())
        np.testing.assert_allclose(col.cpu().numpy(), g["out_colors"][deg], atol=5e-6)


def test_standalone_covariance_and_colours_backward_vs_fp64_autograd():
    from mygauhuman_b200.lbs import eval_sh_colors, get_covariance
    from oracle import lbs_oracle as LO
    gen = torch.Generator().manual_seed(11)
    P = 2000
    s = torch.exp(torch.randn(P, 3, generator=gen) * 0.3 - 4.0)
    q = torch.randn(P, 4, generator=gen) * 1.3
    T = torch.eye(3)[None] + 0.3 * torch.randn(P, 3, 3, generator=gen)
    gc = torch.randn(P, 6, generator=gen)
    for use_T in (True, False):
        a = [t.clone().cuda().requires_grad_(True) for t in (s, q, T)]
        cov = get_covariance(a[0], a[1], 1.3, a[2] if use_T else None)
        cov.backward(gc.cuda())
        b = [t.clone().double().requires_grad_(True) for t in (s, q, T)]
        ref = LO.covariance(b[0], 1.3, b[1], b[2] if use_T else None)
        ref.backward(gc.double())
        assert _rel(cov.detach().cpu().numpy(), ref.detach().numpy()) < 1e-5  # fp32 vs fp64, off-diagonals cancel
        for x, y, n in zip(a, b, ("scaling", "rotation", "transforms")):
            if n == "transforms" and not use_T:
                assert x.grad is None
                continue
            assert _rel(x.grad.cpu().numpy(), y.grad.numpy()) < 1e-4, n
    sh = torch.cat([torch.rand(P, 1, 3, generator=gen) * 6 - 3, torch.randn(P, 15, 3, generator=gen) * 0.3], 1)
    m = torch.randn(P, 3, generator=gen)
    cam = torch.tensor([0.2, 1.0, 3.0])
    gcol = torch.randn(P, 3, generator=gen)
    for deg in range(4):
        a = [t.clone().cuda().requires_grad_(True) for t in (sh, m)]
        col = eval_sh_colors(deg, a[0], a[1], cam.cuda())
        col.backward(gcol.cuda())
        b = [t.clone().double().requires_grad_(True) for t in (sh, m)]
        ref = LO.view_colors(deg, b[0], b[1], cam.double())
        ref.backward(gcol.double())
        np.testing.assert_allclose(col.detach().cpu().numpy(), ref.detach().numpy(), atol=5e-6)
        assert (col == 0).any()  # some channels are clamped: their gradient must vanish
        assert _rel(a[0].grad.cpu().numpy(), b[0].grad.numpy()) < 1e-4, deg
        if b[1].grad is None:  # degree 0: no view dependence
            assert float(a[1].grad.abs().max()) == 0.0
        else:
            assert _rel(a[1].grad.cpu().numpy(), b[1].grad.numpy()) < 1e-4, deg
    with pytest.raises(RuntimeError):
        get_covariance(s, q)  # CPU tensors: no fallback


def _rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))
