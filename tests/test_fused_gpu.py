"""-m gpu: the fused per-Gaussian chain (mygauhuman_b200.fused.render_deformed: ghr_forward_deform_preprocess /
ghr_backward_deformed) against the two-operator chain it replaces (SMPLDeformer.deform_gaussians -> GaussianRasterizer,
each parity-tested against the oracle / the reference CUDA build elsewhere).

Bars: the intermediate values (means3D, cov3D, colours, opacity), radii, tiles_touched, point_offsets and the three
images BIT-EXACT (same device functions, the projection's fp32 sequence is pinned); gradients per element within
1e-4 rel (tests.common.grad_close) -- the fused backward skips the fp32 round trip of four intermediate gradients.
"""
import numpy as np
import pytest
import torch

from mygauhuman_b200 import synthetic as S
from tests.common import grad_close

pytestmark = pytest.mark.gpu

W, H = 200, 152


def _scene(smpl_model, P, seed, view_id=1, radius=3.0, W=W, H=H):
    from mygauhuman_b200.lbs import SMPLDeformer
    dev = torch.device("cuda")
    big = S.big_pose_params()
    tv = S.big_pose_vertices(smpl_model, big)
    g = S.make_gaussians(P, tv.numpy(), seed=seed, sh_degree=3)
    leaves = {k: g[k].to(dev).requires_grad_(True) for k in ("xyz", "log_scale", "rotation", "features", "opacity_logit",
                                                             "normals")}
    cam = S.make_camera(W, H, angle=0.6 * view_id, radius=radius, seed=view_id, focal=537.0 * W / 512)
    pose = {k: v.to(dev) for k, v in S.make_pose_params(view_id).items()}
    pose["Th"] = pose["Th"] * 0.2
    deformer = SMPLDeformer(smpl_model, dev)
    return dev, deformer, leaves, {k: v.to(dev) for k, v in big.items()}, tv.to(dev), cam, pose


def _settings(cam, dev, W=W, H=H):
    from mygauhuman_b200 import diff_gaussian_rasterization as dgr
    return dgr.GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
        bg=torch.tensor([0.1, 0.2, 0.3], device=dev), scale_modifier=1.0,
        viewmatrix=cam["world_view_transform"].to(dev), projmatrix=cam["full_proj_transform"].to(dev), sh_degree=3,
        campos=cam["camera_center"].to(dev), prefiltered=False, debug=False)


def _two_operator(deformer, L, pose, big, tv, rs, lbs_weights=None, correct_Rs=None):
    from mygauhuman_b200 import diff_gaussian_rasterization as dgr
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    d = deformer.deform_gaussians(L["xyz"], pose, big, tv, normals=L["normals"], log_scaling=L["log_scale"],
                                  rotation=L["rotation"], features=L["features"], sh_degree=3, campos=rs.campos,
                                  opacity_logit=L["opacity_logit"], lbs_weights=lbs_weights, correct_Rs=correct_Rs)
    means2D = torch.zeros_like(d["means3D"], requires_grad=True)
    color, radii, depth, alpha = dgr.GaussianRasterizer(rs)(means3D=d["means3D"], means2D=means2D, opacities=d["opacity"],
                                                            colors_precomp=d["colors"], cov3D_precomp=d["cov3D"])
    return color, radii, depth, alpha, d, means2D, _C._last_geom


def _geom_ints(geom, P, W, H):
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    st = _C.state_views(P, 0, W, H, geom, torch.empty(0, dtype=torch.uint8, device=geom.device),
                        torch.zeros(1 << 22, dtype=torch.uint8, device=geom.device))
    return st["tiles_touched"].clone(), st["point_offsets"].clone()


@pytest.mark.parametrize("P,view_id,radius", [(6000, 1, 3.0), (20000, 2, 1.1), (3000, 3, 0.35)])
def test_forward_is_bit_exact_with_the_two_operator_chain(smpl_model, P, view_id, radius):
    """radius 1.1 / 0.35: the camera sits close to / inside the body, so many Gaussians are frustum- or near-culled."""
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    from mygauhuman_b200.fused import render_deformed
    dev, deformer, L, big, tv, cam, pose = _scene(smpl_model, P, seed=view_id, view_id=view_id, radius=radius)
    rs = _settings(cam, dev)
    with torch.no_grad():
        c0, r0, d0, a0, mid, _, geom0 = _two_operator(deformer, L, pose, big, tv, rs)
        t0 = _geom_ints(geom0, P, W, H)
        c1, r1, d1, a1, aux = render_deformed(deformer, L["xyz"], pose, big, tv, rs, log_scaling=L["log_scale"],
                                              rotation=L["rotation"], features=L["features"],
                                              opacity_logit=L["opacity_logit"], normals=L["normals"], return_aux=True)
        t1 = _geom_ints(_C._last_geom, P, W, H)
    vis = r0 > 0
    assert 0 < int(vis.sum()) and (radius > 2 or int(vis.sum()) < P)
    assert torch.equal(r0, r1)
    assert torch.equal(t0[0], t1[0]) and torch.equal(t0[1], t1[1])
    assert torch.equal(aux["means3D"], mid["means3D"])
    assert torch.equal(aux["opacity"], mid["opacity"])
    assert torch.equal(aux["world_normals"], mid["world_normals"])
    assert torch.equal(aux["cov3D"][vis], mid["cov3D"][vis])
    assert torch.equal(aux["colors"][vis], mid["colors"][vis])
    assert torch.equal(c0, c1) and torch.equal(d0, d1) and torch.equal(a0, a1)


@pytest.mark.parametrize("P,view_id,radius,nets", [(6000, 1, 3.0, False), (5000, 2, 1.1, False), (4000, 4, 3.0, True)])
def test_backward_matches_the_two_operator_chain(smpl_model, P, view_id, radius, nets):
    """nets=True: LBS-weight logit offsets (LBSOffsetDecoder's output) and pose corrections (BodyPoseRefiner's) with
    gradients, i.e. the motion_offset_flag branch of render()."""
    from mygauhuman_b200.fused import render_deformed
    dev, deformer, L, big, tv, cam, pose = _scene(smpl_model, P, seed=10 + view_id, view_id=view_id, radius=radius)
    rs = _settings(cam, dev)
    gen = torch.Generator().manual_seed(7)
    w = [torch.randn(n, H, W, generator=gen).to(dev) for n in (3, 1, 1)]
    lw = cr = None
    if nets:
        lw = (0.3 * torch.randn(P, 24, generator=gen)).to(dev).requires_grad_(True)
        from mygauhuman_b200.lbs import batch_rodrigues
        cr = batch_rodrigues((0.05 * torch.randn(23, 3, generator=gen)).to(dev)).requires_grad_(True)
    extra = [t for t in (lw, cr) if t is not None]

    c0, r0, d0, a0, _, m2d0, _ = _two_operator(deformer, L, pose, big, tv, rs, lbs_weights=lw, correct_Rs=cr)
    ((c0 * w[0]).sum() + (d0 * w[1]).sum() + (a0 * w[2]).sum()).backward()
    want = {k: v.grad.clone() for k, v in L.items() if v.grad is not None}
    want["means2D"] = m2d0.grad.clone()
    want_extra = [t.grad.clone() for t in extra]
    for t in list(L.values()) + extra:
        t.grad = None

    m2d1 = torch.zeros(P, 3, device=dev, requires_grad=True)
    c1, r1, d1, a1 = render_deformed(deformer, L["xyz"], pose, big, tv, rs, log_scaling=L["log_scale"],
                                     rotation=L["rotation"], features=L["features"], opacity_logit=L["opacity_logit"],
                                     lbs_weights=lw, correct_Rs=cr, means2D=m2d1)
    assert torch.equal(c0, c1) and torch.equal(r0, r1)
    ((c1 * w[0]).sum() + (d1 * w[1]).sum() + (a1 * w[2]).sum()).backward()
    got = {k: v.grad for k, v in L.items() if v.grad is not None}
    got["means2D"] = m2d1.grad
    for k in ("xyz", "log_scale", "rotation", "features", "opacity_logit", "means2D"):
        grad_close(k, got[k].cpu().numpy(), want[k].cpu().numpy())
    culled = (r0 == 0)
    if int(culled.sum()):
        for k in ("xyz", "log_scale", "rotation", "features", "opacity_logit"):
            assert float(got[k][culled].abs().max()) == 0.0, k
    for name, t, wg in zip(("lbs_weights", "correct_Rs"), extra, want_extra):
        grad_close(name, t.grad.cpu().numpy(), wg.cpu().numpy())


def test_graphed_step_fused_equals_unfused(smpl_model):
    """The timed configuration (CUDA-graph replay, gradients accumulated into a flat bucket) on both chains."""
    from mygauhuman_b200.graphed import GraphedRenderStep, pack_view
    P = 6000
    dev, deformer, L, big, tv, cam, pose = _scene(smpl_model, P, seed=3)
    vec = pack_view(cam["world_view_transform"], cam["full_proj_transform"], cam["camera_center"],
                    {k: v.cpu() for k, v in pose.items()}).to(dev)
    targets = torch.randn(5, H, W, generator=torch.Generator().manual_seed(5)).to(dev)
    loss_fn = lambda color, depth, alpha, t: (torch.cat([color, depth, alpha]) * t).sum()
    res = {}
    for fused in (False, True):
        step = GraphedRenderStep(deformer, L, big, tv, H, W, cam["tanfovx"], cam["tanfovy"], torch.zeros(3, device=dev),
                                 loss_fn, accumulate=True, fused=fused)
        step.capture([vec], [targets])
        step.zero_grads()
        loss = step(vec, targets)
        step(vec, targets)  # two views accumulated
        torch.cuda.synchronize()
        res[fused] = (float(loss), {k: v.clone() for k, v in step.flat_views.items()}, step.kernels_per_replay)
    assert abs(res[True][0] - res[False][0]) <= 1e-6 * max(1.0, abs(res[False][0]))
    for k in L:
        grad_close(k, res[True][1][k].cpu().numpy(), res[False][1][k].cpu().numpy())
    assert res[True][2] == res[False][2] - 2, res  # two libghr kernels fewer per view


@pytest.mark.parametrize("n_views,with_corrections", [(1, False), (3, True), (8, False), (11, True)])
def test_batched_pose_stage_equals_the_per_view_kernels(smpl_model, n_views, with_corrections):
    """ghr_lbs_pose_views (one launch for the views of a batch; 11 views = two launches) against ghr_lbs_pose +
    ghr_lbs_vertex_offsets per view: pose buffers and vertex offsets bit for bit."""
    from mygauhuman_b200 import _lib
    from mygauhuman_b200.lbs import SMPLDeformer, batch_rodrigues
    dev = torch.device("cuda")
    deformer = SMPLDeformer(smpl_model, dev)
    big = {k: v.to(dev) for k, v in S.big_pose_params().items()}
    gen = torch.Generator().manual_seed(n_views)
    views, want = [], []
    for i in range(n_views):
        pose = {k: v.to(dev).float().contiguous() for k, v in S.make_pose_params(20 + i).items()}
        cr = None
        if with_corrections and i % 2 == 0:
            cr = batch_rodrigues((0.05 * torch.randn(23, 3, generator=gen)).to(dev))
        buf = deformer.pose_buffer(pose, cr)
        off = deformer.vertex_offsets(deformer._big_pose_buffer(big), buf, pose["shapes"])
        want.append((buf, off))
        views.append(dict(poses=pose["poses"].reshape(-1), shapes=pose["shapes"].reshape(-1), R=pose["R"].reshape(-1),
                          Th=pose["Th"].reshape(-1), correct_Rs=cr))
    outs = [(torch.full((_lib.POSE_FLOATS,), float("nan"), device=dev), torch.full((deformer.V, 3), float("nan"), device=dev))
            for _ in range(n_views)]
    deformer.pose_views(big, views, outs)
    torch.cuda.synchronize()
    segments = ((_lib.POSE_A, 288), (_lib.POSE_ROT, 216), (_lib.POSE_PF, 207), (_lib.POSE_JOINTS, 72), (_lib.POSE_RGINV, 9),
                (_lib.POSE_RG, 9), (_lib.POSE_TH, 3), (_lib.POSE_G, 288), (_lib.POSE_ROD, 216))  # what the kernels write
    for (gb, go), (wb, wo) in zip(outs, want):
        for o, n in segments:
            assert torch.equal(gb[o:o + n], wb[o:o + n]), o
        assert torch.equal(go, wo)


def test_edge_cases_empty_single_and_fully_culled(smpl_model):
    """P = 0 (the reference returns its zero-filled outputs, rasterize_points.cu:83), P = 1, and a camera that looks away
    from the body (every Gaussian culled: background image, zero gradients, nothing added)."""
    from mygauhuman_b200.fused import render_deformed
    dev, deformer, L, big, tv, cam, pose = _scene(smpl_model, 300, seed=4)
    rs = _settings(cam, dev)
    kw = lambda n: dict(log_scaling=L["log_scale"][:n], rotation=L["rotation"][:n], features=L["features"][:n],
                        opacity_logit=L["opacity_logit"][:n])
    c, r, d, a = render_deformed(deformer, L["xyz"][:0], pose, big, tv, rs, **kw(0))
    assert r.numel() == 0 and float(c.abs().max()) == 0.0 and float(a.abs().max()) == 0.0
    c1, r1, d1, a1 = render_deformed(deformer, L["xyz"][:1], pose, big, tv, rs, **kw(1))
    c2, r2, _, _, _, _, _ = _two_operator(deformer, {k: v[:1] for k, v in L.items()}, pose, big, tv, rs)
    assert torch.equal(r1, r2) and torch.equal(c1, c2)
    # camera behind the body, looking away: flip the view direction by moving the camera through the body
    away = S.make_camera(W, H, angle=0.6, radius=-3.0, seed=1, focal=537.0 * W / 512)
    rs2 = _settings(away, dev)
    for t in L.values():
        t.grad = None
    c3, r3, d3, a3 = render_deformed(deformer, L["xyz"], pose, big, tv, rs2, **kw(300))
    if int((r3 > 0).sum()) == 0:
        assert torch.allclose(c3, rs2.bg[:, None, None].expand_as(c3)) and float(a3.abs().max()) == 0.0
        (c3.sum() + a3.sum()).backward()
        for k in ("xyz", "log_scale", "rotation", "features", "opacity_logit"):
            assert float(L[k].grad.abs().max()) == 0.0, k
    else:  # the synthetic camera still sees part of the body: at least the chains must agree
        c4, r4, _, _, _, _, _ = _two_operator(deformer, L, pose, big, tv, rs2)
        assert torch.equal(r3, r4) and torch.equal(c3, c4)


def test_densification_statistics_fused_into_the_backward(smpl_model):
    """render_deformed(densify=...) updates max_radii2D / xyz_gradient_accum / denom inside the backward kernel; the
    reference does it with three indexed statements after backward (train.py:401-404, gaussian_model.py:764-766), which
    optim.add_densification_stats (ghr_densify_stats, pinned against those statements) restates."""
    from mygauhuman_b200.fused import render_deformed
    from mygauhuman_b200.optim import add_densification_stats
    P = 5000
    dev, deformer, L, big, tv, cam, pose = _scene(smpl_model, P, seed=6, view_id=2, radius=1.3)
    rs = _settings(cam, dev)
    w = torch.randn(3, H, W, generator=torch.Generator().manual_seed(3)).to(dev)
    got = [torch.zeros(P, device=dev), torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev)]
    want = [t.clone() for t in got]
    got[0][::7] = 40.0  # some Gaussians already hold a larger radius than this view gives them
    want[0][::7] = 40.0
    for _ in range(2):  # two views accumulate
        m2d = torch.zeros(P, 3, device=dev, requires_grad=True)
        color, radii, _, _ = render_deformed(deformer, L["xyz"], pose, big, tv, rs, log_scaling=L["log_scale"],
                                             rotation=L["rotation"], features=L["features"], opacity_logit=L["opacity_logit"],
                                             means2D=m2d, densify=tuple(got))
        (color * w).sum().backward()
        add_densification_stats(radii, m2d.grad.contiguous(), *want)
    vis = radii > 0
    assert 0 < int(vis.sum()) < P
    assert torch.equal(got[0], want[0]) and torch.equal(got[2], want[2])
    assert float((got[1] - want[1]).abs().max()) <= 1e-6 * float(want[1].abs().max())
    assert float(got[2][~vis].abs().max()) == 0.0 and float(got[2][vis].min()) == 2.0
