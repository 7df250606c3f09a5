"""-m gpu: the LBS-weight offset network on the tensor cores (mygauhuman_b200.nets.LBSOffsetDecoder ->
ghr_lbs_mlp_forward, csrc/ghr_mlp.cu) against

  * tests/golden/lbs_mlp.npz: the reference's own module run in fp32 on the CPU (oracle/gen_golden_mlp.py);
  * the same layers in fp64 on the GPU (the truth) and in fp32 under torch's DEFAULT settings, which is how the
    reference runs them: cuDNN convolutions with TF32 allowed.

The kernels compute with 11-bit-significand operands (FP16 by default, TF32 with GHR_MLP_KERNEL=tf32) and fp32
accumulation, so the bar is stated against the truth: max |error| <=
3e-3 of the largest logit (TF32 keeps 11 significant bits; five layers of 63..191-term dot products), and not worse than
the reference's own default-precision error by more than a small factor.  The measured errors are printed.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "lbs_mlp.npz")
REL = 3e-3


def _module_from_golden(dev):
    from mygauhuman_b200.nets import LBSOffsetDecoder
    g = np.load(GOLDEN)
    m = LBSOffsetDecoder()
    m.load_state_dict({k[6:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("param.")})
    return m.to(dev), g


def _truth(m, pts):
    ps = [p.detach().double() for p in m._params()]
    with torch.no_grad():
        return m.torch_forward(pts.double(), ps)


def test_golden_vector_of_the_reference_module():
    dev = torch.device("cuda")
    m, g = _module_from_golden(dev)
    pts = torch.from_numpy(g["pts"]).to(dev)
    with torch.no_grad():
        out = m(pts)
    want = torch.from_numpy(g["out"]).to(dev)
    assert out.shape == want.shape == (1, 24, pts.shape[1])
    scale = float(want.abs().max())
    err = float((out - want).abs().max())
    print(f"golden: max |err| {err:.3e} of scale {scale:.3f}")
    assert err <= REL * scale
    # the caller's permute gives row-major [1, P, 24] logits (what the deform kernels read without a copy)
    assert out.permute(0, 2, 1)[0].is_contiguous()


@pytest.mark.parametrize("P", [1, 127, 128, 129, 1000, 50_000])
def test_matches_fp64_truth_within_tf32_and_the_reference_default_precision(P):
    dev = torch.device("cuda")
    m, _ = _module_from_golden(dev)
    gen = torch.Generator().manual_seed(P)
    pts = ((torch.rand(1, P, 3, generator=gen) * 2 - 1) * torch.tensor([0.9, 1.1, 0.4])).to(dev)
    with torch.no_grad():
        out = m(pts)
        truth = _truth(m, pts)
        ref_default = m.torch_forward(pts)  # torch defaults: cuDNN conv with TF32 allowed, as the reference runs it
    torch.cuda.synchronize()
    scale = float(truth.abs().max())
    err = float((out.double() - truth).abs().max())
    err_ref = float((ref_default.double() - truth).abs().max())
    print(f"P={P}: ours {err:.3e}, reference default precision {err_ref:.3e}, scale {scale:.3f}")
    assert torch.isfinite(out).all()
    assert err <= REL * scale
    if P >= 127:  # (a single point does not take cuDNN's TF32 path: torch computes it in plain fp32, error 2e-7)
        assert err <= 2.0 * err_ref + 2e-4 * scale


def test_repacks_when_the_parameters_change_and_backward_matches_torch():
    dev = torch.device("cuda")
    m, _ = _module_from_golden(dev)
    pts = ((torch.rand(1, 300, 3, generator=torch.Generator().manual_seed(1)) * 2 - 1)).to(dev)
    with torch.no_grad():
        a = m(pts).clone()
        m.bw_fc.bias.add_(0.25)                      # an optimiser step
        m.bw_linears[1].weight.mul_(1.5)
        b = m(pts)
        want = _truth(m, pts)
    assert float((b - a).abs().max()) > 0.2
    assert float((b.double() - want).abs().max()) <= REL * float(want.abs().max())
    # backward: parameter gradients through the recomputed torch layers == torch autograd of the same layers
    w = torch.randn(1, 24, 300, generator=torch.Generator().manual_seed(2)).to(dev)
    (m(pts) * w).sum().backward()
    got = [p.grad.clone() for p in m._params()]
    for p in m.parameters():
        p.grad = None
    (m.torch_forward(pts) * w).sum().backward()
    for g_, p in zip(got, m._params()):
        assert torch.allclose(g_, p.grad, rtol=1e-4, atol=1e-6 * float(p.grad.abs().max()))


def test_logits_drive_the_deformation(smpl_model):
    """End of the row: the kernel's [P,24] logits as `lbs_weights` of the fused deformation, against fp32 torch logits."""
    from mygauhuman_b200 import synthetic as S
    from mygauhuman_b200.lbs import SMPLDeformer
    dev = torch.device("cuda")
    m, _ = _module_from_golden(dev)
    big = S.big_pose_params()
    tv = S.big_pose_vertices(smpl_model, big)
    P = 4000
    g = S.make_gaussians(P, tv.numpy(), seed=2, sh_degree=3)
    xyz = g["xyz"].to(dev)
    pose = {k: v.to(dev) for k, v in S.make_pose_params(3).items()}
    deformer = SMPLDeformer(smpl_model, dev)
    bigd = {k: v.to(dev) for k, v in big.items()}
    with torch.no_grad():
        lw = m(xyz[None]).permute(0, 2, 1)[0]
        lw_ref = _truth(m, xyz[None]).permute(0, 2, 1)[0].float()
        a = deformer.deform_gaussians(xyz, pose, bigd, tv.to(dev), lbs_weights=lw)
        b = deformer.deform_gaussians(xyz, pose, bigd, tv.to(dev), lbs_weights=lw_ref)
    d = float((a["means3D"] - b["means3D"]).abs().max())
    print(f"world points move by {d:.3e} m under the TF32 logits")
    assert d <= 2e-3


def test_body_pose_refiner_matches_the_reference_golden_and_feeds_the_pose_stage(smpl_model):
    """BodyPoseRefiner (nets/mlp_delta_body_pose.py) as one kernel: fp32 throughout, so the bar is fp32 round-off (1e-5 abs
    on rotation-matrix entries) against the reference module's CPU output; gradients through the torch recompute."""
    from mygauhuman_b200 import synthetic as S
    from mygauhuman_b200.lbs import SMPLDeformer
    from mygauhuman_b200.nets import BodyPoseRefiner
    dev = torch.device("cuda")
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "pose_refiner.npz"))
    m = BodyPoseRefiner(mlp_width=128, mlp_depth=2)
    m.load_state_dict({k[6:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("param.")})
    m = m.to(dev)
    pose = torch.from_numpy(g["pose"]).to(dev)
    Rs = m(pose)["Rs"]
    want = torch.from_numpy(g["Rs"]).to(dev)
    assert Rs.shape == want.shape == (1, 23, 3, 3)
    err = float((Rs - want).abs().max())
    print(f"pose refiner: max |err| {err:.3e}")
    assert err <= 1e-5
    assert float((want - torch.eye(3, device=dev)).abs().max()) > 0.1  # the golden is not the identity
    w = torch.randn(1, 23, 3, 3, generator=torch.Generator().manual_seed(4)).to(dev)
    (Rs * w).sum().backward()
    got = [p.grad.clone() for p in m._params()]
    for p in m.parameters():
        p.grad = None
    (m.torch_forward(pose) * w).sum().backward()
    for a, p in zip(got, m._params()):
        assert torch.allclose(a, p.grad, rtol=1e-4, atol=1e-6 * float(p.grad.abs().max()))
    # its output is the correct_Rs input of the pose stage
    deformer = SMPLDeformer(smpl_model, dev)
    params = {k: v.to(dev) for k, v in S.make_pose_params(2).items()}
    buf = deformer.pose_buffer(params, Rs.detach()[0])
    assert torch.isfinite(buf[:288]).all()
