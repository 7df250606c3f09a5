"""-m gpu: capture()/restore() interchange with the reference (scene/gaussian_model.py:106-150).  A run is
started on the reference's own GaussianModel + torch.optim.Adam (imported from baseline/_ref), snapshotted with its
capture() through torch.save, resumed on libghr (checkpoint.restore -> FusedAdam), and the two continue in lockstep; then
the other way round.  Tolerance: 2e-7 of the operands per step (one fp32 rounding), as in tests/test_optim_gpu.py."""
import io
import types

import pytest
import torch

from mygauhuman_b200 import synthetic as S

pytestmark = pytest.mark.gpu

P = 4000


def _training_args():
    return types.SimpleNamespace(percent_dense=0.01, position_lr_init=0.00016, position_lr_final=0.0000016,
                                 position_lr_delay_mult=0.01, position_lr_max_steps=30_000, feature_lr=0.0025,
                                 opacity_lr=0.05, scaling_lr=0.005, rotation_lr=0.001, pose_refine_lr=0.00005,
                                 lbs_offset_lr=0.00005, normal_lr=0.0002)


def _ref_model(seed):
    from oracle import ref_py
    assert ref_py.available(), "baseline/_ref is missing (oracle/install_ref_py.sh, run by build())"
    ns = ref_py.load()
    smpl = S.make_smpl_model(0)
    tv = S.big_pose_vertices(smpl, S.big_pose_params())
    g = S.make_gaussians(P, tv.numpy(), seed=seed, sh_degree=3)
    gen = torch.Generator().manual_seed(seed)
    g["albedo"] = torch.randn(P, 3, generator=gen)
    g["roughness"] = torch.randn(P, 1, generator=gen)
    par = lambda t: torch.nn.Parameter(t.cuda().contiguous())
    leaves = {k: par(v) for k, v in g.items()}
    m = ref_py.make_model(smpl, leaves, "cuda", motion_offset=True)
    m._features_dc = par(g["features"][:, :1])
    m._features_rest = par(g["features"][:, 1:])
    m.spatial_lr_scale = 1.7
    m.max_radii2D = torch.zeros(P, device="cuda")
    torch.manual_seed(seed)
    m.pose_decoder = ns.nets_pose.BodyPoseRefiner(total_bones=24, embedding_size=69, mlp_width=128, mlp_depth=2).cuda()
    m.lweight_offset_decoder = ns.nets_lbs.LBSOffsetDecoder(total_bones=24).cuda()
    return m


def _params_of(model_or_state):
    g = (lambda k: getattr(model_or_state, "_" + k)) if not isinstance(model_or_state, dict) else (lambda k: model_or_state[k])
    ps = [g(k) for k in ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")]
    dec = (model_or_state["pose_decoder"], model_or_state["lweight_offset_decoder"]) if isinstance(model_or_state, dict) \
        else (model_or_state.pose_decoder, model_or_state.lweight_offset_decoder)
    ps += list(dec[0].parameters()) + list(dec[1].parameters())
    ps += [g(k) for k in ("normal", "albedo", "roughness")]
    return ps


def _set_grads(params, seed):
    gen = torch.Generator().manual_seed(seed)
    for p in params:
        if p.requires_grad:
            p.grad = (torch.randn(p.shape, generator=gen) * 1e-2).to(p.device)


def _close(a, b):
    for x, y in zip(a, b):
        assert x.shape == y.shape
        tol = 4e-7 * max(1.0, float(y.abs().max()))
        assert float((x.detach() - y.detach()).abs().max()) <= tol


def test_reference_checkpoint_resumes_on_libghr_and_back():
    from mygauhuman_b200 import checkpoint as Ck
    from mygauhuman_b200.optim import FusedAdam
    ta = _training_args()
    ref = _ref_model(1)
    ref.training_setup(ta)
    assert isinstance(ref.optimizer, torch.optim.Adam)
    for it in range(2):
        _set_grads(_params_of(ref), 10 + it)
        ref.optimizer.step()
    buf = io.BytesIO()
    torch.save((ref.capture(), 2), buf)                      # train.py's snapshot
    buf.seek(0)
    model_args, first_iter = torch.load(buf, weights_only=False)
    assert first_iter == 2 and len(model_args) == len(Ck.FIELDS)
    state, opt = Ck.restore(model_args, ta)
    assert isinstance(opt, FusedAdam) and [g["name"] for g in opt.param_groups] == [g["name"] for g in ref.optimizer.param_groups]
    assert [g["lr"] for g in opt.param_groups] == [g["lr"] for g in ref.optimizer.param_groups]
    ours = _params_of(state)
    _close(ours, _params_of(ref))
    for it in range(3):                                     # both continue: same gradients, same parameters after
        _set_grads(_params_of(ref), 20 + it)
        _set_grads(ours, 20 + it)
        ref.optimizer.step()
        opt.step()
        _close(ours, _params_of(ref))
    # the other way round: our capture() -> the reference's restore()
    buf = io.BytesIO()
    torch.save((Ck.capture(state, opt), 5), buf)
    buf.seek(0)
    model_args, _ = torch.load(buf, weights_only=False)
    back = _ref_model(2)
    back.restore(model_args, ta)
    assert isinstance(back.optimizer, torch.optim.Adam)
    _close(_params_of(back), ours)
    for it in range(2):
        _set_grads(_params_of(back), 30 + it)
        _set_grads(ours, 30 + it)
        back.optimizer.step()
        opt.step()
        _close(ours, _params_of(back))
    with pytest.raises(ValueError):
        Ck.restore(model_args[:5], ta)
