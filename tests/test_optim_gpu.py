"""-m gpu parity of the optimiser-side kernels: ghr_adam_step through FusedAdam vs the numpy oracle on
torch.optim.Adam's golden vectors and vs torch.optim.Adam on the GPU; ghr_densify_stats vs the oracle / the reference's
statements.  Tolerance: fp32 rounding of each update (2e-7 of the operands' magnitude), written per assert."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden", "optim_adam.npz")
LR = {"xyz": 0.00016, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}
SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}


def close(a, b, scale, eps=2e-7):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return bool(np.all(np.abs(a - b) <= eps * np.maximum(np.abs(b), scale)))


def reference_groups(params):
    return [{"params": [params[n]], "lr": LR[n], "name": n} for n in LR]  # scene/gaussian_model.py:256-263


def test_fused_adam_on_torch_adam_goldens():
    from mygauhuman_b200.optim import FusedAdam
    from oracle import optim_oracle as OO
    z = np.load(GOLD)
    params = {n: torch.nn.Parameter(torch.from_numpy(z[f"p0_{n}"]).cuda()) for n in LR}
    opt = FusedAdam(reference_groups(params), lr=0.0, eps=1e-15)
    for it in range(4):
        before = {n: (params[n].detach().cpu().numpy().copy(),) for n in LR}
        for n in LR:
            params[n].grad = torch.from_numpy(z[f"g{it}_{n}"]).cuda()
        opt.param_groups[0]["lr"] = float(z["xyz_lr"][it])
        opt.step()
        for n in LR:
            st = opt.state[params[n]]
            g, p0 = z[f"g{it}_{n}"], before[n][0]
            m0 = z[f"m{it}_{n}"] if it else np.zeros_like(p0)
            v0 = z[f"v{it}_{n}"] if it else np.zeros_like(p0)
            lr = float(z["xyz_lr"][it]) if n == "xyz" else LR[n]
            assert int(st["step"]) == it + 1
            # vs torch.optim.Adam's own output
            assert close(st["exp_avg"].cpu().numpy(), z[f"m{it + 1}_{n}"], np.maximum(np.abs(m0), np.abs(g) * 0.1)), (n, it)
            assert close(st["exp_avg_sq"].cpu().numpy(), z[f"v{it + 1}_{n}"], np.abs(v0)), (n, it)
            assert close(params[n].detach().cpu().numpy(), z[f"p{it + 1}_{n}"], np.maximum(np.abs(p0), lr), 4e-7), (n, it)
            # vs the oracle on the same inputs
            po, mo, vo = OO.adam_step(p0, g, m0, v0, it + 1, lr)
            assert close(params[n].detach().cpu().numpy(), po, np.maximum(np.abs(p0), lr), 4e-7), (n, it)
            # chain from the golden state: rounding differences must not accumulate into the next comparison
            with torch.no_grad():
                params[n].copy_(torch.from_numpy(z[f"p{it + 1}_{n}"]))
                st["exp_avg"].copy_(torch.from_numpy(z[f"m{it + 1}_{n}"]))
                st["exp_avg_sq"].copy_(torch.from_numpy(z[f"v{it + 1}_{n}"]))


@pytest.mark.parametrize("P", [1, 511, 50_000])
def test_fused_adam_tracks_torch_adam_on_gpu(P):
    from mygauhuman_b200.optim import FusedAdam
    g = torch.Generator().manual_seed(P)
    init = {n: torch.randn((P,) + s, generator=g) for n, s in SHAPES.items()}
    ours = {n: torch.nn.Parameter(t.clone().cuda()) for n, t in init.items()}
    ref = {n: torch.nn.Parameter(t.clone().cuda()) for n, t in init.items()}
    o1 = FusedAdam(reference_groups(ours), lr=0.0, eps=1e-15)
    o2 = torch.optim.Adam(reference_groups(ref), lr=0.0, eps=1e-15)
    for it in range(6):
        for n in LR:
            grad = (torch.randn(init[n].shape, generator=g) * 10.0 ** float(torch.randint(-6, 1, (1,), generator=g))).cuda()
            grad[torch.rand(P, generator=g).cuda() < 0.3] = 0
            ours[n].grad, ref[n].grad = grad.clone(), grad.clone()
        lr = 0.00016 * 0.9 ** it
        o1.param_groups[0]["lr"] = lr
        o2.param_groups[0]["lr"] = lr
        o1.step(zero_grads=(it % 2 == 1))
        o2.step()
        for n in LR:
            if it % 2 == 1:
                assert float(ours[n].grad.abs().max()) == 0.0  # cleared in the same pass, tensor kept
            # trajectories: six steps of fp32 rounding in the moments, each step's update is <= ~lr
            d = (ours[n].detach() - ref[n].detach()).abs().max().item()
            assert d <= 2e-6 * max(LR[n], 1e-4) * (it + 1) + 3e-7 * ref[n].detach().abs().max().item(), (n, it, d)
            for k in ("exp_avg", "exp_avg_sq"):  # relative to the tensor's scale: single sums can cancel
                a, b = o1.state[ours[n]][k], o2.state[ref[n]][k]
                assert (a - b).abs().max().item() <= 3e-7 * (it + 1) * b.abs().max().item(), (n, it, k)


def test_fused_adam_unaligned_views_ragged_tail_and_many_tensors():
    from mygauhuman_b200.optim import FusedAdam
    g = torch.Generator().manual_seed(3)
    base = torch.randn(100_000, generator=g).cuda()
    sizes = [1, 2, 3, 5, 2047, 2048, 2049, 4097, 7, 9, 11, 13, 15, 17, 19, 21, 23, 8190, 33]  # > 16 tensors: two launches
    ours, ref, o = [], [], 1  # offset 1: 4-byte aligned only -> scalar path
    for s in sizes:
        ours.append(torch.nn.Parameter(base[o:o + s].clone()))
        ref.append(torch.nn.Parameter(base[o:o + s].clone()))
        o += s
    big = torch.randn(2048 * 5 + 1, generator=g).cuda()
    pv = big[1:]                         # misaligned data pointer
    ours.append(torch.nn.Parameter(pv.clone()))
    ours[-1].data = big.clone()[1:]      # parameter storage itself starts 4 bytes into an allocation
    ref.append(torch.nn.Parameter(pv.clone()))
    o1 = FusedAdam(ours, lr=1e-2, betas=(0.8, 0.99), eps=1e-8)
    o2 = torch.optim.Adam(ref, lr=1e-2, betas=(0.8, 0.99), eps=1e-8)
    for it in range(3):
        for a, b in zip(ours, ref):
            gr = torch.randn(a.shape, generator=g).cuda()
            a.grad, b.grad = gr.clone(), gr.clone()
        o1.step()
        o2.step()
    for a, b in zip(ours, ref):
        torch.testing.assert_close(a.detach(), b.detach(), rtol=2e-6, atol=2e-7)


def test_fused_adam_state_survives_densification_style_rekeying():
    """_prune_optimizer / cat_tensors_to_optimizer (scene/gaussian_model.py:423-483) cut and extend the state tensors
    and re-key them under a new nn.Parameter; the optimiser must carry on from them."""
    from mygauhuman_b200.optim import FusedAdam
    g = torch.Generator().manual_seed(5)
    P = 1000
    x1 = torch.nn.Parameter(torch.randn(P, 3, generator=g).cuda())
    x2 = torch.nn.Parameter(x1.detach().clone())
    o1 = FusedAdam([{"params": [x1], "lr": 1e-3, "name": "xyz"}], lr=0.0, eps=1e-15)
    o2 = torch.optim.Adam([{"params": [x2], "lr": 1e-3, "name": "xyz"}], lr=0.0, eps=1e-15)

    def step(xa, xb, n):
        gr = torch.randn(n, 3, generator=g).cuda()
        xa.grad, xb.grad = gr.clone(), gr.clone()
        o1.step()
        o2.step()

    step(x1, x2, P)
    mask = (torch.rand(P, generator=g) < 0.7).cuda()
    ext = torch.randn(50, 3, generator=g).cuda()
    new = []
    for opt, x in ((o1, x1), (o2, x2)):
        group = opt.param_groups[0]
        st = opt.state.get(group["params"][0], None)
        st["exp_avg"] = torch.cat((st["exp_avg"][mask], torch.zeros_like(ext)), dim=0)
        st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"][mask], torch.zeros_like(ext)), dim=0)
        del opt.state[group["params"][0]]
        group["params"][0] = torch.nn.Parameter(torch.cat((x.detach()[mask], ext), dim=0).requires_grad_(True))
        opt.state[group["params"][0]] = st
        new.append(group["params"][0])
    step(new[0], new[1], new[0].shape[0])
    assert int(o1.state[new[0]]["step"]) == 2
    torch.testing.assert_close(new[0].detach(), new[1].detach(), rtol=2e-6, atol=1e-7)


def test_fused_adam_rejects_cpu_parameters():
    from mygauhuman_b200.optim import FusedAdam
    p = torch.nn.Parameter(torch.zeros(4))
    p.grad = torch.ones(4)
    with pytest.raises(ValueError):
        FusedAdam([p], lr=1e-3).step()


def test_adam_abi_argument_checks():
    from mygauhuman_b200 import _lib
    l = _lib.lib()
    t = (_lib.AdamGroup * 1)()
    t[0].count, t[0].step = 4, 1
    assert l.ghr_adam_step(1, t, 0.9, 0.999, 1e-15, 0, None) == _lib.GHR_ERR_BAD_ARG      # null pointers
    assert l.ghr_adam_step(17, t, 0.9, 0.999, 1e-15, 0, None) == _lib.GHR_ERR_BAD_ARG     # too many groups
    assert l.ghr_adam_step(0, None, 0.9, 0.999, 1e-15, 0, None) == _lib.GHR_OK
    x = torch.zeros(4, device="cuda")
    t[0].param = t[0].grad = t[0].exp_avg = t[0].exp_avg_sq = x.data_ptr()
    t[0].step = 0
    assert l.ghr_adam_step(1, t, 0.9, 0.999, 1e-15, 0, None) == _lib.GHR_ERR_BAD_ARG      # step counts from 1
    t[0].step = 1
    assert l.ghr_adam_step(1, t, 1.0, 0.999, 1e-15, 0, None) == _lib.GHR_ERR_BAD_ARG      # beta1 < 1


def test_densify_stats_matches_reference_statements_and_oracle():
    from mygauhuman_b200.optim import add_densification_stats
    from oracle import optim_oracle as OO
    z = np.load(GOLD)
    cu = lambda k: torch.from_numpy(z[k]).cuda()
    mr, acc, dn = cu("ds_max0"), cu("ds_acc0"), cu("ds_den0")
    add_densification_stats(cu("ds_radii"), cu("ds_vgrad"), mr, acc, dn)
    assert np.array_equal(mr.cpu().numpy(), z["ds_max1"])
    assert np.array_equal(dn.cpu().numpy(), z["ds_den1"])
    np.testing.assert_allclose(acc.cpu().numpy(), z["ds_acc1"], rtol=2e-7)
    o = OO.densify_stats(z["ds_radii"], z["ds_vgrad"], z["ds_max0"], z["ds_acc0"], z["ds_den0"])
    np.testing.assert_allclose(acc.cpu().numpy(), o[1], rtol=2e-7)
    # optional outputs, empty input
    acc2 = cu("ds_acc0")
    add_densification_stats(cu("ds_radii"), cu("ds_vgrad"), None, acc2, None)
    assert torch.equal(acc2, acc)
    add_densification_stats(torch.zeros(0, dtype=torch.int32, device="cuda"), torch.zeros(0, 3, device="cuda"))


def test_densify_stats_after_a_real_backward():
    """End of a rendered iteration: radii from the forward, viewspace gradient from the backward (means2D.grad)."""
    from mygauhuman_b200.optim import add_densification_stats
    from tests import common as Cm
    case = Cm.raster_case(3000, 128, 96, seed=4, mode="precomp")
    tc = Cm.to_torch(case, "cuda:0")
    fw = Cm.ours_forward(tc)
    grads = Cm.ours_backward(tc, fw)
    radii, dmean2D = fw[4], grads[0]
    P = radii.shape[0]
    mr, acc, dn = torch.zeros(P, device="cuda"), torch.zeros(P, 1, device="cuda"), torch.zeros(P, 1, device="cuda")
    for _ in range(2):
        add_densification_stats(radii, dmean2D, mr, acc, dn)
    vis = radii > 0
    assert vis.any()  # (culled Gaussians are covered by the golden-vector test above)
    assert torch.equal(mr, torch.where(vis, radii.float(), torch.zeros_like(mr)))
    assert torch.equal(dn[:, 0], 2.0 * vis.float())
    want = 2.0 * torch.norm(dmean2D[:, :2], dim=-1, keepdim=True) * vis[:, None]
    torch.testing.assert_close(acc, want, rtol=1e-6, atol=0)
