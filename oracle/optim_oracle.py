"""ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by the product path).

CPU restatement (numpy, fp32 element arithmetic) of the optimiser side of the reference's training iteration:

* ``adam_step``      torch.optim.Adam(l, lr=0.0, eps=1e-15).step() as the reference configures it
                     (/root/reference/scene/gaussian_model.py:283, train.py:418): torch/optim/adam.py
                     ``_single_tensor_adam`` with weight_decay=0, amsgrad=False, maximize=False.  torch is a
                     third-party dependency of the reference (not vendored; 2.11.0 in this image); its published
                     algorithm is restated here.
* ``densify_stats``  train.py:401-404 + GaussianModel.add_densification_stats (scene/gaussian_model.py:764-766).

PINNED: tests/test_optim_oracle_cpu.py checks both against tests/golden/optim_adam.npz, produced by running
torch.optim.Adam itself and the reference's add_densification_stats statements (oracle/gen_golden_optim.py).
"""
import numpy as np

F = np.float32


def adam_step(param, grad, exp_avg, exp_avg_sq, step, lr, beta1=0.9, beta2=0.999, eps=1e-15):
    """One step for one tensor; returns (param, exp_avg, exp_avg_sq) as new fp32 arrays.  `step` is the 1-based
    count AFTER the increment (torch/optim/adam.py: step_t += 1 comes first)."""
    p, g, m, v = (np.asarray(a, dtype=F) for a in (param, grad, exp_avg, exp_avg_sq))
    m = m + (g - m) * F(1.0 - beta1)                       # exp_avg.lerp_(grad, 1 - beta1)
    v = v * F(beta2) + F(1.0 - beta2) * g * g              # exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    bc1 = 1.0 - beta1 ** step                              # Python doubles
    bc2 = 1.0 - beta2 ** step
    step_size = lr / bc1
    denom = np.sqrt(v) / F(bc2 ** 0.5) + F(eps)            # (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)
    p = p + F(-step_size) * (m / denom)                    # param.addcdiv_(exp_avg, denom, value=-step_size)
    return p.astype(F), m.astype(F), v.astype(F)


def densify_stats(radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom):
    """Returns the three updated arrays (copies).  radii int32 [P]; viewspace_grad [P,3]; max_radii2D [P];
    xyz_gradient_accum, denom [P,1]."""
    vis = np.asarray(radii) > 0
    mr = np.array(max_radii2D, dtype=F, copy=True)
    acc = np.array(xyz_gradient_accum, dtype=F, copy=True)
    dn = np.array(denom, dtype=F, copy=True)
    mr[vis] = np.maximum(mr[vis], np.asarray(radii)[vis].astype(F))
    g = np.asarray(viewspace_grad, dtype=F)[vis, :2]
    acc[vis] += np.sqrt((g * g).sum(-1, keepdims=True, dtype=F))
    dn[vis] += F(1)
    return mr, acc, dn
