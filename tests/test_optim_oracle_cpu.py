"""The optimiser oracle (oracle/optim_oracle.py) pinned against golden vectors produced by torch.optim.Adam itself in
the reference's configuration and by the reference's densification-statistics statements (oracle/gen_golden_optim.py)."""
import os

import numpy as np
import pytest

from oracle import optim_oracle as OO

GOLD = os.path.join(os.path.dirname(__file__), "golden", "optim_adam.npz")
LR = {"xyz": None, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}


def close(a, b, scale):
    """fp32 rounding only: |a-b| <= 2e-7 * max(|b|, scale) elementwise; `scale` = magnitude of the operands the value
    was summed from (a sum can cancel below its operands, its rounding error cannot)."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return bool(np.all(np.abs(a - b) <= 2e-7 * np.maximum(np.abs(b), scale)))


def test_adam_oracle_matches_torch_adam_goldens():
    z = np.load(GOLD)
    for name, lr in LR.items():
        p = z[f"p0_{name}"]
        m, v = np.zeros_like(p), np.zeros_like(p)
        for it in range(4):
            lr_it = z["xyz_lr"][it] if name == "xyz" else lr
            g, m0, v0, p0 = z[f"g{it}_{name}"], m, v, p
            p, m, v = OO.adam_step(p, g, m, v, it + 1, lr_it)
            # moments: fp32 rounding (FMA contraction inside torch's vectorised CPU kernels differs from numpy's)
            assert close(m, z[f"m{it + 1}_{name}"], np.maximum(np.abs(m0), np.abs(g) * 0.1)), (name, it)
            assert close(v, z[f"v{it + 1}_{name}"], np.abs(v0)), (name, it)
            assert close(p, z[f"p{it + 1}_{name}"], np.maximum(np.abs(p0), lr_it)), (name, it)
            # chain from the golden state so that rounding differences cannot accumulate into a false pass/fail
            p, m, v = z[f"p{it + 1}_{name}"], z[f"m{it + 1}_{name}"], z[f"v{it + 1}_{name}"]


def test_adam_oracle_zero_gradient_rows_do_not_move_at_first_step():
    z = np.load(GOLD)
    g = z["g0_xyz"]
    p, m, v = OO.adam_step(z["p0_xyz"], g, np.zeros_like(g), np.zeros_like(g), 1, 0.00016)
    still = (g == 0)
    assert still.any() and np.array_equal(p[still], z["p0_xyz"][still])  # 0 / (0 + 1e-15) = 0


def test_densify_stats_oracle_matches_reference_statements():
    z = np.load(GOLD)
    mr, acc, dn = OO.densify_stats(z["ds_radii"], z["ds_vgrad"], z["ds_max0"], z["ds_acc0"], z["ds_den0"])
    assert np.array_equal(mr, z["ds_max1"])
    assert np.array_equal(dn, z["ds_den1"])
    np.testing.assert_allclose(acc, z["ds_acc1"], rtol=2e-7)
    hidden = z["ds_radii"] <= 0
    assert hidden.any() and np.array_equal(acc[hidden], z["ds_acc0"][hidden])
