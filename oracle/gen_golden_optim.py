"""ORACLE -- TEST INFRASTRUCTURE ONLY.

Generates tests/golden/optim_adam.npz: the reference's optimiser set-up (scene/gaussian_model.py:250-283: six
parameter groups, lr per group from arguments/__init__.py:78-88, torch.optim.Adam(l, lr=0.0, eps=1e-15)) stepped by
torch.optim.Adam ITSELF on the CPU for a few iterations with seeded gradients, and the densification statistics
statements of train.py:401-404 / scene/gaussian_model.py:764-766 executed verbatim on seeded inputs.

Run:  python oracle/gen_golden_optim.py
"""
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# arguments/__init__.py:78-88 (spatial_lr_scale = 1)
GROUPS = [("xyz", (3,), 0.00016), ("f_dc", (1, 3), 0.0025), ("f_rest", (15, 3), 0.0025 / 20.0), ("opacity", (1,), 0.05),
          ("scaling", (3,), 0.005), ("rotation", (4,), 0.001)]


def main(P=257, steps=4, seed=0):
    g = torch.Generator().manual_seed(seed)
    params = {n: torch.nn.Parameter(torch.randn((P,) + shp, generator=g)) for n, shp, _ in GROUPS}
    l = [{"params": [params[n]], "lr": lr, "name": n} for n, _, lr in GROUPS]
    opt = torch.optim.Adam(l, lr=0.0, eps=1e-15)   # scene/gaussian_model.py:283
    out = {f"p0_{n}": params[n].detach().numpy().copy() for n in params}
    for it in range(steps):
        for n, _, _ in GROUPS:
            grad = torch.randn(params[n].shape, generator=g) * (10.0 ** float(torch.randint(-6, 1, (1,), generator=g)))
            grad[torch.rand(P, generator=g) < 0.3] = 0.0     # Gaussians culled this iteration get zero gradients
            params[n].grad = grad
            out[f"g{it}_{n}"] = grad.numpy().copy()
        if it == 2:  # update_learning_rate moves the xyz lr (gaussian_model.py:290-295)
            opt.param_groups[0]["lr"] = 0.00012
        opt.step()
        for n in params:
            st = opt.state[params[n]]
            out[f"p{it + 1}_{n}"] = params[n].detach().numpy().copy()
            out[f"m{it + 1}_{n}"] = st["exp_avg"].numpy().copy()
            out[f"v{it + 1}_{n}"] = st["exp_avg_sq"].numpy().copy()
    out["xyz_lr"] = np.array([0.00016, 0.00016, 0.00012, 0.00012], dtype=np.float64)
    # densification statistics, statements as in the reference
    radii = torch.randint(-0, 40, (P,), generator=g, dtype=torch.int32)
    radii[torch.rand(P, generator=g) < 0.4] = 0
    vgrad = torch.randn(P, 3, generator=g) * 1e-3
    vgrad[:, 2] = 0
    max_radii2D = torch.randint(0, 30, (P,), generator=g).float()
    xyz_gradient_accum = torch.rand(P, 1, generator=g)
    denom = torch.randint(0, 5, (P, 1), generator=g).float()
    out.update(ds_radii=radii.numpy().copy(), ds_vgrad=vgrad.numpy().copy(), ds_max0=max_radii2D.numpy().copy(),
               ds_acc0=xyz_gradient_accum.numpy().copy(), ds_den0=denom.numpy().copy())
    visibility_filter = radii > 0                                   # gaussian_renderer/__init__.py:287
    max_radii2D[visibility_filter] = torch.max(max_radii2D[visibility_filter], radii[visibility_filter])  # train.py:403
    xyz_gradient_accum[visibility_filter] += torch.norm(vgrad[visibility_filter, :2], dim=-1, keepdim=True)  # :765
    denom[visibility_filter] += 1                                   # :766
    out.update(ds_max1=max_radii2D.numpy(), ds_acc1=xyz_gradient_accum.numpy(), ds_den1=denom.numpy())
    path = os.path.join(ROOT, "tests", "golden", "optim_adam.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
