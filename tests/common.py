"""Shared builders for the parity tests: seeded raster cases as numpy dicts."""
import math

import numpy as np

from mygauhuman_b200 import synthetic as S

_SMPL = None


def smpl():
    global _SMPL
    if _SMPL is None:
        _SMPL = S.make_smpl_model(0)
    return _SMPL


def raster_case(P, W, H, seed=0, mode="precomp", degree=3, angle=0.3, radius=3.0, scale=0.01, bg=(0.1, 0.2, 0.3),
                cloud=None):
    """mode: 'precomp' (colors_precomp + cov3D_precomp, what render() uses),
    'sh_scale_rot' (baking.py path), 'sh_cov', 'color_scale_rot'."""
    base = smpl()["v_template"].numpy() if cloud is None else cloud
    g = S.make_gaussians(P, base, seed=seed, sh_degree=3, scale=scale)
    cam = S.make_camera(W, H, angle=angle, radius=radius, seed=seed)
    xyz = g["xyz"].numpy()
    scales = np.exp(g["log_scale"].numpy()).astype(np.float32)
    rot = g["rotation"].numpy()
    opacity = (1.0 / (1.0 + np.exp(-g["opacity_logit"].numpy()))).astype(np.float32)
    sh = g["features"].numpy()
    rng = np.random.default_rng(seed + 77)
    case = dict(P=P, W=W, H=H, degree=degree, bg=np.asarray(bg, np.float32), means3D=xyz, opacity=opacity,
                view=cam["world_view_transform"].numpy().copy(), proj=cam["full_proj_transform"].numpy().copy(),
                campos=cam["camera_center"].numpy().copy(), tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                scale_modifier=1.0, colors=None, cov3D=None, scales=None, rotations=None, sh=None)
    if mode in ("precomp", "color_scale_rot"):
        case["colors"] = rng.uniform(0, 1, size=(P, 3)).astype(np.float32)
    else:
        case["sh"] = sh
    if mode in ("precomp", "sh_cov"):
        # cov = R S S^T R^T with normalised q, as reference utils/general_utils.py:79-117 builds it
        q = rot / np.linalg.norm(rot, axis=1, keepdims=True)
        r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                      2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                      2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).reshape(P, 3, 3)
        L = R * scales[:, None, :]
        cov = L @ L.transpose(0, 2, 1)
        case["cov3D"] = np.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]],
                                 1).astype(np.float32)
    else:
        case["scales"] = scales
        case["rotations"] = rot.astype(np.float32)
    case["dL_dcolor"] = rng.normal(size=(3, H, W)).astype(np.float32)
    case["dL_ddepth"] = rng.normal(size=(1, H, W)).astype(np.float32)
    case["dL_dalpha"] = rng.normal(size=(1, H, W)).astype(np.float32)
    return case


def oracle_forward(case):
    import oracle as O
    c = case
    return O.raster_forward(c["bg"], c["means3D"], c["colors"], c["opacity"], c["scales"], c["rotations"],
                            c["scale_modifier"], c["cov3D"], c["view"], c["proj"], c["tanfovx"], c["tanfovy"], c["H"],
                            c["W"], c["sh"], c["degree"], c["campos"])


def oracle_backward(case, st):
    import oracle as O
    c = case
    return O.raster_backward(st, c["bg"], c["means3D"], c["colors"], c["scales"], c["rotations"], c["scale_modifier"],
                             c["cov3D"], c["view"], c["proj"], c["tanfovx"], c["tanfovy"], c["dL_dcolor"],
                             c["dL_ddepth"], c["dL_dalpha"], c["sh"], c["degree"], c["campos"])


def to_torch(case, device):
    import torch
    t = {}
    for k, v in case.items():
        if isinstance(v, np.ndarray):
            t[k] = torch.from_numpy(v).to(device)
        else:
            t[k] = v
    return t


def empty():
    import torch
    return torch.Tensor([])


def ours_forward(tc):
    """Call the drop-in _C.rasterize_gaussians with the reference's 19 positional args."""
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    e = empty
    return _C.rasterize_gaussians(
        tc["bg"], tc["means3D"], tc["colors"] if tc["colors"] is not None else e(), tc["opacity"],
        tc["scales"] if tc["scales"] is not None else e(), tc["rotations"] if tc["rotations"] is not None else e(),
        tc["scale_modifier"], tc["cov3D"] if tc["cov3D"] is not None else e(), tc["view"], tc["proj"], tc["tanfovx"],
        tc["tanfovy"], tc["H"], tc["W"], tc["sh"] if tc["sh"] is not None else e(), tc["degree"], tc["campos"], False,
        False)


def ours_backward(tc, fw):
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    e = empty
    n, color, depth, alpha, radii, geom, binning, img = fw
    return _C.rasterize_gaussians_backward(
        tc["bg"], tc["means3D"], radii, tc["colors"] if tc["colors"] is not None else e(),
        tc["scales"] if tc["scales"] is not None else e(), tc["rotations"] if tc["rotations"] is not None else e(),
        tc["scale_modifier"], tc["cov3D"] if tc["cov3D"] is not None else e(), tc["view"], tc["proj"], tc["tanfovx"],
        tc["tanfovy"], tc["dL_dcolor"], tc["dL_ddepth"], tc["dL_dalpha"], tc["sh"] if tc["sh"] is not None else e(),
        tc["degree"], tc["campos"], geom, n, binning, img, alpha, False)


def rel_err(a, b):
    """max |a-b| / max(|b|_inf, tiny): the 'relative to the tensor scale' error used for gradients."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)) if a.size else 0.0


def grad_close(name, a, b, rel=1e-4):
    """Per-element gradient criterion of the parity tests (north_star: parameter gradients within 1e-4 rel):
    |a - b| <= rel * |b| + 0.1 * rel * max|b| for EVERY element -- the second term is the floor below which the order
    of fp32 accumulation (atomics on both sides) decides."""
    a = np.asarray(a, np.float64).reshape(-1); b = np.asarray(b, np.float64).reshape(-1)
    assert a.shape == b.shape, name
    scale = max(np.abs(b).max(), 1e-30) if b.size else 1.0
    err = np.abs(a - b)
    bad = err > rel * np.abs(b) + rel * 1e-1 * scale
    assert not bad.any(), f"{name}: {bad.sum()} of {a.size} elements off, max err {err.max():.3e} (scale {scale:.3e})"
