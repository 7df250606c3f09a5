"""Seeded synthetic inputs of the named shapes: an SMPL-topology
body model (6,890 vertices, 24 joints, random blend shapes / skinning weights),
Gaussians initialised on it, and cameras built with the reference's own
conventions (reference utils/graphics_utils.py:38-50,82-105 and
scene/cameras.py:58-67).  No datasets, no SMPL checkpoints.

Everything is generated on the CPU with a seeded generator, so CPU baselines
and GPU runs see bit-identical inputs.
"""
import math

import numpy as np
import torch

SMPL_PARENTS = [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21]
NUM_VERTS = 6890
NUM_JOINTS = 24


def _rodrigues_np(r):
    th = np.linalg.norm(r)
    if th < 1e-12:
        return np.eye(3)
    k = r / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + math.sin(th) * K + (1 - math.cos(th)) * K @ K


def make_smpl_model(seed=0, num_verts=NUM_VERTS):
    """A random but well-conditioned body model with SMPL's tensor shapes
    (reference scene/gaussian_model.py:880-892 SMPL_to_tensor keys)."""
    g = np.random.default_rng(seed)
    # rest joints of a 1.7 m humanoid (y up), children placed along plausible bone directions
    J = np.zeros((NUM_JOINTS, 3))
    offs = {1: (0.09, -0.09, 0), 2: (-0.09, -0.09, 0), 3: (0, 0.11, 0), 4: (0, -0.40, 0), 5: (0, -0.40, 0),
            6: (0, 0.13, 0), 7: (0, -0.42, 0), 8: (0, -0.42, 0), 9: (0, 0.10, 0), 10: (0, -0.06, 0.12),
            11: (0, -0.06, 0.12), 12: (0, 0.17, 0), 13: (0.10, 0.08, 0), 14: (-0.10, 0.08, 0), 15: (0, 0.12, 0),
            16: (0.12, 0.02, 0), 17: (-0.12, 0.02, 0), 18: (0.26, 0, 0), 19: (-0.26, 0, 0), 20: (0.25, 0, 0),
            21: (-0.25, 0, 0), 22: (0.09, 0, 0), 23: (-0.09, 0, 0)}
    J[0] = (0, 0.95, 0)
    for j in range(1, NUM_JOINTS):
        J[j] = J[SMPL_PARENTS[j]] + np.array(offs[j])
    # vertices: capsules around bones
    per = num_verts // NUM_JOINTS
    counts = [per] * NUM_JOINTS
    counts[0] += num_verts - per * NUM_JOINTS
    verts, owner = [], []
    for j in range(NUM_JOINTS):
        a = J[SMPL_PARENTS[j]] if j else J[0] + np.array([0, -0.08, 0])
        b = J[j]
        t = g.random((counts[j], 1))
        c = a * (1 - t) + b * t
        d = g.normal(size=(counts[j], 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        rad = 0.11 if j in (0, 3, 6, 9) else (0.09 if j in (12, 15) else 0.05)
        verts.append(c + d * rad * g.uniform(0.7, 1.0, size=(counts[j], 1)))
        owner += [j] * counts[j]
    v_template = np.concatenate(verts).astype(np.float32)
    owner = np.array(owner)
    # skinning weights: <= 4 non-zeros per row (owner, parent, random others), Dirichlet
    weights = np.zeros((num_verts, NUM_JOINTS), np.float32)
    for i in range(num_verts):
        js = {owner[i], max(SMPL_PARENTS[owner[i]], 0)}
        while len(js) < 4:
            js.add(int(g.integers(0, NUM_JOINTS)))
        js = sorted(js)
        w = g.dirichlet([4.0 if j == owner[i] else 0.5 for j in js])
        weights[i, js] = w
    weights /= weights.sum(1, keepdims=True)
    # sparse row-normalised joint regressor: each joint regresses from vertices it owns
    Jreg = np.zeros((NUM_JOINTS, num_verts), np.float32)
    for j in range(NUM_JOINTS):
        idx = np.nonzero(owner == j)[0]
        pick = g.choice(idx, size=min(32, len(idx)), replace=False)
        Jreg[j, pick] = g.random(len(pick)).astype(np.float32) + 0.1
        Jreg[j] /= Jreg[j].sum()
    shapedirs = (g.normal(size=(num_verts, 3, 10)) * 0.01).astype(np.float32)
    posedirs = (g.normal(size=(num_verts * 3, 207)) * 0.001).astype(np.float32)
    return {
        "v_template": torch.from_numpy(v_template),
        "shapedirs": torch.from_numpy(shapedirs),
        "posedirs": torch.from_numpy(posedirs),
        "J_regressor": torch.from_numpy(Jreg),
        "weights": torch.from_numpy(weights),
        "kintree_table": torch.tensor([SMPL_PARENTS, list(range(NUM_JOINTS))], dtype=torch.long),
    }


def big_pose_params():
    """reference scene/dataset_readers.py:586-594: arms and legs spread."""
    poses = np.zeros((1, 72), np.float32)
    poses[0, 5] = 45 / 180 * math.pi
    poses[0, 8] = -45 / 180 * math.pi
    poses[0, 23] = -30 / 180 * math.pi
    poses[0, 26] = 30 / 180 * math.pi
    return {"poses": torch.from_numpy(poses), "shapes": torch.zeros(1, 10), "R": torch.eye(3), "Th": torch.zeros(1, 3)}


def make_pose_params(seed=0):
    g = np.random.default_rng(seed + 1000)
    poses = (g.normal(size=(1, 72)) * 0.2).astype(np.float32)
    shapes = g.normal(size=(1, 10)).astype(np.float32)
    R = _rodrigues_np(g.normal(size=3) * 0.3).astype(np.float32)
    Th = g.uniform(-0.5, 0.5, size=(1, 3)).astype(np.float32)
    return {"poses": torch.from_numpy(poses), "shapes": torch.from_numpy(shapes), "R": torch.from_numpy(R),
            "Th": torch.from_numpy(Th)}


def make_gaussians(P, base_points, seed=0, sh_degree=3, jitter=0.01, scale=0.01):
    """P Gaussians scattered around base_points[V,3]."""
    g = np.random.default_rng(seed + 2000)
    V = base_points.shape[0]
    vid = np.arange(P) % V
    xyz = base_points[vid] + g.normal(size=(P, 3)) * jitter
    M = (sh_degree + 1) ** 2
    log_scale = np.log(scale) + g.normal(size=(P, 3)) * 0.3
    rot = g.normal(size=(P, 4))
    rot /= np.linalg.norm(rot, axis=1, keepdims=True)
    opacity_logit = g.normal(size=(P, 1)) * 1.5
    sh = np.concatenate([g.uniform(-1, 1, size=(P, 1, 3)), g.normal(size=(P, M - 1, 3)) * 0.05], 1)
    normals = g.normal(size=(P, 3))
    normals /= np.linalg.norm(normals, axis=1, keepdims=True)
    f = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
    return {"xyz": f(xyz), "log_scale": f(log_scale), "rotation": f(rot), "opacity_logit": f(opacity_logit),
            "features": f(sh), "normals": f(normals)}


# ---- cameras (reference conventions) ----------------------------------------

def world2view(R, t):
    """getWorld2View2 with translate=0, scale=1 (reference utils/graphics_utils.py:38-50)."""
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.transpose()
    Rt[:3, 3] = t
    Rt[3, 3] = 1.0
    return np.float32(np.linalg.inv(np.linalg.inv(Rt)))


def projection_refine(K, H, W, znear=0.001, zfar=1000.0):
    """getProjectionMatrix_refine (reference utils/graphics_utils.py:82-105)."""
    P = torch.zeros(4, 4, dtype=torch.float32)
    P[0, 0] = 2 * K[0, 0] / W
    P[0, 1] = 2 * K[0, 1] / W
    P[0, 2] = -1 + 2 * (K[0, 2] / W)
    P[1, 1] = 2 * K[1, 1] / H
    P[1, 2] = -1 + 2 * (K[1, 2] / H)
    P[2, 2] = (zfar + znear) / (zfar - znear)
    P[2, 3] = -1 * 2 * zfar * znear / (zfar - znear)
    P[3, 2] = 1.0
    return P


def make_camera(W, H, angle=0.0, radius=3.0, target=(0.0, 0.9, 0.0), elev=0.1, seed=0, focal=None):
    """Ring camera looking at `target`; intrinsics f=537*(W/512), principal point centre +- 2 px."""
    g = np.random.default_rng(seed + 3000)
    target = np.asarray(target, dtype=np.float64)
    eye = target + radius * np.array([math.sin(angle) * math.cos(elev), math.sin(elev), math.cos(angle) * math.cos(elev)])
    fwd = target - eye
    fwd /= np.linalg.norm(fwd)
    up = np.array([0.0, -1.0, 0.0])  # image y points down
    right = np.cross(up, fwd)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    c2w_R = np.stack([right, down, fwd], 1)  # columns = camera axes in world
    w2c_R = c2w_R.T
    T = -w2c_R @ eye
    R = w2c_R.T  # the reference stores R transposed (scene/cameras.py + getWorld2View2)
    f = 537.0 * (W / 512.0) if focal is None else focal
    K = torch.tensor([[f, 0.0, W / 2 + g.uniform(-2, 2)], [0.0, f, H / 2 + g.uniform(-2, 2)], [0.0, 0.0, 1.0]],
                     dtype=torch.float32)
    wvt = torch.tensor(world2view(R, T)).transpose(0, 1).contiguous()
    proj = projection_refine(K, H, W).transpose(0, 1)
    full = wvt.unsqueeze(0).bmm(proj.unsqueeze(0)).squeeze(0).contiguous()
    center = wvt.inverse()[3, :3].contiguous()
    fovx = 2 * math.atan(W / (2 * f))
    fovy = 2 * math.atan(H / (2 * f))
    return {"W": W, "H": H, "K": K, "world_view_transform": wvt, "full_proj_transform": full,
            "camera_center": center, "FoVx": fovx, "FoVy": fovy, "tanfovx": math.tan(fovx * 0.5),
            "tanfovy": math.tan(fovy * 0.5)}


def make_scene_cloud(P, seed=0, box=10.0, scale=0.02, sh_degree=3):
    """Config 5: uniform cloud in a box in front of a camera at the origin looking down +z."""
    g = np.random.default_rng(seed + 4000)
    pts = g.uniform(-box / 2, box / 2, size=(P, 3))
    pts[:, 2] += box / 2 + 1.0
    return make_gaussians(P, pts.astype(np.float32), seed=seed, sh_degree=sh_degree, jitter=0.0, scale=scale)


def big_pose_vertices(smpl, big):
    """SMPL surface skinned into the big pose: the `big_pose_world_vertex` every reference camera carries
    (reference scene/dataset_readers.py:586-600).  Host-side numpy (fp64) -- input generation, not the hot path."""
    poses = big["poses"].numpy().reshape(24, 3).astype(np.float64)
    betas = big["shapes"].numpy().reshape(-1).astype(np.float64)
    vt = smpl["v_template"].numpy().astype(np.float64)
    v_shaped = vt + smpl["shapedirs"].numpy().astype(np.float64) @ betas
    joints = smpl["J_regressor"].numpy().astype(np.float64) @ v_shaped
    R = np.stack([_rodrigues_np(p) for p in poses])
    pf = (R[1:] - np.eye(3)[None]).reshape(-1)
    v_posed = v_shaped + (smpl["posedirs"].numpy().astype(np.float64).reshape(-1, 207) @ pf).reshape(-1, 3)
    G = np.zeros((24, 4, 4))
    for i in range(24):
        M = np.eye(4)
        M[:3, :3] = R[i]
        M[:3, 3] = joints[i] - (joints[SMPL_PARENTS[i]] if i else 0.0)
        G[i] = M if i == 0 else G[SMPL_PARENTS[i]] @ M
    A = G.copy()
    A[:, :3, 3] -= np.einsum("jab,jb->ja", G[:, :3, :3], joints)
    T = np.einsum("vj,jab->vab", smpl["weights"].numpy().astype(np.float64), A)
    verts = np.einsum("vab,vb->va", T[:, :3, :3], v_posed) + T[:, :3, 3]
    return torch.from_numpy(verts.astype(np.float32))
