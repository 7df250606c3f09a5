"""ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by mygauhuman_b200/).

CPU restatement (PyTorch, any dtype; fp32 to mirror the reference, fp64 as the
accuracy yardstick) of the LBS deformation part of the hot path.  Each function
cites the reference lines it follows (paths relative to /root/reference).
Pinned against tests/golden/lbs_*.npz, which oracle/gen_golden_lbs.py produced
by running the reference's own Python functions in the dev container.

The reference finds the nearest SMPL vertex with the third-party ``knn_cuda``
(KNN_CUDA 0.2, not vendored; scene/gaussian_model.py:23,87,775).  Its published
algorithm is exact brute-force k-NN on squared Euclidean distance; here it is
``nearest_vertex`` (argmin over all vertices, lowest index on ties).
"""
import torch

from mygauhuman_b200.synthetic import SMPL_PARENTS


def rodrigues(rvec):
    """scene/gaussian_model.py:894-912 batch_rodrigues_torch == :982-1013 batch_rodrigues
    == smplx/lbs.py:299-333: angle = |r + 1e-8|, k = r/angle, R = I + sin K + (1-cos) K^2."""
    n = rvec.shape[0]
    angle = torch.linalg.norm(rvec + 1e-8, dim=1, keepdim=True)
    k = rvec / angle
    kx, ky, kz = k[:, 0], k[:, 1], k[:, 2]
    zero = torch.zeros_like(kx)
    K = torch.stack([zero, -kz, ky, kz, zero, -kx, -ky, kx, zero], dim=1).reshape(n, 3, 3)
    s = torch.sin(angle)[:, :, None]
    c = torch.cos(angle)[:, :, None]
    eye = torch.eye(3, dtype=rvec.dtype, device=rvec.device)[None]
    return eye + s * K + (1.0 - c) * (K @ K)


def rigid_chain(rot, joints, parents=SMPL_PARENTS):
    """scene/gaussian_model.py:914-944 get_rigid_transformation_torch (== smplx/lbs.py:349-405):
    G_0 = [R_0 | j_0], G_i = G_parent(i) [R_i | j_i - j_parent(i)], A_i = G_i with translation - G_i^{3x3} j_i."""
    J = joints.shape[0]
    par = torch.as_tensor(list(parents), dtype=torch.long, device=joints.device)
    rel = joints.clone()
    rel[1:] = rel[1:] - joints[par[1:]]                                    # :923-924
    top = torch.cat([rot, rel[:, :, None]], dim=-1)                        # :927 [J,3,4]
    bottom = torch.zeros(J, 1, 4, dtype=rot.dtype, device=rot.device)
    bottom[..., 3] = 1
    mats = torch.cat([top, bottom], dim=-2)                                # :928-930 [J,4,4]
    G = [mats[0]]
    for i in range(1, J):                                                  # :933-936 serial chain
        G.append(G[parents[i]] @ mats[i])
    A = torch.stack(G)
    A = A.clone()
    A[:, :3, 3] = A[:, :3, 3] - torch.einsum("jab,jb->ja", A[:, :3, :3], joints)
    # the reference subtracts sum_k G[:, r, k] * [j,0][k] from column 3 of every row, incl. row 3 (which is 0 0 0 1 . [j,0] = 0)
    return A


def rest_joints(smpl, betas):
    """scene/gaussian_model.py:951-970: v_shaped = v_template + shapedirs[..., :nb] . betas; joints = J_regressor v_shaped."""
    nb = betas.shape[-1]
    v_shaped = smpl["v_template"] + (smpl["shapedirs"][..., :nb] * betas.reshape(1, 1, nb)).sum(-1)
    return smpl["J_regressor"] @ v_shaped


def transform_params(smpl, params, rot=None, correct_Rs=None):
    """scene/gaussian_model.py:947-980 get_transform_params_torch for batch size 1 -> (A[24,4,4], rot[24,3,3], joints)."""
    if rot is None:
        rot = rodrigues(params["poses"].reshape(-1, 3))
        if correct_Rs is not None:
            rot = torch.cat([rot[0:1], rot[1:] @ correct_Rs.reshape(-1, 3, 3)], 0)
    joints = rest_joints(smpl, params["shapes"].reshape(-1))
    return rigid_chain(rot, joints), rot, joints


def nearest_vertex(t_vertices, query):
    """KNN(k=1, transpose_mode=True)(ref=t_vertices, query) -> index of the closest vertex (gaussian_model.py:775)."""
    out = torch.empty(query.shape[0], dtype=torch.long, device=query.device)
    tv = t_vertices.double()
    q = query.double()
    for s in range(0, q.shape[0], 4096):
        d = ((q[s:s + 4096, None, :] - tv[None]) ** 2).sum(-1)
        out[s:s + 4096] = d.argmin(1)
    return out


def pose_feature(rot):
    """(rot_mats[1:] - I) flattened, gaussian_model.py:811,838."""
    return (rot[1:] - torch.eye(3, dtype=rot.dtype, device=rot.device)[None]).reshape(-1)


def coarse_deform_c2source(smpl, query_pts, params, t_params, t_vertices, lbs_weights=None, correct_Rs=None,
                           return_transl=False, normals=None, vert_ids=None):
    """scene/gaussian_model.py:768-872 for batch size 1.
    query_pts [P,3], normals [P,3], lbs_weights [P,24] or None, correct_Rs [23,3,3] or None.
    Returns (smpl_pts[P,3], world_pts[P,3], bweights[P,24], transforms[P,3,3], translation[P,3]|None,
    world_normals[P,3], vert_ids[P])."""
    dt = query_pts.dtype
    P = query_pts.shape[0]
    V = t_vertices.shape[0]
    if vert_ids is None:
        vert_ids = nearest_vertex(t_vertices, query_pts)
    bw = smpl["weights"][vert_ids]                                       # :777-779
    if lbs_weights is not None:
        bw = torch.softmax(torch.log(bw + 1e-9) + lbs_weights, dim=-1)   # :780-781
    # big pose -> T pose (:784-799)
    A_big, rot_big, _ = transform_params(smpl, t_params)
    Ab = (bw @ A_big.reshape(24, 16)).reshape(P, 4, 4)
    x = query_pts - Ab[:, :3, 3]
    Rinv = torch.linalg.inv(Ab[:, :3, :3])
    x = (Rinv @ x[:, :, None])[:, :, 0]
    n = (Rinv @ normals[:, :, None])[:, :, 0] if normals is not None else None
    transforms = Rinv
    translation = None
    if return_transl:
        translation = (Rinv @ (-Ab[:, :3, 3])[:, :, None])[:, :, 0]
    # blend-shape offsets gathered at the nearest vertex (:802-844)
    posedirs = smpl["posedirs"].reshape(V * 3, -1)
    off_big = (posedirs @ pose_feature(rot_big)).reshape(V, 3)[vert_ids]
    x = x - off_big
    nb = params["shapes"].shape[-1]
    off_shape = (smpl["shapedirs"][..., :nb] @ params["shapes"].reshape(nb))[vert_ids]
    x = x + off_shape
    rot = rodrigues(params["poses"].reshape(-1, 3))
    if correct_Rs is not None:
        rot = torch.cat([rot[0:1], rot[1:] @ correct_Rs.reshape(-1, 3, 3)], 0)
    off_pose = (posedirs @ pose_feature(rot)).reshape(V, 3)[vert_ids]
    x = x + off_pose
    if return_transl:
        translation = translation - off_big + off_shape + off_pose
    # T pose -> target pose (:846-860)
    A_t, _, _ = transform_params(smpl, params, rot=rot)
    At = (bw @ A_t.reshape(24, 16)).reshape(P, 4, 4)
    smpl_pts = (At[:, :3, :3] @ x[:, :, None])[:, :, 0] + At[:, :3, 3]
    smpl_n = (At[:, :3, :3] @ n[:, :, None])[:, :, 0] if n is not None else None
    transforms = At[:, :3, :3] @ transforms
    if return_transl:
        translation = (At[:, :3, :3] @ translation[:, :, None])[:, :, 0] + At[:, :3, 3]
    # smpl space -> world (:862-870)
    R = params["R"].to(dt)
    Th = params["Th"].to(dt).reshape(1, 3)
    R_inv = torch.linalg.inv(R)
    world_pts = smpl_pts @ R_inv + Th
    world_n = smpl_n @ R_inv if smpl_n is not None else None
    transforms = R[None] @ transforms
    if return_transl:
        translation = translation @ R_inv + Th
    return smpl_pts, world_pts, bw, transforms, translation, world_n, vert_ids


def quat_to_rot(q):
    """utils/general_utils.py:76-99 build_rotation (normalises q)."""
    q = q / torch.sqrt((q * q).sum(1, keepdim=True))
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    return torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).reshape(-1, 3, 3)


def covariance(scaling, scaling_modifier, rotation, transform=None):
    """scene/gaussian_model.py:35-42 + utils/general_utils.py:60-117:
    L = R(q) diag(mod*s); Sigma = L L^T; Sigma' = T Sigma T^T; packed [xx,xy,xz,yy,yz,zz]."""
    L = quat_to_rot(rotation) * (scaling_modifier * scaling)[:, None, :]
    S = L @ L.transpose(1, 2)
    if transform is not None:
        S = transform @ S @ transform.transpose(1, 2)
    return torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1)


SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
SH_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
         1.445305721320277, -0.5900435899266435]


def eval_sh(deg, sh, dirs):
    """utils/sh_utils.py:135-190 (deg <= 3).  sh [P,3,M] (channel-major), dirs [P,3] unit."""
    res = SH_C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[:, 0:1], dirs[:, 1:2], dirs[:, 2:3]
        res = res - SH_C1 * y * sh[..., 1] + SH_C1 * z * sh[..., 2] - SH_C1 * x * sh[..., 3]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            res = (res + SH_C2[0] * xy * sh[..., 4] + SH_C2[1] * yz * sh[..., 5] +
                   SH_C2[2] * (2.0 * zz - xx - yy) * sh[..., 6] + SH_C2[3] * xz * sh[..., 7] +
                   SH_C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                res = (res + SH_C3[0] * y * (3 * xx - yy) * sh[..., 9] + SH_C3[1] * xy * z * sh[..., 10] +
                       SH_C3[2] * y * (4 * zz - xx - yy) * sh[..., 11] +
                       SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12] +
                       SH_C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + SH_C3[5] * z * (xx - yy) * sh[..., 14] +
                       SH_C3[6] * x * (xx - 3 * yy) * sh[..., 15])
    return res


def view_colors(deg, features, means3D, campos):
    """gaussian_renderer/__init__.py:124-125,191-195: dirs from the deformed means; clamp_min(eval_sh + 0.5, 0).
    features [P,M,3] (the reference's get_features layout)."""
    d = means3D - campos.reshape(1, 3)
    d = d / d.norm(dim=1, keepdim=True)
    return torch.clamp_min(eval_sh(deg, features.transpose(1, 2), d) + 0.5, 0.0)


def smplx_lbs(betas, pose, v_template, shapedirs, posedirs, J_regressor, parents, lbs_weights):
    """smplx/lbs.py:156-252 lbs(..., pose2rot=True) for batch size 1 (the 4-value variant of this fork):
    returns (verts[V,3], J_transformed[24,3], A[24,4,4], T[V,4,4])."""
    v_shaped = v_template + (shapedirs * betas.reshape(1, 1, -1)).sum(-1)           # :197 blend_shapes
    J = J_regressor @ v_shaped                                                       # :200 vertices2joints
    rot = rodrigues(pose.reshape(-1, 3))                                             # :205-206
    pf = pose_feature(rot)                                                           # :208
    v_posed = v_shaped + (pf @ posedirs).reshape(-1, 3)                              # :210-212 (posedirs [207, V*3])
    A = rigid_chain(rot, J, parents)                                                 # :222
    G = A.clone()
    G[:, :3, 3] = G[:, :3, 3] + torch.einsum("jab,jb->ja", A[:, :3, :3], J)          # posed joints = chain translation
    T = (lbs_weights @ A.reshape(24, 16)).reshape(-1, 4, 4)                          # :226-230
    verts = (T[:, :3, :3] @ v_posed[:, :, None])[:, :, 0] + T[:, :3, 3]              # :232-238
    return verts, G[:, :3, 3], A, T
