"""View-batched forward rasterisation and the cube-map pass of the reference's occlusion baking.

Reference: baking.py::bake_set (:136-309).  For every occupied cell of a 10^3 grid over the Gaussians it renders the
six 90-degree 32x32 faces of a cube map centred on the cell with ONE ``_C.rasterize_gaussians`` call per face, on
boolean-mask gathers of every Gaussian tensor that leave out the cell's own Gaussians (:237-246), and keeps only
the alpha map (:283).  ``rasterize_views`` renders any number of views of the same Gaussians through
``ghr_forward_views`` (six kernel launches per batch of views; the per-cell mask is a label compare inside the
projection kernel), and ``bake_opacity_cubemaps`` is the bake_set loop on top of it.

The cube-map -> lat-long resampling (``dr.texture(..., filter_mode="nearest", boundary_mode="cube")``, baking.py
:286-297) is a plain gather (direction -> face + texel); ``cubemap_sample_nearest`` is that call for any number of cube
maps, ``texture`` has nvdiffrast's call signature (so the reference's unmodified bake_set runs on it), and
``bake_occlusion`` is all of bake_set: cube faces, lookup, per-Gaussian ``1 - alpha`` and the normal-hemisphere mask
(:231,305-307) in three kernels (``ghr_forward_views`` batch, ``ghr_bake_occlusion``).

No CPU fallback: CUDA tensors only.
"""
import math

import torch

from . import _lib
from .diff_gaussian_rasterization._C import _params, _stream

CAMERA_FLOATS = 36  # viewmatrix 16 | projmatrix 16 | campos 3 | pad


def projection_matrix(znear, zfar, fovX, fovY):
    """utils/graphics_utils.py:51-70 getProjectionMatrix."""
    tanHalfFovY, tanHalfFovX = math.tan(fovY / 2), math.tan(fovX / 2)
    top, right = tanHalfFovY * znear, tanHalfFovX * znear
    bottom, left = -top, -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -2 * (zfar * znear) / (zfar - znear)
    return P


def cube_face_rotations():
    """The six camera-to-world rotations of baking.py:146-195 (looking along -x, +x, -y, +y, -z, +z), [6,4,4]."""
    return torch.tensor([
        [[0.0, 0.0, 1.0, 0.0], [0.0, -1.0, 0.0, 0.0], [-1.0, 0.0, 0.0, 0.0], [0.0, 0.0, 0.0, 1.0]],
        [[0.0, 0.0, -1.0, 0.0], [0.0, -1.0, 0.0, 0.0], [1.0, 0.0, 0.0, 0.0], [0.0, 0.0, 0.0, 1.0]],
        [[1.0, 0.0, 0.0, 0.0], [0.0, 0.0, 1.0, 0.0], [0.0, 1.0, 0.0, 0.0], [0.0, 0.0, 0.0, 1.0]],
        [[1.0, 0.0, 0.0, 0.0], [0.0, 0.0, -1.0, 0.0], [0.0, -1.0, 0.0, 0.0], [0.0, 0.0, 0.0, 1.0]],
        [[1.0, 0.0, 0.0, 0.0], [0.0, -1.0, 0.0, 0.0], [0.0, 0.0, 1.0, 0.0], [0.0, 0.0, 0.0, 1.0]],
        [[-1.0, 0.0, 0.0, 0.0], [0.0, -1.0, 0.0, 0.0], [0.0, 0.0, -1.0, 0.0], [0.0, 0.0, 0.0, 1.0]],
    ])


def pack_cameras(world_view_transform, full_proj_transform, camera_center):
    """[n,4,4], [n,4,4], [n,3] (the reference's transposed matrices, scene/cameras.py:63-67) -> [n,36]."""
    n = world_view_transform.shape[0]
    out = torch.zeros(n, CAMERA_FLOATS, dtype=torch.float32, device=world_view_transform.device)
    out[:, 0:16] = world_view_transform.reshape(n, 16)
    out[:, 16:32] = full_proj_transform.reshape(n, 16)
    out[:, 32:35] = camera_center.reshape(n, 3)
    return out


def cube_cameras(positions, znear=0.01, zfar=5.0):
    """Six cube-face cameras per position, as baking.py:196-202,248-258 builds them: c2w = rotation with the position
    as translation, w2c = c2w^-1, world_view_transform = w2c^T, full_proj = world_view_transform @ projection^T,
    camera_center = position.  positions [C,3] -> packed cameras [C*6, 36] (face index fastest)."""
    dev = positions.device
    C = positions.shape[0]
    rot = cube_face_rotations().to(dev)                       # [6,4,4]
    c2w = rot[None].repeat(C, 1, 1, 1)                        # [C,6,4,4]
    c2w[:, :, :3, 3] = positions[:, None, :].float()
    w2c = torch.inverse(c2w)
    wvt = w2c.transpose(-1, -2).contiguous()                  # getWorld2View(R, T)^T
    proj = projection_matrix(znear, zfar, math.pi * 0.5, math.pi * 0.5).transpose(0, 1).to(dev)
    full = wvt @ proj
    center = torch.inverse(wvt)[:, :, 3, :3]                  # world_view_transform.inverse()[3, :3]
    return pack_cameras(wvt.reshape(C * 6, 4, 4), full.reshape(C * 6, 4, 4), center.reshape(C * 6, 3))


def pc_to_grid(pc, res):
    """baking.py:104-134: occupied cells of a res^3 grid over the points.  Returns (grid_centers [C,3], grid_sizes [3],
    pc_grid_indices [P] = cell of every point, unique_indices [C,3])."""
    min_coords, max_coords = pc.min(dim=0)[0], pc.max(dim=0)[0]
    grid_sizes = (max_coords - min_coords) / res
    idx = torch.floor((pc - min_coords) / grid_sizes).long().clamp_(0, res - 1)
    unique_indices, inverse = torch.unique(idx, return_inverse=True, dim=0)
    centers = min_coords[None, :] + unique_indices * grid_sizes[None, :] + grid_sizes[None, :] / 2
    return centers, grid_sizes, inverse, unique_indices


_empty = torch.Tensor([])
_state = {"pairs_per_slot": 1.0}


def rasterize_views(cameras, means3D, opacity, scales=None, rotations=None, cov3D_precomp=None, shs=None,
                    colors_precomp=None, sh_degree=0, image_height=32, image_width=32, tanfovx=1.0, tanfovy=1.0, bg=None,
                    scale_modifier=1.0, labels=None, skip_labels=None, want_color=False, want_depth=False,
                    want_radii=False, views_per_batch=None):
    """Forward-render n views (cameras [n,36], pack_cameras / cube_cameras) of the same Gaussians.

    Per view the result equals ``_C.rasterize_gaussians`` on that view (same argument meaning).  labels [P] int32 +
    skip_labels [n] int32: view v leaves out the Gaussians with labels == skip_labels[v].
    Returns dict(alpha [n,1,H,W], color [n,3,H,W] | None, depth [n,1,H,W] | None, radii [n,P] int32 | None,
    num_rendered int)."""
    if not means3D.is_cuda:
        raise RuntimeError("libghr: tensor arguments must live on a CUDA device")
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:58-60
    if (scales is None or rotations is None) == (cov3D_precomp is None):
        raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
    if want_color and (shs is None) == (colors_precomp is None):
        raise Exception("Please provide excatly one of either SHs or precomputed colors!")
    if want_depth and not want_color:
        raise ValueError("depth is accumulated next to the colour: want_depth needs want_color")
    if cameras.dim() != 2 or cameras.shape[1] != CAMERA_FLOATS or cameras.dtype != torch.float32:
        raise ValueError("cameras must be a float32 [n,36] tensor (pack_cameras)")
    if (labels is None) != (skip_labels is None):
        raise ValueError("labels and skip_labels come together")
    dev = means3D.device
    n, P, H, W = cameras.shape[0], means3D.shape[0], int(image_height), int(image_width)
    cameras = cameras.to(dev).contiguous()
    if labels is not None:
        labels = labels.to(device=dev, dtype=torch.int32).contiguous()
        skip_labels = skip_labels.to(device=dev, dtype=torch.int32).contiguous()
        if labels.numel() != P or skip_labels.numel() != n:
            raise ValueError("labels must have one entry per Gaussian and skip_labels one per view")
    e = _empty
    keep = []
    bg_t = bg if bg is not None else torch.zeros(3, device=dev)
    prm = _params(keep, bg_t, means3D, colors_precomp if colors_precomp is not None else e, opacity,
                  scales if scales is not None else e, rotations if rotations is not None else e, scale_modifier,
                  cov3D_precomp if cov3D_precomp is not None else e, e, e, tanfovx, tanfovy, H, W,
                  shs if shs is not None else e, sh_degree, e, False, False)
    alpha = torch.empty(n, 1, H, W, device=dev)
    color = torch.empty(n, 3, H, W, device=dev) if want_color else None
    depth = torch.empty(n, 1, H, W, device=dev) if want_depth else None
    radii = torch.empty(n, P, dtype=torch.int32, device=dev) if want_radii else None
    if n == 0:
        return dict(alpha=alpha, color=color, depth=depth, radii=radii, num_rendered=0)
    if views_per_batch is None:
        # The blend of a batch ends when its longest tile list does (one dependent chain per pixel), so few large
        # batches beat many small ones: up to 64 M (view, Gaussian) slots per batch (3 GB of records on a 180 GB part).
        views_per_batch = max(1, min(n, (1 << 26) // max(P, 1)))
    l = _lib.lib()
    ptr = lambda t, o=0: None if t is None else t[o:].data_ptr()
    total_pairs = 0
    ws = None
    with _lib.on_device(dev):
        stream = _stream(dev)
        for lo in range(0, n, views_per_batch):
            nb = min(views_per_batch, n - lo)
            capacity = max(1 << 16, int(_state["pairs_per_slot"] * 1.25 * nb * P))
            while True:
                capacity = min(capacity, (1 << 30) - 1)
                need = l.ghr_views_workspace_bytes(P, nb, W, H, capacity)
                if need == 0:
                    raise ValueError("rasterize_views: batch too large (n_views * P must stay below 2^31)")
                if ws is None or ws.numel() < need:
                    ws = None
                    ws = torch.empty(need, dtype=torch.uint8, device=dev)
                _lib.check(l.ghr_forward_views(prm, nb, ptr(cameras, lo), ptr(labels), ptr(skip_labels, lo), ws.data_ptr(),
                                               ws.numel(), capacity, ptr(color, lo), ptr(depth, lo), ptr(alpha, lo),
                                               ptr(radii, lo), stream), "ghr_forward_views")
                hdr = ws[:8].view(torch.int32).tolist()  # synchronises: (num_rendered, error)
                if hdr[0] <= capacity:
                    if hdr[1] != 0:
                        raise _lib.GhrError(-hdr[1], "ghr_forward_views (device)")
                    break
                capacity = int(hdr[0] * 1.1) + 1024  # overflow: nothing was written, go again with room
            total_pairs += hdr[0]
            if P:
                _state["pairs_per_slot"] = max(_state["pairs_per_slot"] * 0.5, hdr[0] / float(nb * P))
    return dict(alpha=alpha, color=color, depth=depth, radii=radii, num_rendered=total_pairs)


def bake_opacity_cubemaps(means3D, opacity, scales, rotations, points=None, grid_res=10, res=32, scale_modifier=1.0,
                          znear=0.01, zfar=5.0):
    """The rasterisation part of bake_set (baking.py:204-284): occupied grid cells over `points` (default: means3D), six
    32x32 alpha maps per cell rendered from the cell centre without the cell's own Gaussians.

    Returns (opacity_cubemaps [C,6,res,res,1] -- torch.stack(opacity_cubemap) of :283-288 for every cell --,
    grid_centers [C,3], pc_grid_indices [P])."""
    points = means3D if points is None else points
    centers, _, cell_of, _ = pc_to_grid(points, grid_res)
    C = centers.shape[0]
    cams = cube_cameras(centers, znear, zfar)
    skip = torch.arange(C, device=means3D.device, dtype=torch.int32).repeat_interleave(6)
    out = rasterize_views(cams, points, opacity, scales=scales, rotations=rotations, image_height=res, image_width=res,
                          tanfovx=1.0, tanfovy=1.0, scale_modifier=scale_modifier, labels=cell_of.to(torch.int32),
                          skip_labels=skip)
    return out["alpha"].view(C, 6, res, res, 1), centers, cell_of


def envmap_dirs(res=(16, 32), device="cuda"):
    """baking.py:35-58 get_envmap_dirs: (solid_angles [H,W,1], directions [H,W,3]) of the lat-long grid."""
    H, W = res
    gy, gx = torch.meshgrid(torch.linspace(0.0, 1.0, H, device=device), torch.linspace(-1.0, 1.0, W, device=device),
                            indexing="ij")
    d_theta, d_phi = math.pi / H, 2 * math.pi / W
    sintheta, costheta = torch.sin(gy * math.pi), torch.cos(gy * math.pi)
    sinphi, cosphi = torch.sin(gx * math.pi), torch.cos(gx * math.pi)
    reflvec = torch.stack((sintheta * sinphi, costheta, -sintheta * cosphi), dim=-1)
    solid_angles = ((costheta - torch.cos(gy * math.pi + d_theta)) * d_phi)[..., None]
    return solid_angles, reflvec


def cubemap_sample_nearest(tex, dirs):
    """tex [n,6,res,res,C], dirs [...,3] -> [n, ..., C]: nearest-texel cube-map lookup (OpenGL / nvdiffrast face
    convention, ghr_cubemap_sample_nearest)."""
    if not tex.is_cuda:
        raise RuntimeError("libghr: tensor arguments must live on a CUDA device (there is no CPU fallback)")
    if tex.dim() != 5 or tex.shape[1] != 6 or tex.shape[2] != tex.shape[3] or dirs.shape[-1] != 3:
        raise ValueError("tex must be [n,6,res,res,C] and dirs [...,3]")
    tex = tex.float().contiguous()
    d = dirs.to(tex.device).float().reshape(-1, 3).contiguous()
    n, res, C = tex.shape[0], tex.shape[2], tex.shape[4]
    out = torch.empty(n, d.shape[0], C, device=tex.device)
    with _lib.on_device(tex.device):
        _lib.check(_lib.lib().ghr_cubemap_sample_nearest(n, res, C, tex.data_ptr(), d.shape[0], d.data_ptr(), out.data_ptr(),
                                                         _stream(tex.device)), "ghr_cubemap_sample_nearest")
    return out.view(n, *dirs.shape[:-1], C)


def texture(tex, uv, filter_mode="nearest", boundary_mode="cube"):
    """nvdiffrast.torch.texture for the one configuration bake_set uses (baking.py:290-297): tex [n,6,res,res,C],
    uv = directions [n,H,W,3], nearest filter, cube boundary -> [n,H,W,C].  Anything else raises."""
    if filter_mode != "nearest" or boundary_mode != "cube":
        raise NotImplementedError("only filter_mode='nearest', boundary_mode='cube' (what bake_set uses)")
    if uv.shape[0] != tex.shape[0]:
        raise ValueError("one direction grid per cube map")
    return torch.cat([cubemap_sample_nearest(tex[i:i + 1], uv[i]) for i in range(tex.shape[0])])


def bake_occlusion(means3D, normal, opacity, scales, rotations, H=16, W=32, grid_res=10, res=32, scale_modifier=1.0):
    """bake_set (baking.py:136-309) end to end: occlusion [P,H,W,1] = (envmap_dirs . normal > 0) * (1 - alpha of the
    Gaussian's grid cell seen from the cell centre in direction envmap_dirs).  means3D [P,3] world positions, normal
    [P,3], opacity [P,1], scales [P,3] (activated), rotations [P,4]."""
    cubes, centers, cell_of = bake_opacity_cubemaps(means3D, opacity, scales, rotations, grid_res=grid_res, res=res,
                                                    scale_modifier=scale_modifier)
    _, dirs = envmap_dirs((H, W), means3D.device)
    d = dirs.reshape(-1, 3).contiguous()
    P, C = means3D.shape[0], cubes.shape[0]
    nrm = normal.float().contiguous()
    cell = cell_of.to(torch.int64).contiguous()
    cubes = cubes.reshape(C, 6, res, res).contiguous()
    out = torch.empty(P, H * W, device=means3D.device)
    with _lib.on_device(means3D.device):
        _lib.check(_lib.lib().ghr_bake_occlusion(P, H * W, C, res, cubes.data_ptr(), cell.data_ptr(), nrm.data_ptr(),
                                                 d.data_ptr(), out.data_ptr(), _stream(means3D.device)),
                   "ghr_bake_occlusion")
    return out.view(P, H, W, 1)
