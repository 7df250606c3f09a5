"""render()'s per-Gaussian chain and rasteriser call as ONE differentiable operator on the fused kernels of libghr.so.

The reference's render() (gaussian_renderer/__init__.py:94-272) does, per view,

    coarse_deform_c2source -> get_covariance -> eval_sh + clamp -> get_opacity -> GaussianRasterizer(...)

``SMPLDeformer.deform_gaussians`` + ``GaussianRasterizer`` reproduce that as two operators with the deformed means, world
covariances, colours and opacities as tensors in between.  ``render_deformed`` is the same computation with those four
tensors never materialised: ghr_forward_deform_preprocess (LBS + covariance + SH + projection + tile scan in one kernel),
the rasteriser's binning / sort / blend, and ghr_backward_deformed (blend backward + one kernel for projection backward,
SH / covariance / skinning backward).  Same numerics as the two-operator chain: radii, tile lists and images bit for
bit, parameter gradients to fp32 round-off (tests/test_fused_gpu.py).  No CPU fallback.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import DeformParams, RasterParams, check, lib, state_layout
from .diff_gaussian_rasterization import _C
from .lbs import NJ, _f32c, _ptr, _stream


def _view_params(keep, P, rs, dev):
    """The camera / image half of ghr_raster_params (per-Gaussian pointers stay NULL)."""
    v = RasterParams()
    v.P, v.D, v.M = P, int(rs.sh_degree), 0
    v.W, v.H = int(rs.image_width), int(rs.image_height)
    v.scale_modifier = float(rs.scale_modifier)
    v.tan_fovx, v.tan_fovy = float(rs.tanfovx), float(rs.tanfovy)
    v.prefiltered, v.debug = int(bool(rs.prefiltered)), int(bool(rs.debug))
    for name, t in (("background", rs.bg), ("viewmatrix", rs.viewmatrix), ("projmatrix", rs.projmatrix),
                    ("campos", rs.campos)):
        t = _f32c(t.detach().to(dev)).reshape(-1)
        keep.append(t)
        setattr(v, name, t.data_ptr())
    return v


class _RenderDeformedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, deformer, params, t_params, t_vertices, rs, scales_are_log, want_aux, densify, means2D, xyz, normals,
                lbs_weights, correct_Rs, scaling, rotation, features, opacity_logit):
        self = deformer
        dev = self.device
        ctx.set_materialize_grads(False)
        if not xyz.is_cuda:
            raise RuntimeError("libghr: tensor arguments must live on a CUDA device (there is no CPU fallback)")
        P, H, W = xyz.shape[0], int(rs.image_height), int(rs.image_width)
        f = dict(dtype=torch.float32, device=dev)
        xyz_c = _f32c(xyz.detach())
        nrm = None if normals is None else _f32c(normals.detach())
        lw = None if lbs_weights is None else lbs_weights.detach().float()
        sc = _f32c(scaling.detach())
        rot = _f32c(rotation.detach())
        if rot.data_ptr() % 16:
            rot = rot.clone()
        shs = _f32c(features.detach())
        ol = _f32c(opacity_logit.detach())
        M = shs.shape[1]
        out_color = torch.empty(3, H, W, **f)
        out_depth = torch.empty(1, H, W, **f)
        out_alpha = torch.empty(1, H, W, **f)
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        if P == 0:
            ctx.P = 0
            for t in (out_color, out_depth, out_alpha):
                t.zero_()
            ctx.mark_non_differentiable(radii)
            return out_color, radii, out_depth, out_alpha, None, None, None, None, None
        pose_big, pose_tgt, off, vid = self.prepare_pose(xyz_c, params, t_params, t_vertices, correct_Rs)
        keep = []
        view = _view_params(keep, P, rs, dev)
        p = DeformParams()
        p.P, p.V, p.sh_degree, p.M = P, self.V, int(rs.sh_degree), M
        p.scale_modifier = float(rs.scale_modifier)
        p.xyz, p.normals, p.vert_ids, p.weights = xyz_c.data_ptr(), _ptr(nrm), vid.data_ptr(), self.weights.data_ptr()
        if lw is not None:
            p.lbs_offsets = lw.data_ptr()
            p.lbs_stride_p, p.lbs_stride_j = lw.stride(0), lw.stride(1)
        p.vertex_offsets, p.pose_big, p.pose_target = off.data_ptr(), pose_big.data_ptr(), pose_tgt.data_ptr()
        p.scales, p.rotations, p.shs, p.campos = sc.data_ptr(), rot.data_ptr(), shs.data_ptr(), view.campos
        p.scales_are_log = int(bool(scales_are_log))
        p.opacity_logits = ol.data_ptr()

        aux = [None] * 5  # world_pts, world_normals, cov3D, colors, opacity
        if want_aux:
            aux = [torch.empty(P, 3, **f), torch.empty(P, 3, **f) if nrm is not None else None, torch.empty(P, 6, **f),
                   torch.empty(P, 3, **f), torch.empty(P, 1, **f)]
        with _lib.on_device(dev):
            L0 = state_layout(P, 0, W, H)
            geom = torch.empty(L0.geom_bytes, dtype=torch.uint8, device=dev)
            img = torch.empty(L0.img_bytes, dtype=torch.uint8, device=dev)
            l, s = lib(), _stream()

            def preprocess(counter_ptr):
                check(l.ghr_forward_deform_preprocess(C.byref(p), C.byref(view), radii.data_ptr(), geom.data_ptr(),
                                                      geom.numel(), *[_ptr(t) for t in aux], counter_ptr, s),
                      "ghr_forward_deform_preprocess")

            def render(capacity):
                L = state_layout(P, capacity, W, H)
                binning = torch.empty(L.binning_bytes, dtype=torch.uint8, device=dev)
                check(l.ghr_forward_render_deformed(C.byref(view), geom.data_ptr(), geom.numel(), binning.data_ptr(),
                                                    binning.numel(), capacity, img.data_ptr(), img.numel(),
                                                    out_color.data_ptr(), out_depth.data_ptr(), out_alpha.data_ptr(), s),
                      "ghr_forward_render_deformed")
                return binning

            num_rendered, binning = _C.launch_with_capacity(dev, P, W, H, preprocess, render)
            if rs.debug:
                _C.read_status(geom, raise_on_error=True)
        _C._last_geom = geom
        ctx.P, ctx.deformer, ctx.p, ctx.view, ctx.densify = P, self, p, view, densify
        ctx.keep = (keep, xyz_c, nrm, lw, sc, rot, shs, ol, pose_big, pose_tgt, off, vid)
        ctx.state = (radii, geom, binning, img)
        ctx.correct_Rs = correct_Rs
        ctx.has = (normals is not None, lbs_weights is not None, correct_Rs is not None)
        ctx.lw_shape = None if lbs_weights is None else lbs_weights.shape
        ctx.ol_shape = opacity_logit.shape
        ctx.mark_non_differentiable(radii, *[t for t in aux if t is not None])
        return (out_color, radii, out_depth, out_alpha, *aux)

    @staticmethod
    def backward(ctx, g_color, g_radii, g_depth, g_alpha, *g_aux):
        n_in = 17
        if ctx.P == 0:
            return (None,) * n_in
        self, p, view = ctx.deformer, ctx.p, ctx.view
        keep, xyz_c, nrm, lw, sc, rot, shs, ol, pose_big, pose_tgt, off, vid = ctx.keep
        radii, geom, binning, img = ctx.state
        has_n, has_lw, has_cr = ctx.has
        dev = self.device
        P, M, H, W = p.P, p.M, view.H, view.W
        f = dict(dtype=torch.float32, device=dev)
        zero = lambda *shape: torch.zeros(*shape, **f)
        g = [zero(*shape) if t is None else _f32c(t)
             for t, shape in ((g_color, (3, H, W)), (g_depth, (1, H, W)), (g_alpha, (1, H, W)))]
        capacity = _C._capacity_of(binning, P, W, H)
        need = ctx.needs_input_grad[1:]  # (after dropping `densify`) 7 = means2D, 8.. = xyz, normals, lbs_weights, correct_Rs, ...
        stats = [None, None, None]
        if ctx.densify is not None:
            stats = list(ctx.densify)
            for t in stats:
                assert t is None or (t.is_cuda and t.dtype == torch.float32 and t.numel() == P and t.is_contiguous())
        accum = zero(P * 12)
        d_m2d = torch.empty(P, 3, **f)
        sink = self.grad_sink
        if sink is not None:
            assert not has_lw, "grad_sink does not cover lbs_weights"
            d_xyz, d_sc, d_rot, d_sh, d_ol = (sink[k] for k in ("xyz", "scaling", "rotation", "features", "opacity_logit"))
            for t in (d_xyz, d_sc, d_rot, d_sh, d_ol):
                assert t.is_contiguous() and t.dtype == torch.float32 and t.data_ptr() % 16 == 0
            d_lw = None
            p.accumulate_grads = 2 if self.grad_sink_atomic else 1
        else:
            p.accumulate_grads = 0
            d_xyz, d_sc, d_rot = torch.empty(P, 3, **f), torch.empty(P, 3, **f), torch.empty(P, 4, **f)
            d_sh, d_ol = torch.empty(P, M, 3, **f), torch.empty(P, **f)
            d_lw = torch.empty(P, NJ, **f) if has_lw else None
        p.dL_dopacity_logits = d_ol.data_ptr()
        want_pose = has_cr and need[11]
        d_A = zero(NJ * 12) if want_pose else None
        d_off = zero(self.V, 3) if want_pose else None
        with _lib.on_device(dev):
            check(lib().ghr_backward_deformed(
                C.byref(p), C.byref(view), radii.data_ptr(), geom.data_ptr(), geom.numel(), binning.data_ptr(),
                binning.numel(), capacity, img.data_ptr(), img.numel(), g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(),
                accum.data_ptr(), d_m2d.data_ptr(), d_xyz.data_ptr(), _ptr(d_lw), d_sc.data_ptr(), d_rot.data_ptr(),
                d_sh.data_ptr(), _ptr(d_A), _ptr(d_off), *[_ptr(t) for t in stats], _stream()), "ghr_backward_deformed")
            d_cr = None
            if want_pose:
                buf = torch.empty(23 * 9 + 207, **f)
                check(lib().ghr_lbs_pose_backward(C.byref(self._model), pose_tgt.data_ptr(), None, d_A.data_ptr(),
                                                  d_off.data_ptr(), buf.data_ptr(), _stream()), "ghr_lbs_pose_backward")
                d_cr = buf[:207].view(23, 3, 3).reshape(ctx.correct_Rs.shape)
        g_m2d = d_m2d if need[7] else None
        if sink is not None:  # the per-Gaussian gradients were added to the sink tensors
            return (None,) * 8 + (g_m2d, None, None, None, d_cr, None, None, None, None)
        zn = torch.zeros_like(nrm) if has_n and need[9] else None  # normals do not reach the images
        return (None,) * 8 + (g_m2d, d_xyz, zn, None if d_lw is None else d_lw.view(ctx.lw_shape), d_cr, d_sc, d_rot, d_sh,
                              d_ol.view(ctx.ol_shape))


def render_deformed(deformer, xyz, params, t_params, t_vertices, raster_settings, *, rotation, features, opacity_logit,
                    scaling=None, log_scaling=None, normals=None, lbs_weights=None, correct_Rs=None, means2D=None,
                    return_aux=False, densify=None):
    """Deform the canonical Gaussians to the pose `params` and rasterise them with `raster_settings`
    (GaussianRasterizationSettings; its sh_degree is the active SH degree, its campos the SH view point).

    xyz [P,3]; scaling [P,3] activated or log_scaling [P,3] raw (exp inside); rotation [P,4] raw; features [P,M,3];
    opacity_logit [P,1] raw (sigmoid inside); normals [P,3] optional (only for the `world_normals` aux output);
    lbs_weights [P,24] logit offsets (LBSOffsetDecoder) and correct_Rs [23,3,3] (BodyPoseRefiner) optional;
    means2D [P,3]: optional leaf whose .grad receives the screen-space gradient (densification statistics).
    densify: optional (max_radii2D, xyz_gradient_accum, denom) fp32 tensors with P entries (any of them None), updated IN
    the backward kernel for the visible Gaussians as train.py:401-404 / add_densification_stats do after it.

    Returns (color[3,H,W], radii[P], depth[1,H,W], alpha[1,H,W]) and, with return_aux=True, a dict of the intermediate
    values (means3D, world_normals, cov3D, colors, opacity) as non-differentiable copies."""
    if (scaling is None) == (log_scaling is None):
        raise ValueError("pass either scaling or log_scaling")
    out = _RenderDeformedFn.apply(deformer, params, t_params, t_vertices, raster_settings, log_scaling is not None,
                                  bool(return_aux), densify, means2D, xyz, normals, lbs_weights, correct_Rs,
                                  log_scaling if scaling is None else scaling, rotation, features, opacity_logit)
    color, radii, depth, alpha = out[:4]
    if not return_aux:
        return color, radii, depth, alpha
    return color, radii, depth, alpha, dict(zip(("means3D", "world_normals", "cov3D", "colors", "opacity"), out[4:]))
