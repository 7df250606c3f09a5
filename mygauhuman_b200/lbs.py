"""LBS deformation of SMPL-initialised Gaussians on libghr.so -- the host-side mirror of the
reference's LBS functions (same names, arguments and return conventions):

* ``SMPLDeformer.coarse_deform_c2source``   reference scene/gaussian_model.py:768-872
* ``SMPLDeformer.get_transform_params_torch`` reference scene/gaussian_model.py:947-980
* ``batch_rodrigues``                         reference scene/gaussian_model.py:894-912,982-1013; smplx/lbs.py:299-333
* ``build_covariance_from_scaling_rotation``  reference scene/gaussian_model.py:35-42
* ``eval_sh``                                 reference utils/sh_utils.py:135-190
* ``SMPLDeformer.lbs``                        reference smplx/lbs.py:156-252 (4-value variant of this fork)
* ``SMPLDeformer.deform_gaussians``           the fused fast path render() needs: LBS + T Sigma T^T + SH colour
                                              in one kernel (reference gaussian_renderer/__init__.py:94-195)

All arithmetic runs in CUDA kernels of libghr.so (ghr_lbs_pose.cu, ghr_lbs_deform.cu,
ghr_knn.cu); there is no PyTorch/CPU fallback.  The reference's third-party ``knn_cuda``
nearest-vertex search is replaced by an exact grid search (``ghr_nearest_vertex``).
"""
import ctypes as C

import torch

from . import _lib
from ._lib import DeformParams, LbsModel, check, lib

NJ = 24


def _stream():
    return _lib.raw_stream(torch._C._cuda_getDevice())


def _f32c(t):
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _ptr(t):
    return None if t is None else t.data_ptr()


class SMPLDeformer:
    """Device-resident SMPL-topology body model + the LBS operators of the hot path.

    ``smpl`` is the reference's ``SMPL_NEUTRAL`` dict (SMPL_to_tensor keys, gaussian_model.py:880-892):
    v_template [V,3], shapedirs [V,3,NB], posedirs [V*3,207] (or [V,3,207]), J_regressor [24,V],
    weights [V,24], kintree_table [2,24].
    """

    def __init__(self, smpl, device="cuda"):
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("SMPLDeformer needs a CUDA device (libghr has no CPU fallback)")
        if dev.index is None:  # "cuda" = the current device, fixed here so every later call agrees on it
            dev = torch.device("cuda", torch.cuda.current_device())
        lib()
        self.device = dev
        f = lambda k: _f32c(smpl[k].to(dev))
        self.v_template = f("v_template")
        self.shapedirs = f("shapedirs")
        self.V = self.v_template.shape[0]
        self.NB = self.shapedirs.shape[-1]
        self.posedirs = f("posedirs").reshape(self.V * 3, -1).contiguous()
        assert self.posedirs.shape[1] == 207, "posedirs must have 207 pose-feature columns"
        self.weights = f("weights")
        assert self.weights.shape == (self.V, NJ)
        self.J_regressor = f("J_regressor")
        self.parents = smpl["kintree_table"][0].to(dev).to(torch.int32).contiguous()
        self.joint_template = torch.empty(NJ, 3, device=dev)
        self.joint_shapedirs = torch.empty(NJ, 3, max(self.NB, 1), device=dev)
        with _lib.on_device(dev):
            check(lib().ghr_lbs_prepare(self.V, self.NB, self.v_template.data_ptr(), self.shapedirs.data_ptr(),
                                        self.J_regressor.data_ptr(), self.joint_template.data_ptr(),
                                        self.joint_shapedirs.data_ptr(), _stream()), "ghr_lbs_prepare")
        self._model = LbsModel(self.V, self.NB, self.weights.data_ptr(), self.posedirs.data_ptr(),
                               self.shapedirs.data_ptr(), self.joint_template.data_ptr(),
                               self.joint_shapedirs.data_ptr(), self.parents.data_ptr())
        self._grid_key, self._grid = None, None
        self._pose_key, self._pose_cached, self._pose_src = None, None, ()
        self._side_stream = torch.cuda.Stream(device=self.device) if torch.cuda.is_available() else None
        # Optional gradient sink (multi-view accumulation, see graphed.py): dict name -> fp32 contiguous tensor for
        # "xyz", "normals", "scaling", "rotation", "features", "opacity_logit".  When set, the backward kernels ADD the
        # per-Gaussian gradients of those inputs into the given tensors and autograd receives None for them.
        self.grad_sink = None
        self.vert_ids_override = None  # int32 [P]: nearest-vertex ids computed by the caller (once per multi-view batch)
        self.grad_sink_atomic = False  # True: add with atomics (several streams share the sink, MultiLaneRenderStep)
        self.pose_override = None  # (pose buffer, vertex offsets) of the target pose computed by the caller (pose_views)

    # ---- pose stage --------------------------------------------------------
    def pose_buffer(self, params, correct_Rs=None):
        """ghr_lbs_pose: one single-warp kernel -> pose buffer (layout in include/ghr.h)."""
        dev = self.device
        poses = _f32c(params["poses"].to(dev)).reshape(-1)
        shapes = _f32c(params["shapes"].to(dev)).reshape(-1)[: self.NB]
        R = _f32c(params["R"].to(dev)).reshape(-1)
        Th = _f32c(params["Th"].to(dev)).reshape(-1)
        cr = None if correct_Rs is None else _f32c(correct_Rs.detach().to(dev)).reshape(-1)
        out = torch.empty(_lib.POSE_FLOATS, device=dev)
        with _lib.on_device(dev):
            check(lib().ghr_lbs_pose(C.byref(self._model), poses.data_ptr(), shapes.data_ptr(), _ptr(cr), R.data_ptr(),
                                     Th.data_ptr(), out.data_ptr(), _stream()), "ghr_lbs_pose")
        return out

    def _big_pose_buffer(self, t_params):
        """Pose buffer of the big pose, cached while the caller keeps passing the same (unmodified) tensors.  The key
        tensors are held alive (`_pose_src`), so the allocator cannot hand their addresses to new per-view tensors with
        other contents; a miss during CUDA-graph capture is computed but not cached (a graph-pool buffer must not
        leak into later eager calls)."""
        src = tuple(t_params[k] for k in ("poses", "shapes", "R", "Th"))
        key = tuple((t.data_ptr(), t._version) for t in src)
        if key == self._pose_key and all(a is b for a, b in zip(src, self._pose_src)):
            return self._pose_cached
        buf = self.pose_buffer(t_params)
        if not torch.cuda.is_current_stream_capturing():
            self._pose_cached, self._pose_key, self._pose_src = buf, key, src
        return buf

    def get_transform_params_torch(self, params, rot_mats=None, correct_Rs=None):
        """(A[1,24,4,4], R, Th, joints[1,24,3]) as the reference returns them."""
        if rot_mats is not None:
            raise NotImplementedError("rot_mats override: pass poses (+ correct_Rs) instead")
        buf = self.pose_buffer(params, correct_Rs)
        A = torch.zeros(1, NJ, 4, 4, device=self.device)
        A[0, :, :3, :] = buf[_lib.POSE_A:_lib.POSE_A + 288].view(NJ, 3, 4)
        A[0, :, 3, 3] = 1.0
        joints = buf[_lib.POSE_JOINTS:_lib.POSE_JOINTS + 72].view(1, NJ, 3).clone()
        return A, params["R"], params["Th"], joints

    def vertex_offsets(self, pose_big, pose_target, shapes):
        off = torch.empty(self.V, 3, device=self.device)
        sh = None if shapes is None else _f32c(shapes.to(self.device)).reshape(-1)[: self.NB]
        with _lib.on_device(self.device):
            check(lib().ghr_lbs_vertex_offsets(C.byref(self._model), _ptr(pose_big), pose_target.data_ptr(), _ptr(sh),
                                               off.data_ptr(), _stream()), "ghr_lbs_vertex_offsets")
        return off

    def pose_views(self, t_params, views, outs):
        """The pose stage of several views in one launch (ghr_lbs_pose_views): views = list of dicts with device tensors
        poses [72], shapes [>=NB], R [9], Th [3] (fp32, contiguous; slices of a packed view vector are fine) and
        optionally correct_Rs [23,3,3]; outs = list of (pose buffer [POSE_FLOATS], vertex offsets [V,3]) to fill.  The big
        pose `t_params` is the one the offsets are taken against."""
        assert len(views) == len(outs)
        pose_big = self._big_pose_buffer(t_params)
        with _lib.on_device(self.device):
            for s0 in range(0, len(views), _lib.POSE_VIEWS_MAX):
                chunk = list(zip(views[s0:s0 + _lib.POSE_VIEWS_MAX], outs[s0:s0 + _lib.POSE_VIEWS_MAX]))
                arr = (_lib.PoseView * len(chunk))()
                for a, (v, (pose_out, off_out)) in zip(arr, chunk):
                    for k in ("poses", "shapes", "R", "Th"):
                        t = v[k]
                        assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous(), k
                        setattr(a, k, t.data_ptr())
                    cr = v.get("correct_Rs")
                    a.correct_Rs = None if cr is None else _f32c(cr.detach()).data_ptr()
                    assert pose_out.numel() >= _lib.POSE_FLOATS and off_out.numel() == self.V * 3
                    a.pose_out, a.offsets_out = pose_out.data_ptr(), off_out.data_ptr()
                check(lib().ghr_lbs_pose_views(C.byref(self._model), len(chunk), arr, pose_big.data_ptr(), _stream()),
                      "ghr_lbs_pose_views")

    # ---- nearest vertex ----------------------------------------------------
    def nearest_vertex(self, t_vertices, query, out=None):
        """Exact nearest vertex index [P] int32 (the reference's knn_cuda KNN(k=1), gaussian_model.py:775)."""
        tv = _f32c(t_vertices.detach()).reshape(-1, 3)
        key = (tv.data_ptr(), t_vertices._version, tv.shape[0])
        with _lib.on_device(self.device):
            if key != self._grid_key:
                nbytes = lib().ghr_vertex_grid_workspace_bytes(tv.shape[0])
                ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
                check(lib().ghr_vertex_grid_build(tv.shape[0], tv.data_ptr(), ws.data_ptr(), ws.numel(), _stream()),
                      "ghr_vertex_grid_build")
                self._grid, self._grid_key, self._grid_src = ws, key, tv
            q = _f32c(query.detach()).reshape(-1, 3)
            vid = torch.empty(q.shape[0], dtype=torch.int32, device=self.device) if out is None else out
            check(lib().ghr_nearest_vertex(q.shape[0], q.data_ptr(), self._grid_key[2], self._grid.data_ptr(), vid.data_ptr(),
                                           _stream()),
                  "ghr_nearest_vertex")
        return vid

    def prepare_pose(self, xyz_c, params, t_params, t_vertices, correct_Rs=None):
        """Everything the per-Gaussian kernels need besides the Gaussians: (pose buffer of the big pose, pose buffer of the
        target pose, pose blend-shape vertex offsets [V,3], nearest-vertex ids [P] int32)."""
        dev = self.device
        P = xyz_c.shape[0]
        pose_big = self._big_pose_buffer(t_params)
        # The nearest-vertex search (depends on xyz only) runs on a side stream next to the pose chain and the pose
        # blend-shape offsets (depend on the pose only); both are short, latency-bound kernels.  Fork/join through
        # stream waits, so the overlap is kept when the step is captured into a CUDA graph.  Every buffer the side
        # stream touches is allocated on the caller's stream beforehand.
        vid = self.vert_ids_override
        have_vid = vid is not None
        if have_vid:
            assert vid.dtype == torch.int32 and vid.numel() == P and vid.is_contiguous() and vid.device == xyz_c.device
        else:
            vid = torch.empty(P, dtype=torch.int32, device=dev)
        tv = _f32c(t_vertices.detach())
        cur = torch.cuda.current_stream(dev)
        side = self._side_stream if not have_vid else None
        if have_vid:
            pass
        elif P and side is not None:
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                self.nearest_vertex(tv, xyz_c, out=vid)
        if self.pose_override is not None:  # computed for the whole batch by pose_views (MultiLaneRenderStep)
            pose_tgt, off = self.pose_override
        else:
            pose_tgt = self.pose_buffer(params, correct_Rs)
            off = self.vertex_offsets(pose_big, pose_tgt, params["shapes"])
        if have_vid:
            pass
        elif P and side is not None:
            cur.wait_stream(side)
        elif P:
            self.nearest_vertex(tv, xyz_c, out=vid)
        return pose_big, pose_tgt, off, vid

    # ---- the operators -----------------------------------------------------
    def deform_gaussians(self, xyz, params, t_params, t_vertices, normals=None, lbs_weights=None, correct_Rs=None,
                         scaling=None, rotation=None, scaling_modifier=1.0, features=None, sh_degree=0, campos=None,
                         return_transl=False, log_scaling=None, opacity_logit=None):
        """Fused LBS + world covariance + view-dependent colour.  xyz [P,3]; lbs_weights [P,24] (any strides);
        scaling [P,3] (activated), rotation [P,4] (raw), features [P,M,3].
        Optionally the reference's parameter activations are fused in as well: pass `log_scaling` [P,3] instead of
        `scaling` (exp, get_scaling) and/or `opacity_logit` [P,1] (sigmoid, get_opacity -> returned as "opacity").
        Returns dict(smpl_pts, means3D, bweights, transforms, translation, world_normals, cov3D, colors, opacity)."""
        if log_scaling is not None:
            if scaling is not None:
                raise ValueError("pass either scaling or log_scaling")
            scaling = log_scaling
        out = _DeformFn.apply(self, params, t_params, t_vertices, float(scaling_modifier), int(sh_degree), campos,
                              bool(return_transl), log_scaling is not None, xyz, normals, lbs_weights, correct_Rs, scaling,
                              rotation, features, opacity_logit)
        keys = ("smpl_pts", "means3D", "bweights", "transforms", "translation", "world_normals", "cov3D", "colors",
                "opacity")
        return dict(zip(keys, out))

    def coarse_deform_c2source(self, query_pts, params, t_params, t_vertices, lbs_weights=None, correct_Rs=None,
                               return_transl=False, normals=None):
        """Drop-in for GaussianModel.coarse_deform_c2source (batch size 1): returns
        (smpl_src_pts[1,P,3], world_src_pts[1,P,3], bweights[1,P,24], transforms[1,P,3,3], translation|None,
        world_normals[1,P,3]|None)."""
        assert query_pts.dim() == 3 and query_pts.shape[0] == 1, "batch size 1, as the reference uses it"
        d = self.deform_gaussians(query_pts[0], params, t_params, t_vertices[0] if t_vertices.dim() == 3 else t_vertices,
                                  normals=None if normals is None else normals[0],
                                  lbs_weights=None if lbs_weights is None else lbs_weights[0], correct_Rs=correct_Rs,
                                  return_transl=return_transl)
        u = lambda t: None if t is None else t[None]
        return (u(d["smpl_pts"]), u(d["means3D"]), u(d["bweights"]), u(d["transforms"]), u(d["translation"]),
                u(d["world_normals"]))

    def lbs(self, betas, pose):
        """smplx lbs() over the model's own vertices (BASELINE.json config 1), batch size 1:
        (verts[1,V,3], J_transformed[1,24,3], A[1,24,4,4], T[1,V,4,4])."""
        dev = self.device
        params = {"poses": pose, "shapes": betas, "R": torch.eye(3, device=dev), "Th": torch.zeros(3, device=dev)}
        buf = self.pose_buffer(params)
        off = self.vertex_offsets(None, buf, betas)
        verts = torch.empty(self.V, 3, device=dev)
        T12 = torch.empty(self.V, 12, device=dev)
        with _lib.on_device(dev):
            check(lib().ghr_smpl_lbs_vertices(C.byref(self._model), self.v_template.data_ptr(), buf.data_ptr(),
                                              off.data_ptr(), verts.data_ptr(), T12.data_ptr(), _stream()),
                  "ghr_smpl_lbs_vertices")
        A = torch.zeros(1, NJ, 4, 4, device=dev)
        A[0, :, :3, :] = buf[_lib.POSE_A:_lib.POSE_A + 288].view(NJ, 3, 4)
        A[0, :, 3, 3] = 1.0
        T = torch.zeros(1, self.V, 4, 4, device=dev)
        T[0, :, :3, :] = T12.view(self.V, 3, 4)
        T[0, :, 3, 3] = 1.0
        Jt = buf[_lib.POSE_G:_lib.POSE_G + 288].view(NJ, 3, 4)[:, :, 3].clone()[None]
        return verts[None], Jt, A, T


class _DeformFn(torch.autograd.Function):
    """ghr_lbs_deform_forward / _backward (+ ghr_lbs_pose_backward for correct_Rs)."""

    @staticmethod
    def forward(ctx, deformer, params, t_params, t_vertices, scaling_modifier, sh_degree, campos, return_transl,
                scales_are_log, xyz, normals, lbs_weights, correct_Rs, scaling, rotation, features, opacity_logit):
        self = deformer
        dev = self.device
        ctx.set_materialize_grads(False)  # unused outputs arrive as None in backward instead of zero-filled tensors
        if not xyz.is_cuda:
            raise RuntimeError("libghr: tensor arguments must live on a CUDA device (there is no CPU fallback)")
        P = xyz.shape[0]
        xyz_c = _f32c(xyz.detach())
        nrm = None if normals is None else _f32c(normals.detach())
        lw = None if lbs_weights is None else lbs_weights.detach().float()
        sc = None if scaling is None else _f32c(scaling.detach())
        rot = None if rotation is None else _f32c(rotation.detach())
        if rot is not None and rot.data_ptr() % 16:
            rot = rot.clone()
        shs = None if features is None else _f32c(features.detach())
        M = 0 if shs is None else shs.shape[1]
        cam = None if campos is None else _f32c(campos.detach().to(dev)).reshape(-1)
        pose_big, pose_tgt, off, vid = self.prepare_pose(xyz_c, params, t_params, t_vertices, correct_Rs)

        p = DeformParams()
        p.P, p.V, p.sh_degree, p.M = P, self.V, sh_degree, M
        p.scale_modifier = scaling_modifier
        p.xyz, p.normals, p.vert_ids, p.weights = xyz_c.data_ptr(), _ptr(nrm), vid.data_ptr(), self.weights.data_ptr()
        if lw is not None:
            p.lbs_offsets = lw.data_ptr()
            p.lbs_stride_p, p.lbs_stride_j = lw.stride(0), lw.stride(1)
        p.vertex_offsets, p.pose_big, p.pose_target = off.data_ptr(), pose_big.data_ptr(), pose_tgt.data_ptr()
        p.scales, p.rotations, p.shs, p.campos = _ptr(sc), _ptr(rot), _ptr(shs), _ptr(cam)
        p.scales_are_log = int(bool(scales_are_log) and sc is not None)
        ol = None if opacity_logit is None else _f32c(opacity_logit.detach())

        f = dict(dtype=torch.float32, device=dev)
        smpl_pts = torch.empty(P, 3, **f)
        world = torch.empty(P, 3, **f)
        bw = torch.empty(P, NJ, **f)
        transforms = torch.empty(P, 3, 3, **f)
        transl = torch.empty(P, 3, **f) if return_transl else None
        wn = torch.empty(P, 3, **f) if nrm is not None else None
        cov = torch.empty(P, 6, **f) if sc is not None else None
        col = torch.empty(P, 3, **f) if shs is not None else None
        clamped = torch.empty(P, dtype=torch.uint8, device=dev) if shs is not None else None
        opac = None
        if ol is not None:
            opac = torch.empty(ol.shape, **f)
            p.opacity_logits, p.opacity_out = ol.data_ptr() if P else None, opac.data_ptr() if P else None
        if P:
            with _lib.on_device(dev):
                check(lib().ghr_lbs_deform_forward(C.byref(p), smpl_pts.data_ptr(), world.data_ptr(), bw.data_ptr(),
                                                   transforms.data_ptr(), _ptr(transl), _ptr(wn), _ptr(cov), _ptr(col),
                                                   _ptr(clamped), _stream()), "ghr_lbs_deform_forward")
        ctx.deformer, ctx.p = self, p
        ctx.keep = (xyz_c, nrm, lw, sc, rot, shs, cam, pose_big, pose_tgt, off, vid, clamped)
        ctx.ol = ol
        ctx.correct_Rs = correct_Rs
        ctx.has = (normals is not None, lbs_weights is not None, correct_Rs is not None, scaling is not None,
                   rotation is not None, features is not None)
        ctx.lw_shape = None if lbs_weights is None else lbs_weights.shape
        outs = (smpl_pts, world, bw, transforms, transl, wn, cov, col, opac)
        ctx.mark_non_differentiable(*[t for t in (smpl_pts, bw, transl) if t is not None])
        return outs

    @staticmethod
    def backward(ctx, g_smpl, g_world, g_bw, g_transforms, g_transl, g_wn, g_cov, g_col, g_opac):
        self, p = ctx.deformer, ctx.p
        xyz_c, nrm, lw, sc, rot, shs, cam, pose_big, pose_tgt, off, vid, clamped = ctx.keep
        has_n, has_lw, has_cr, has_sc, has_rot, has_sh = ctx.has
        dev = self.device
        P, M = p.P, p.M
        f = dict(dtype=torch.float32, device=dev)
        c = lambda t: None if t is None else _f32c(t)
        g_world, g_transforms, g_wn, g_cov, g_col = c(g_world), c(g_transforms), c(g_wn), c(g_cov), c(g_col)
        need = ctx.needs_input_grad  # positions 9.. = xyz, normals, lbs_weights, correct_Rs, scaling, rotation, features, logit
        d_ol = None
        if ctx.ol is not None and P:
            g_opac = c(g_opac)
            d_ol = torch.empty(ctx.ol.shape, **f)
            p.dL_dopacity, p.dL_dopacity_logits = _ptr(g_opac), d_ol.data_ptr()
        sink = self.grad_sink
        if sink is not None:
            pick = lambda k, on: sink[k] if on else None
            d_xyz, d_n, d_lw = sink["xyz"], pick("normals", has_n), None
            d_sc, d_rot, d_sh = pick("scaling", has_sc), pick("rotation", has_sc), pick("features", has_sh)
            assert not has_lw, "grad_sink does not cover lbs_weights"
            for t in (d_xyz, d_n, d_sc, d_rot, d_sh):
                assert t is None or (t.is_contiguous() and t.dtype == torch.float32 and t.data_ptr() % 16 == 0)
            p.accumulate_grads = 2 if self.grad_sink_atomic else 1
            if ctx.ol is not None and P:
                d_ol = sink["opacity_logit"]
                p.dL_dopacity_logits = d_ol.data_ptr()
        else:
            p.accumulate_grads = 0
            d_xyz = torch.empty(P, 3, **f)
            d_n = torch.empty(P, 3, **f) if has_n else None
            d_lw = torch.empty(P, NJ, **f) if has_lw else None
            d_sc = torch.empty(P, 3, **f) if has_sc else None
            d_rot = torch.empty(P, 4, **f) if has_sc else None
            d_sh = torch.empty(P, M, 3, **f) if has_sh else None
        want_pose = has_cr and need[12]
        d_A = torch.zeros(NJ * 12, **f) if want_pose else None
        d_off = torch.zeros(self.V, 3, **f) if want_pose else None
        if P:
            with _lib.on_device(dev):
                check(lib().ghr_lbs_deform_backward(
                    C.byref(p), _ptr(clamped), _ptr(g_world), _ptr(g_transforms), _ptr(g_wn), _ptr(g_cov), _ptr(g_col),
                    d_xyz.data_ptr(), _ptr(d_n), _ptr(d_lw), _ptr(d_sc), _ptr(d_rot), _ptr(d_sh), _ptr(d_A), _ptr(d_off),
                    _stream()), "ghr_lbs_deform_backward")
        d_cr = None
        if want_pose:
            buf = torch.empty(23 * 9 + 207, **f)
            with _lib.on_device(dev):
                check(lib().ghr_lbs_pose_backward(C.byref(self._model), pose_tgt.data_ptr(), None, d_A.data_ptr(),
                                                  d_off.data_ptr(), buf.data_ptr(), _stream()), "ghr_lbs_pose_backward")
            d_cr = buf[:207].view(23, 3, 3).reshape(ctx.correct_Rs.shape)
        if d_lw is not None:
            d_lw = d_lw.view(ctx.lw_shape)
        if sink is not None:  # already added to the sink tensors
            return (None,) * 12 + (d_cr, None, None, None, None)
        return (None,) * 9 + (d_xyz, d_n, d_lw, d_cr, d_sc, d_rot if has_rot else None, d_sh, d_ol)


# ---- module-level mirrors of the small reference helpers ---------------------

def batch_rodrigues(rot_vecs, epsilon=1e-8, dtype=torch.float32):
    """Rotation matrices [N,3,3] of N axis-angle vectors, via the pose kernel (24 at a time)."""
    dev = rot_vecs.device
    if dev.type != "cuda":
        raise RuntimeError("libghr: tensor arguments must live on a CUDA device (there is no CPU fallback)")
    rv = _f32c(rot_vecs.detach()).reshape(-1, 3)
    N = rv.shape[0]
    out = torch.empty(N, 3, 3, device=dev)
    model = _rodrigues_model(dev)
    eye, zero = torch.eye(3, device=dev).reshape(-1), torch.zeros(3, device=dev)
    for s in range(0, N, NJ):
        chunk = torch.zeros(NJ, 3, device=dev)
        n = min(NJ, N - s)
        chunk[:n] = rv[s:s + n]
        buf = torch.empty(_lib.POSE_FLOATS, device=dev)
        with _lib.on_device(dev):
            check(lib().ghr_lbs_pose(C.byref(model[0]), chunk.data_ptr(), zero.data_ptr(), None, eye.data_ptr(),
                                     zero.data_ptr(), buf.data_ptr(), _stream()), "ghr_lbs_pose")
        out[s:s + n] = buf[_lib.POSE_ROD:_lib.POSE_ROD + 216].view(NJ, 3, 3)[:n]
    return out.to(dtype)


_RODRIGUES_MODELS = {}


def _rodrigues_model(dev):
    """A dummy one-vertex model so ghr_lbs_pose can be used for its Rodrigues stage alone."""
    m = _RODRIGUES_MODELS.get(dev)
    if m is None:
        w = torch.zeros(1, NJ, device=dev)
        pd = torch.zeros(3, 207, device=dev)
        jt = torch.zeros(NJ, 3, device=dev)
        par = torch.zeros(NJ, dtype=torch.int32, device=dev)
        model = LbsModel(1, 0, w.data_ptr(), pd.data_ptr(), None, jt.data_ptr(), None, par.data_ptr())
        m = (model, (w, pd, jt, par))
        _RODRIGUES_MODELS[dev] = m
    return m


# ---- stand-alone forms of the fused covariance / colour math ----------------------------------

def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("libghr: tensor arguments must live on a CUDA device (there is no CPU fallback)")


class _CovarianceFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scaling, rotation, scaling_modifier, transforms):
        _need_cuda(scaling, rotation, transforms)
        sc, rot = _f32c(scaling.detach()), _f32c(rotation.detach())
        if rot.data_ptr() % 16:
            rot = rot.clone()
        tr = None if transforms is None else _f32c(transforms.detach()).reshape(-1, 3, 3)
        P = sc.shape[0]
        cov = torch.empty(P, 6, dtype=torch.float32, device=sc.device)
        with _lib.on_device(sc.device):
            check(lib().ghr_covariance_forward(P, sc.data_ptr(), rot.data_ptr(), float(scaling_modifier), _ptr(tr),
                                               cov.data_ptr(), _stream()), "ghr_covariance_forward")
        ctx.save_for_backward(sc, rot, tr)
        ctx.mod, ctx.tr_shape = float(scaling_modifier), None if transforms is None else transforms.shape
        return cov

    @staticmethod
    def backward(ctx, g_cov):
        sc, rot, tr = ctx.saved_tensors
        P = sc.shape[0]
        g = _f32c(g_cov)
        d_sc, d_rot = torch.empty_like(sc), torch.empty_like(rot)
        d_tr = torch.empty_like(tr) if (tr is not None and ctx.needs_input_grad[3]) else None
        with _lib.on_device(sc.device):
            check(lib().ghr_covariance_backward(P, sc.data_ptr(), rot.data_ptr(), ctx.mod, _ptr(tr), g.data_ptr(),
                                                d_sc.data_ptr(), d_rot.data_ptr(), _ptr(d_tr), _stream()),
                  "ghr_covariance_backward")
        return d_sc, d_rot, None, None if d_tr is None else d_tr.view(ctx.tr_shape)


def get_covariance(scaling, rotation, scaling_modifier=1.0, transforms=None):
    """GaussianModel.get_covariance(scaling_modifier, transforms) (scene/gaussian_model.py:35-42,208-209): scaling [P,3]
    = get_scaling (activated), rotation [P,4] = the raw quaternion (normalised inside, general_utils.py:79), transforms
    [P,3,3] or None -> [P,6] packed [xx,xy,xz,yy,yz,zz].  Differentiable in scaling, rotation and transforms."""
    return _CovarianceFn.apply(scaling, rotation, scaling_modifier, transforms)


class _ShColorsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, sh_degree, features, means3D, campos):
        _need_cuda(features, means3D, campos)
        shs, mean = _f32c(features.detach()), _f32c(means3D.detach())
        cam = _f32c(campos.detach()).reshape(-1)
        P, M = shs.shape[0], shs.shape[1]
        colors = torch.empty(P, 3, dtype=torch.float32, device=shs.device)
        clamped = torch.empty(P, dtype=torch.uint8, device=shs.device)
        with _lib.on_device(shs.device):
            check(lib().ghr_sh_colors_forward(P, int(sh_degree), M, shs.data_ptr(), mean.data_ptr(), cam.data_ptr(),
                                              colors.data_ptr(), clamped.data_ptr(), _stream()), "ghr_sh_colors_forward")
        ctx.save_for_backward(shs, mean, cam, clamped)
        ctx.deg = int(sh_degree)
        return colors

    @staticmethod
    def backward(ctx, g_colors):
        shs, mean, cam, clamped = ctx.saved_tensors
        P, M = shs.shape[0], shs.shape[1]
        g = _f32c(g_colors)
        d_shs = torch.empty_like(shs)
        d_mean = torch.empty_like(mean) if ctx.needs_input_grad[2] else None
        with _lib.on_device(shs.device):
            check(lib().ghr_sh_colors_backward(P, ctx.deg, M, shs.data_ptr(), mean.data_ptr(), cam.data_ptr(),
                                               clamped.data_ptr(), g.data_ptr(), d_shs.data_ptr(), _ptr(d_mean), _stream()),
                  "ghr_sh_colors_backward")
        return None, d_shs, d_mean, None


def eval_sh_colors(sh_degree, features, means3D, campos):
    """colors_precomp of render() when pipe.convert_SHs_python is set (gaussian_renderer/__init__.py:124-125,191-195):
    clamp_min(eval_sh(sh_degree, features.transpose(1, 2), normalize(means3D - campos)) + 0.5, 0).  features [P,M,3]
    (pc.get_features), means3D [P,3], campos [3] -> [P,3].  Differentiable in features and means3D."""
    return _ShColorsFn.apply(sh_degree, features, means3D, campos)
