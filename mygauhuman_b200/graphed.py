"""CUDA-graph capture of one full hot-path step (LBS + covariance + SH -> rasteriser forward -> loss ->
rasteriser backward -> LBS backward).

At the ZJU shape the step is ~45 kernel launches of a few microseconds each: issued one by one from Python
the host, not the GPU, sets the pace (measured: 0.85 ms/step of host issue time vs 0.55 ms of GPU work).
The reference has the same problem many times over (~60 ATen launches for LBS alone).  Here the
whole step is captured once into a CUDA graph and replayed per view with one launch:

* the per-view inputs (camera matrices, camera position, SMPL pose/shape/R/Th: 129 floats, and whatever the
  loss needs, e.g. the ground-truth image) live in static device buffers that are overwritten before a replay;
* the rasteriser runs in its sync-free mode with a fixed binning capacity (no host read of num_rendered);
  the largest num_rendered ever seen is tracked ON THE DEVICE inside the graph, so an overflow of that
  capacity (which would silently drop splats) is detected at the next host synchronisation and the graph is
  re-captured with a larger capacity;
* gradients land in the leaves' ``.grad`` tensors, which are static across replays.

Nothing here changes numerics: the captured kernels are the same ones the eager API launches.
"""
import torch

from . import _lib
from . import diff_gaussian_rasterization as dgr
from .diff_gaussian_rasterization import _C

PACK = dict(view=(0, 16), proj=(16, 32), campos=(32, 35), poses=(35, 107), shapes=(107, 117), R=(117, 126), Th=(126, 129))
PACK_FLOATS = 129


def pack_view(world_view_transform, full_proj_transform, camera_center, smpl_param):
    """[view 16 | proj 16 | campos 3 | poses 72 | shapes 10 | R 9 | Th 3] -> one fp32 vector (any device)."""
    v = torch.cat([world_view_transform.reshape(-1), full_proj_transform.reshape(-1), camera_center.reshape(-1),
                   smpl_param["poses"].reshape(-1), smpl_param["shapes"].reshape(-1), smpl_param["R"].reshape(-1),
                   smpl_param["Th"].reshape(-1)]).float()
    assert v.numel() == PACK_FLOATS, "expects 72 pose and 10 shape coefficients"
    return v


def unpack_view(vec):
    g = lambda k: vec[PACK[k][0]:PACK[k][1]]
    cam = dict(view=g("view").view(4, 4), proj=g("proj").view(4, 4), campos=g("campos"))
    pose = dict(poses=g("poses").view(1, 72), shapes=g("shapes").view(1, 10), R=g("R").view(3, 3), Th=g("Th").view(1, 3))
    return cam, pose


class GraphedRenderStep:
    """One training-style step of the hot path as a replayable CUDA graph.

    leaves: dict of leaf tensors (requires_grad) -- xyz [P,3], log_scale [P,3], rotation [P,4], features [P,M,3],
            opacity_logit [P,1], normals [P,3] (optional).  Activations (exp / sigmoid) mirror the reference's
            get_scaling / get_opacity (scene/gaussian_model.py:152-175).
    loss_fn(color, depth, alpha, *loss_inputs) -> scalar tensor; it is captured too, so it must be sync-free.
    """

    def __init__(self, deformer, leaves, big_pose_params, t_vertices, image_height, image_width, tanfovx, tanfovy, bg,
                 loss_fn, n_loss_inputs=1, sh_degree=3, scaling_modifier=1.0, capacity_hint=None, use_graph=True,
                 all_reduce=None, group=None, accumulate=False, shared_flat=None, fused=True):
        """all_reduce: None (single GPU), "after" (one NCCL sum all-reduce of the flat gradient bucket issued right
        after the replay) or "captured" (the all-reduce is part of the graph: every rank replays in lockstep).
        accumulate=True: multi-view batches.  Every call ADDS its gradients to the flat bucket (leaves[k].grad are
        views of it); zero_grads() starts a batch and reduce_grads() sums the bucket over the ranks once per batch
        (all_reduce must then be None: the collective is per batch, not per view).
        shared_flat (accumulate mode): another step's flat bucket to add into, with atomics -- the lanes of a
        MultiLaneRenderStep run concurrently on several streams and sum into one bucket.
        fused=True: the per-Gaussian chain runs as one kernel each way (fused.render_deformed: LBS + covariance + SH +
        projection, and their backward) instead of SMPLDeformer.deform_gaussians -> GaussianRasterizer."""
        self.fused = bool(fused)
        self.deformer, self.leaves = deformer, leaves
        self.big, self.t_vertices = big_pose_params, t_vertices
        self.H, self.W, self.tanfovx, self.tanfovy = int(image_height), int(image_width), float(tanfovx), float(tanfovy)
        self.bg, self.loss_fn, self.sh_degree, self.mod = bg, loss_fn, int(sh_degree), float(scaling_modifier)
        self.use_graph = use_graph
        dev = deformer.device
        self.vec = torch.zeros(PACK_FLOATS, device=dev)
        self.loss_inputs = None
        self.n_loss_inputs = n_loss_inputs
        self.n_seen = torch.zeros(1, dtype=torch.int32, device=dev)  # max num_rendered over all replays (device)
        self.capacity = None if capacity_hint is None else int(capacity_hint)
        assert all_reduce in (None, "after", "captured")
        assert not (accumulate and all_reduce is not None)
        self.all_reduce, self.group, self.accumulate = all_reduce, group, accumulate
        self.flat, self.flat_views = None, None
        if all_reduce is not None or accumulate:
            # gradients of all leaves in one bucket (59 floats per Gaussian at SH degree 3): one collective per step,
            # and leaves[k].grad are views of it, so nothing is copied back after the reduction
            pad = lambda n: (n + 3) // 4 * 4  # every view starts on a 16-byte boundary
            n_flat = sum(pad(t.numel()) for t in leaves.values())
            assert shared_flat is None or (accumulate and shared_flat.numel() == n_flat)
            self.flat = torch.zeros(n_flat, device=dev) if shared_flat is None else shared_flat
            self.flat_views, o = {}, 0
            for k, t in leaves.items():
                self.flat_views[k] = self.flat[o:o + t.numel()].view_as(t)
                o += pad(t.numel())
        self.vert_ids = None  # static int32 [P]: nearest-vertex ids supplied by the owner (MultiLaneRenderStep)
        # (pose buffer, vertex offsets): static buffers the step reads the target pose from instead of running the pose
        # stage itself.  A MultiLaneRenderStep fills them for all its lanes with ONE launch per group (pose_views) and sets
        # _pose_ready; a direct call of this step fills them itself.
        self.pose_inputs, self._pose_ready = None, False
        self._pre_run = None  # owner's hook: makes the current stream wait for a nearest-vertex search still in flight
        self._sink, self._sink_atomic = None, False
        if accumulate:
            # the deform backward kernel adds straight into the bucket: no separate accumulation pass
            names = {"xyz": "xyz", "normals": "normals", "log_scale": "scaling", "rotation": "rotation",
                     "features": "features", "opacity_logit": "opacity_logit"}
            self._sink = {names[k]: v for k, v in self.flat_views.items()}
            for k, t in leaves.items():
                t.grad = self.flat_views[k]
        self._copy_stream, self._staged, self._loss_ring = None, None, None
        self.graph = None
        self.kernels_per_replay = None
        self.loss = None
        self.out = None

    # ---- the eager step (also what gets captured) -------------------------------------------------
    def _fill_pose(self):
        """The pose stage of the view in self.vec into the static pose buffers (one launch)."""
        self.deformer.pose_views(self.big, [_pose_slices(self.vec)], [self.pose_inputs])

    def _eager(self, vec, loss_inputs, capacity, reduce=False, static_pose=False):
        """static_pose: read the target pose from self.pose_inputs (filled beforehand) instead of computing it from vec."""
        self.deformer.pose_override = self.pose_inputs if static_pose else None
        try:
            return self._eager_impl(vec, loss_inputs, capacity, reduce)
        finally:
            self.deformer.pose_override = None

    def _eager_impl(self, vec, loss_inputs, capacity, reduce):
        L = self.leaves
        camd, pose = unpack_view(vec)
        if self.fused:
            return self._eager_fused(camd, pose, loss_inputs, capacity, reduce)
        # exp / sigmoid of the raw parameters are fused into the deform kernels (no separate ATen launches)
        self.deformer.vert_ids_override = self.vert_ids
        try:
            d = self.deformer.deform_gaussians(L["xyz"], pose, self.big, self.t_vertices, normals=L.get("normals"),
                                               log_scaling=L["log_scale"], rotation=L["rotation"],
                                               scaling_modifier=self.mod, features=L["features"],
                                               sh_degree=self.sh_degree, campos=camd["campos"],
                                               opacity_logit=L["opacity_logit"])
        finally:
            self.deformer.vert_ids_override = None
        rs = dgr.GaussianRasterizationSettings(
            image_height=self.H, image_width=self.W, tanfovx=self.tanfovx, tanfovy=self.tanfovy, bg=self.bg,
            scale_modifier=self.mod, viewmatrix=camd["view"], projmatrix=camd["proj"], sh_degree=self.sh_degree,
            campos=camd["campos"], prefiltered=False, debug=False)
        means2D = torch.zeros_like(d["means3D"], requires_grad=True)
        _C.set_sync_free_capacity(capacity)
        try:
            color, radii, depth, alpha = dgr.rasterize_gaussians(
                d["means3D"], means2D, _C_absent(), d["colors"], d["opacity"], _C_absent(),
                _C_absent(), d["cov3D"], rs)
        finally:
            _C.set_sync_free_capacity(None)
        if capacity is not None:
            geom = _C._last_geom
            torch.maximum(self.n_seen, geom[:4].view(torch.int32), out=self.n_seen)
        loss = self.loss_fn(color, depth, alpha, *loss_inputs)
        if self._sink is not None:
            self.deformer.grad_sink, self.deformer.grad_sink_atomic = self._sink, self._sink_atomic
            try:
                loss.backward()
            finally:
                self.deformer.grad_sink, self.deformer.grad_sink_atomic = None, False
            return loss, dict(color=color, depth=depth, alpha=alpha, radii=radii, means2D=means2D)
        for t in L.values():
            t.grad = None
        loss.backward()
        if self.flat is not None:
            with torch.no_grad():
                torch._foreach_copy_([self.flat_views[k] for k in L], [L[k].grad for k in L])  # one multi-tensor kernel
                if reduce:
                    self._reduce()
            for k in L:
                L[k].grad = self.flat_views[k]
        return loss, dict(color=color, depth=depth, alpha=alpha, radii=radii, means2D=means2D)

    def _eager_fused(self, camd, pose, loss_inputs, capacity, reduce):
        from .fused import render_deformed
        L = self.leaves
        rs = dgr.GaussianRasterizationSettings(
            image_height=self.H, image_width=self.W, tanfovx=self.tanfovx, tanfovy=self.tanfovy, bg=self.bg,
            scale_modifier=self.mod, viewmatrix=camd["view"], projmatrix=camd["proj"], sh_degree=self.sh_degree,
            campos=camd["campos"], prefiltered=False, debug=False)
        means2D = torch.zeros_like(L["xyz"], requires_grad=True)
        self.deformer.vert_ids_override = self.vert_ids
        _C.set_sync_free_capacity(capacity)
        try:
            color, radii, depth, alpha = render_deformed(
                self.deformer, L["xyz"], pose, self.big, self.t_vertices, rs, log_scaling=L["log_scale"],
                rotation=L["rotation"], features=L["features"], opacity_logit=L["opacity_logit"], means2D=means2D)
        finally:
            _C.set_sync_free_capacity(None)
            self.deformer.vert_ids_override = None
        if capacity is not None:
            geom = _C._last_geom
            torch.maximum(self.n_seen, geom[:4].view(torch.int32), out=self.n_seen)
        loss = self.loss_fn(color, depth, alpha, *loss_inputs)
        out = dict(color=color, depth=depth, alpha=alpha, radii=radii, means2D=means2D)
        if self._sink is not None:
            self.deformer.grad_sink, self.deformer.grad_sink_atomic = self._sink, self._sink_atomic
            try:
                loss.backward()
            finally:
                self.deformer.grad_sink, self.deformer.grad_sink_atomic = None, False
            return loss, out
        for t in L.values():
            t.grad = None
        loss.backward()
        if "normals" in L and L["normals"].grad is None:  # normals do not reach the images on this path
            L["normals"].grad = torch.zeros_like(L["normals"])
        if self.flat is not None:
            with torch.no_grad():
                torch._foreach_copy_([self.flat_views[k] for k in L], [L[k].grad for k in L])  # one multi-tensor kernel
                if reduce:
                    self._reduce()
            for k in L:
                L[k].grad = self.flat_views[k]
        return loss, out

    def zero_grads(self):
        """accumulate mode: start a batch of views."""
        self.flat.zero_()

    def reduce_grads(self):
        """accumulate mode: end of a batch -- one sum all-reduce of the bucket over the ranks (no-op on one rank)."""
        self._reduce()
        return self.flat_views

    def _max_over_ranks(self, n):
        import torch.distributed as dist
        if (self.all_reduce is None and not self.accumulate) or not (dist.is_available() and dist.is_initialized()):
            return n
        t = torch.tensor([n], dtype=torch.int64, device=self.vec.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return int(t.item())

    def _reduce(self):
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)

    # ---- capture ------------------------------------------------------------------------------------
    def _measure(self, vec, loss_inputs):
        """Exact (blocking) eager step: returns num_rendered of this view."""
        self._eager(vec, loss_inputs, None)
        torch.cuda.synchronize()
        return _C.read_status(_C._last_geom)[0]

    def capture(self, sample_vecs, loss_inputs):
        """Size the binning capacity from `sample_vecs` (a list of packed views), then capture."""
        self.loss_inputs = [t.clone() for t in loss_inputs]
        n_max = self._max_over_ranks(max(self._measure(v, self.loss_inputs) for v in sample_vecs))
        cap = max(int(n_max * 1.5) + 4096, self.capacity or 0)
        self._capture(cap)

    def _capture(self, capacity):
        self.capacity = _C._round_cap(capacity)
        self.n_seen.zero_()
        if not self.use_graph:
            return
        in_graph = self.all_reduce == "captured"
        sp = self.pose_inputs is not None
        if sp:
            self._fill_pose()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):  # warm-up on a side stream, as torch.cuda.graph requires
                self._eager(self.vec, self.loss_inputs, self.capacity, reduce=in_graph, static_pose=sp)
        torch.cuda.current_stream().wait_stream(s)
        for k, t in self.leaves.items():
            t.grad = None if self._sink is None else self.flat_views[k]
        self.graph = torch.cuda.CUDAGraph()
        n0 = _lib.launch_count()
        with torch.cuda.graph(self.graph):
            self.loss, self.out = self._eager(self.vec, self.loss_inputs, self.capacity, reduce=in_graph, static_pose=sp)
        self.kernels_per_replay = _lib.launch_count() - n0  # libghr kernels inside one replay
        self.n_seen.zero_()

    # ---- replay ---------------------------------------------------------------------------------------
    def __call__(self, vec, *loss_inputs):
        """Run the step for the view packed in `vec` (device or pinned-host tensor).  Returns the loss tensor
        (static; overwritten by the next call).  Gradients are in leaves[k].grad."""
        self._load_inputs(vec, loss_inputs)
        return self._run()

    def _load_inputs(self, vec, loss_inputs):
        if self.capacity is None:
            raise RuntimeError("call capture() first")
        self.vec.copy_(vec, non_blocking=True)
        for dst, src in zip(self.loss_inputs, loss_inputs):
            if src is not dst:
                dst.copy_(src, non_blocking=True)

    def _run(self):
        if self._pre_run is not None:
            self._pre_run()
        if self.pose_inputs is not None and not self._pose_ready:
            self._fill_pose()
        if self.graph is None:
            self.loss, self.out = self._eager(self.vec, self.loss_inputs, self.capacity, reduce=self.all_reduce is not None,
                                              static_pose=self.pose_inputs is not None)
        else:
            self.graph.replay()
            if self.all_reduce == "after":
                self._reduce()
        return self.loss

    # ---- host-side pipelining ---------------------------------------------------------------------------
    def prefetch(self, vec, *loss_inputs):
        """Start copying the NEXT view's inputs (pinned host or device tensors) into staging buffers on a copy stream.
        The following run_prefetched() waits for that copy only, so the transfer of view i+1 overlaps the kernels of
        view i instead of preceding its own (a 512x512 target with 5 planes is 5 MB: ~100 us over PCIe)."""
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.vec.device)
            self._staged = [torch.empty_like(self.vec)] + [torch.empty_like(t) for t in self.loss_inputs]
            self._staged_ready = torch.cuda.Event()
            self._staged_free = torch.cuda.Event()
            self._staged_free.record()
        cs = self._copy_stream
        cs.wait_event(self._staged_free)  # the previous run_prefetched() has consumed the staging buffers
        with torch.cuda.stream(cs):
            self._staged[0].copy_(vec, non_blocking=True)
            for dst, src in zip(self._staged[1:], loss_inputs):
                dst.copy_(src, non_blocking=True)
            self._staged_ready.record()

    def run_prefetched(self):
        """Run the step on the inputs handed to the last prefetch()."""
        cur = torch.cuda.current_stream()
        cur.wait_event(self._staged_ready)
        self._load_inputs(self._staged[0], self._staged[1:])  # device-to-device into the graph's static inputs
        self._staged_free.record()                            # the copy stream may refill the staging buffers now
        return self._run()

    def loss_to_host(self):
        """Asynchronous device-to-host copy of the last step's loss into a pinned ring; returns (slot tensor, event).
        Read the value after event.synchronize() -- typically one step later, so the host never stalls the GPU."""
        if self._loss_ring is None:
            self._loss_ring = [(torch.zeros(1, dtype=torch.float32).pin_memory(), torch.cuda.Event()) for _ in range(4)]
            self._loss_i = 0
        slot, ev = self._loss_ring[self._loss_i % 4]
        self._loss_i += 1
        slot.copy_(self.loss.detach().reshape(1), non_blocking=True)
        ev.record()
        return slot, ev

    def check(self):
        """Synchronises; if some replay needed more binning capacity than the graph was captured with, re-captures
        with a larger capacity and returns False (the caller should redo the affected steps)."""
        n = self._max_over_ranks(int(self.n_seen.item()))  # every rank re-captures together, with the same capacity
        if n > self.capacity:
            self._capture(int(n * 1.5) + 4096)
            return False
        return True


class MultiLaneRenderStep:
    """K graphed lanes over the SAME leaves: the views of a multi-view batch run concurrently on K CUDA streams.

    One view of the ZJU shape is ~20 kernels that are each bound by latency or by the instruction stream of a few heavy
    tiles, not by the machine: one replay keeps a fraction of the 148 SMs busy.  Views of a batch are
    independent until their parameter gradients are summed, so lane k owns a GraphedRenderStep (own deformer scratch,
    own static inputs) and a stream, and all lanes add into ONE flat gradient bucket with fp32 atomics
    (``accumulate_grads = 2`` of ghr_lbs_deform_backward: a bucket per lane cost K-1 extra clears and adds of 11.8 MB
    per batch, 11 % of an 8-view batch).  ``run_group`` replays up to K views at once; ``reduce_grads`` is the one
    NCCL all-reduce of the batch.  Measured on B200 (ZJU shape, tools/overlap_probe.py): 1 / 2 / 3 / 4 lanes =
    2 337 / 3 543 / 3 899 / 4 105 views/s.  Numerics are those of the single-lane step; only the order in which the
    views are ADDED into the batch gradient changes.
    """

    def __init__(self, deformers, leaves, big_pose_params, t_vertices, image_height, image_width, tanfovx, tanfovy, bg,
                 loss_fn, n_loss_inputs=1, sh_degree=3, scaling_modifier=1.0, use_graph=True, group=None, fused=True):
        """deformers: one SMPLDeformer per lane (they hold per-call scratch the lanes must not share)."""
        assert len(deformers) >= 1
        self.leaves, self.group = leaves, group
        if len(deformers) > 1:
            # kernels of different views now share the SMs: the blend kernels take the configuration that lets two of
            # them be co-resident (read at launch, i.e. baked into the lanes' graphs at capture): 4 300 -> 5 150 views/s
            _lib.check(_lib.lib().ghr_set_shared_gpu(1), "ghr_set_shared_gpu")
        self.lanes = []
        for d in deformers:
            self.lanes.append(GraphedRenderStep(d, leaves, big_pose_params, t_vertices, image_height, image_width, tanfovx,
                                                tanfovy, bg, loss_fn, n_loss_inputs=n_loss_inputs, sh_degree=sh_degree,
                                                scaling_modifier=scaling_modifier, use_graph=use_graph, group=group,
                                                accumulate=True, fused=fused,
                                                shared_flat=self.lanes[0].flat if self.lanes else None))
        # The nearest SMPL vertex of a Gaussian depends on the canonical positions only, and those do not change inside a
        # batch: the search runs once per batch (zero_grads) instead of once per view (39 us of a 420 us view).
        self.vert_ids = torch.zeros(leaves["xyz"].shape[0], dtype=torch.int32, device=deformers[0].device)
        dev = deformers[0].device
        # The pose stage (joint chain + pose blend-shape offsets: 17 MB of pose basis per view) of all the views of a
        # group runs as ONE launch before the lanes fork, into per-lane static buffers the lanes' graphs read.
        self.batch_pose = len(self.lanes) > 1
        for lane in self.lanes:
            lane._sink_atomic = len(self.lanes) > 1
            lane.vert_ids = self.vert_ids
            lane._pre_run = self._wait_knn
            if self.batch_pose:
                lane.pose_inputs = (torch.zeros(_lib.POSE_FLOATS, device=dev), torch.zeros(deformers[0].V, 3, device=dev))
        self.streams = [torch.cuda.Stream(device=dev) for _ in self.lanes]
        self._side = torch.cuda.Stream(device=dev)   # nearest-vertex search of a batch, next to the bucket reset and the pose stage
        self._knn_done, self._knn_pending = torch.cuda.Event(), False
        self._fork = torch.cuda.Event()
        self._joins = [torch.cuda.Event() for _ in self.lanes]
        self._alias_grads()

    K = property(lambda self: len(self.lanes))
    flat = property(lambda self: self.lanes[0].flat)
    flat_views = property(lambda self: self.lanes[0].flat_views)
    kernels_per_replay = property(lambda self: self.lanes[0].kernels_per_replay)
    capacity = property(lambda self: self.lanes[0].capacity)

    def _alias_grads(self):
        for k, t in self.leaves.items():
            t.grad = self.lanes[0].flat_views[k]

    def capture(self, sample_vecs, loss_inputs):
        """Lane 0 sizes the binning capacity from the sample views; every lane is captured with it."""
        self._nearest_vertices()
        self.lanes[0].capture(sample_vecs, loss_inputs)
        for lane in self.lanes[1:]:
            lane.loss_inputs = [t.clone() for t in loss_inputs]
            lane._capture(self.lanes[0].capacity)
        self._alias_grads()

    def _nearest_vertices(self):
        lane = self.lanes[0]
        lane.deformer.nearest_vertex(lane.t_vertices, self.leaves["xyz"], out=self.vert_ids)

    def _wait_knn(self):
        if self._knn_pending:
            torch.cuda.current_stream().wait_event(self._knn_done)
            self._knn_pending = False

    def zero_grads(self):
        """Start a batch (on the current stream; the lanes fork from it): clears the gradient bucket and refreshes the
        nearest-vertex ids from the current positions.  Call it after every parameter update."""
        cur = torch.cuda.current_stream()
        self._side.wait_stream(cur)  # the positions the optimiser just wrote; the previous batch's lanes have joined `cur`
        with torch.cuda.stream(self._side):
            self._nearest_vertices()
            self._knn_done.record(self._side)
        self._knn_pending = True
        self.lanes[0].flat.zero_()

    # (Tried: ONE CUDA graph for a whole group -- the pose stage and all K lanes captured across their streams, one host
    # launch instead of K.  Measured slower on B200, 7 040 vs 7 180 views/s at the ZJU shape: the K per-lane launches cost
    # ~120 us of host time per group, but the lanes they start staggered overlap better than K branches of one graph.)
    def run_group(self, vecs, loss_inputs=None, prefetched=False, next_vecs=None, next_loss_inputs=None):
        """Replay len(vecs) <= K views concurrently, view j on lane j.  loss_inputs: one tuple per view (None: keep the
        lanes' resident loss inputs).  prefetched=True: run what each lane's last prefetch() staged instead of `vecs`
        (only its length is used); next_vecs / next_loss_inputs: what to prefetch per lane right after issuing (the host
        copies of the next group then overlap this group's kernels).  Returns the lanes' loss tensors (static buffers)."""
        cur = torch.cuda.current_stream()
        if self.batch_pose:
            srcs = []
            for j in range(len(vecs)):
                lane = self.lanes[j]
                if prefetched:
                    cur.wait_event(lane._staged_ready)
                    srcs.append(lane._staged[0])
                elif vecs[j].is_cuda:
                    srcs.append(vecs[j])
                else:
                    lane.vec.copy_(vecs[j], non_blocking=True)
                    srcs.append(lane.vec)
            self.lanes[0].deformer.pose_views(self.lanes[0].big, [_pose_slices(v) for v in srcs],
                                              [self.lanes[j].pose_inputs for j in range(len(vecs))])
        self._wait_knn()
        self._fork.record(cur)
        losses = []
        for j in range(len(vecs)):
            lane, s = self.lanes[j], self.streams[j]
            s.wait_event(self._fork)
            with torch.cuda.stream(s):
                lane._pose_ready = self.batch_pose
                try:
                    if prefetched:
                        losses.append(lane.run_prefetched())
                    else:
                        li = lane.loss_inputs if loss_inputs is None else loss_inputs[j]
                        losses.append(lane(vecs[j], *li))
                finally:
                    lane._pose_ready = False
                self._joins[j].record(s)
        # the host copies of the NEXT group are issued only after every lane of this group has been launched: issuing them
        # lane by lane in between delayed the last lane's replay by ~130 us of host calls (9 % of an 8-view group)
        if next_vecs is not None:
            for j in range(min(len(vecs), len(next_vecs))):
                self.lanes[j].prefetch(next_vecs[j], *(next_loss_inputs[j] if next_loss_inputs is not None else ()))
        for j in range(len(vecs)):
            cur.wait_event(self._joins[j])
        return losses

    def reduce_grads(self):
        """End of a batch: one sum all-reduce of the shared bucket over the ranks (call after run_group has joined)."""
        self.lanes[0]._reduce()
        return self.lanes[0].flat_views

    def check(self):
        """Synchronises; re-captures every lane with a larger capacity if any replay overflowed (returns False then)."""
        n = max(int(lane.n_seen.item()) for lane in self.lanes)
        n = self.lanes[0]._max_over_ranks(n)
        if n > self.lanes[0].capacity:
            for lane in self.lanes:
                lane._capture(int(n * 1.5) + 4096)
            self._alias_grads()
            return False
        return True


def _pose_slices(vec):
    """The SMPL parameters of a packed view vector as contiguous slices (inputs of SMPLDeformer.pose_views)."""
    return {k: vec[PACK[k][0]:PACK[k][1]] for k in ("poses", "shapes", "R", "Th")}


def _C_absent():
    return torch.Tensor([])
