"""Optimiser side of the training iteration on libghr.

``FusedAdam`` is a drop-in for the ``torch.optim.Adam(l, lr=0.0, eps=1e-15)`` the reference builds in
GaussianModel.training_setup (scene/gaussian_model.py:250-283): same constructor, same ``param_groups`` (so
update_learning_rate, :290-295, keeps working), same ``state[param] = {"step", "exp_avg", "exp_avg_sq"}`` entries (so
the densification code that cuts / extends / re-keys them, :423-483, keeps working) -- but ``step()`` is ONE kernel
launch over all parameter tensors (``ghr_adam_step``) instead of ~10 multi-tensor ATen launches, and it can clear the
gradients in the same pass (``step(zero_grads=True)`` = optimizer.step(); optimizer.zero_grad(), train.py:418-419,
without freeing the gradient tensors, which is what a flat gradient bucket / CUDA-graph replay needs).

``add_densification_stats`` fuses train.py:401-404 and GaussianModel.add_densification_stats (:764-766).

No CPU fallback: CUDA fp32 tensors only.
"""
import ctypes as C

import torch

from . import _lib


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        if not 0.0 <= lr:
            raise ValueError(f"Invalid learning rate: {lr}")
        if not 0.0 <= eps:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 0: {betas[0]}")
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 1: {betas[1]}")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @torch.no_grad()
    def step(self, closure=None, zero_grads=False):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # one launch per (betas, eps, device) bucket of <= ADAM_MAX_GROUPS tensors: one launch for the reference's set-up
        buckets = {}
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or p.grad.dtype != torch.float32:
                    raise ValueError("FusedAdam: CUDA fp32 parameters and gradients only (there is no CPU fallback)")
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                if not p.is_contiguous():
                    raise ValueError("FusedAdam: parameters must be contiguous")
                grad = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)  # host scalar, like torch's default
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                if not (st["exp_avg"].is_contiguous() and st["exp_avg_sq"].is_contiguous()):
                    st["exp_avg"], st["exp_avg_sq"] = st["exp_avg"].contiguous(), st["exp_avg_sq"].contiguous()
                key = (float(beta1), float(beta2), float(group["eps"]), p.device.index or 0)
                buckets.setdefault(key, []).append((p, grad, st, float(group["lr"])))
        l = _lib.lib()
        for (beta1, beta2, eps, dev), items in buckets.items():
            with _lib.on_device(torch.device("cuda", dev)):
                stream = _lib.raw_stream(dev)
                for lo in range(0, len(items), _lib.ADAM_MAX_GROUPS):
                    chunk = items[lo:lo + _lib.ADAM_MAX_GROUPS]
                    table = (_lib.AdamGroup * len(chunk))()
                    for e, (p, grad, st, lr) in zip(table, chunk):
                        e.param, e.grad = p.data_ptr(), grad.data_ptr()
                        e.exp_avg, e.exp_avg_sq = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                        e.count, e.lr, e.step = p.numel(), lr, int(st["step"])
                    _lib.check(l.ghr_adam_step(len(chunk), table, beta1, beta2, eps, int(bool(zero_grads)), stream),
                               "ghr_adam_step")
                    if zero_grads:
                        for p, grad, _, _ in chunk:
                            if grad is not p.grad:  # a contiguous copy was stepped on: clear the real one
                                p.grad.zero_()
        return loss


def add_densification_stats(radii, viewspace_grad, max_radii2D=None, xyz_gradient_accum=None, denom=None):
    """In place, for visibility_filter = radii > 0 (gaussian_renderer/__init__.py:287):
    max_radii2D = max(max_radii2D, radii) (train.py:403); xyz_gradient_accum += ||viewspace_grad[:, :2]||;
    denom += 1 (scene/gaussian_model.py:764-766).  radii int32 [P]; viewspace_grad [P,>=2]; max_radii2D [P];
    xyz_gradient_accum / denom [P,1] or [P]."""
    P = radii.shape[0]
    if not radii.is_cuda or radii.dtype != torch.int32 or not radii.is_contiguous():
        raise ValueError("radii must be a contiguous CUDA int32 tensor")
    if viewspace_grad.dtype != torch.float32 or viewspace_grad.dim() != 2 or viewspace_grad.shape[0] != P \
            or viewspace_grad.shape[1] < 2 or not viewspace_grad.is_contiguous() or viewspace_grad.device != radii.device:
        raise ValueError("viewspace_grad must be a contiguous CUDA fp32 [P,>=2] tensor")
    for t in (max_radii2D, xyz_gradient_accum, denom):
        if t is not None and (t.dtype != torch.float32 or t.numel() != P or not t.is_contiguous()
                              or t.device != radii.device):
            raise ValueError("statistics must be contiguous CUDA fp32 tensors with one entry per Gaussian")
    ptr = lambda t: None if t is None else t.data_ptr()
    dev = radii.device
    with _lib.on_device(dev):
        _lib.check(_lib.lib().ghr_densify_stats(P, radii.data_ptr(), viewspace_grad.data_ptr(), viewspace_grad.shape[1],
                                                ptr(max_radii2D), ptr(xyz_gradient_accum), ptr(denom),
                                                _lib.raw_stream(dev.index or 0)), "ghr_densify_stats")


# ---- view-sharded data parallelism: reduce-scatter + Adam + all-gather fused over NVLink peer memory ---------------------

def flat_layout(leaves):
    """The flat layout GraphedRenderStep / MultiLaneRenderStep give the gradient bucket: the leaves in dict order, each
    starting on a 4-float (16-byte) boundary.  -> (total floats, {name: (offset, count)})."""
    pad = lambda n: (n + 3) // 4 * 4
    table, o = {}, 0
    for k, t in leaves.items():
        table[k] = (o, t.numel())
        o += pad(t.numel())
    return o, table


def flatten_leaves(leaves):
    """Re-home the leaves in ONE flat fp32 buffer with that layout (what the sharded step needs: a parameter is addressed
    by its offset on every rank).  -> (flat_params, {name: leaf tensor viewing it, requires_grad})."""
    total, table = flat_layout(leaves)
    any_leaf = next(iter(leaves.values()))
    flat = torch.zeros(total, dtype=torch.float32, device=any_leaf.device)
    out = {}
    for k, t in leaves.items():
        o, n = table[k]
        flat[o:o + n].copy_(t.detach().reshape(-1))
        out[k] = flat[o:o + n].view(t.shape).requires_grad_(True)  # a view of a buffer that does not require grad: a leaf
    return flat, out


def shard_bounds(total, world, rank):
    """Slice [lo, hi) of a `total`-float buffer owned by `rank`: equal 4-float-aligned slices, the last one ragged."""
    per = ((total + world - 1) // world + 3) // 4 * 4
    lo = min(rank * per, total)
    return lo, min(lo + per, total)


_ipc_bases = {}  # cudaIpcMemHandle bytes -> base address of the mapping in this process (one open per handle)


class ShardedFusedAdam:
    """Adam over a flat parameter replica per rank, fused with the gradient reduction (ghr_adam_step_sharded): every rank
    sums ITS slice of all ranks' gradient buckets straight from peer memory, steps it with its shard of the moments
    (ZeRO-1) and writes the new parameters into every replica -- one kernel instead of NCCL all-reduce + Adam, with
    bit-identical replicas.  Single node (CUDA IPC over NVLink), <= 8 ranks, torch.distributed initialised with NCCL.

    groups: list of dict(name=, offset=, count=, lr=) in the flat layout (flat_layout); lr may be changed between steps
    like a torch param_group (update_learning_rate, scene/gaussian_model.py:290-295)."""

    def __init__(self, flat_params, flat_grads, groups, betas=(0.9, 0.999), eps=1e-8, group=None, sync="flags"):
        """sync: "flags" = the two cross-rank barriers of a step are flag exchanges in peer memory INSIDE the kernel (no
        NCCL call in step()); "nccl" = one-element NCCL all-reduces around the kernel."""
        import torch.distributed as dist
        if sync not in ("flags", "nccl"):
            raise ValueError("sync must be 'flags' or 'nccl'")
        self.sync = sync
        if not (flat_params.is_cuda and flat_grads.is_cuda and flat_params.dtype == torch.float32
                and flat_grads.dtype == torch.float32 and flat_params.numel() == flat_grads.numel()
                and flat_params.is_contiguous() and flat_grads.is_contiguous() and flat_params.numel() % 4 == 0):
            raise ValueError("ShardedFusedAdam: flat contiguous CUDA fp32 buffers of equal, 4-float-padded size")
        self.dist, self.group = dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > _lib.MAX_PEERS:
            raise ValueError(f"ShardedFusedAdam: at most {_lib.MAX_PEERS} ranks (one node)")
        self.params, self.grads = flat_params, flat_grads
        self.param_groups = [dict(g) for g in groups]
        self.betas, self.eps, self.steps = betas, eps, 0
        self.lo, self.hi = shard_bounds(flat_params.numel(), self.world, self.rank)
        dev = flat_params.device
        self.exp_avg = torch.zeros(max(self.hi - self.lo, 4), device=dev)
        self.exp_avg_sq = torch.zeros(max(self.hi - self.lo, 4), device=dev)
        # Exchange the buffers' CUDA IPC handles and map the peers' allocations under OUR device (a mapping opened under
        # the owner's device, as torch's own rebuild does, is not reachable from kernels running here).
        def describe(t):
            st = t.untyped_storage()
            _, handle, _, storage_off, *_ = st._share_cuda_()
            handle = bytes(handle)
            if len(handle) in (65, 66):  # newer torch prefixes the 64-byte cudaIpcMemHandle_t with (version,) kind bytes
                if handle[-65:-64] != b"c":
                    raise RuntimeError("ShardedFusedAdam: the buffers must come from cudaMalloc segments (expandable "
                                       "segments are not supported: unset PYTORCH_CUDA_ALLOC_CONF=expandable_segments)")
                handle = handle[-64:]
            if len(handle) != 64:
                raise RuntimeError(f"ShardedFusedAdam: unexpected CUDA IPC handle of {len(handle)} bytes (expandable "
                                   "segments are not supported: unset PYTORCH_CUDA_ALLOC_CONF=expandable_segments)")
            return handle, int(storage_off) + 4 * int(t.storage_offset())
        self._flags = torch.zeros(_lib.lib().ghr_sharded_flag_words(), dtype=torch.int32, device=dev)
        self._shared = [flat_params.detach(), flat_grads.detach(), self._flags]  # producers keep what they shared alive
        mine = [describe(t) for t in self._shared]
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=group)
        p_ptrs, g_ptrs, f_ptrs = [], [], []
        with _lib.on_device(dev):
            for r in range(self.world):
                if r == self.rank:
                    p_ptrs.append(flat_params.data_ptr())
                    g_ptrs.append(flat_grads.data_ptr())
                    f_ptrs.append(self._flags.data_ptr())
                    continue
                ptrs = []
                for handle, off in everyone[r]:
                    base = _ipc_bases.get(handle)
                    if base is None:
                        out = C.c_void_p()
                        _lib.check(_lib.lib().ghr_ipc_open(handle, C.byref(out)), "ghr_ipc_open")
                        base = _ipc_bases[handle] = int(out.value)
                    ptrs.append(base + off)
                p_ptrs.append(ptrs[0])
                g_ptrs.append(ptrs[1])
                f_ptrs.append(ptrs[2])
        self._p_ptrs = (C.c_void_p * self.world)(*p_ptrs)
        self._g_ptrs = (C.c_void_p * self.world)(*g_ptrs)
        self._f_ptrs = (C.c_void_p * self.world)(*f_ptrs)
        self._token = torch.zeros(1, device=dev)
        dist.barrier(group=group)
        self.check()  # every rank starts at the same epoch, no error latched

    def check(self):
        """Synchronises this rank's device and verifies the in-kernel barriers' state: raises if a wait gave up (a peer
        never arrived within GHR_SHARDED_TIMEOUT_MS, or some rank aborted) or if the ranks disagree on the step count
        (the epoch of the flag exchange).  Collective: every rank must call it.  Cheap enough for once per N steps."""
        err_word = _lib.lib().ghr_sharded_error_word()
        state = torch.tensor([self.steps, -self.steps, int(self._flags[err_word].item())], dtype=torch.int64,
                             device=self.params.device)
        self.dist.all_reduce(state, op=self.dist.ReduceOp.MAX, group=self.group)
        hi, neg_lo, err = (int(x) for x in state.tolist())
        if err:
            why = {1: "timed out waiting for a peer's gradient bucket", 2: "timed out waiting for a peer's replica write",
                   3: "aborted by a peer"}.get(err, f"error {err}")
            raise RuntimeError(f"ShardedFusedAdam: an in-kernel barrier gave up at step {hi} ({why}); parameters of the "
                               "ranks may differ -- restore from a checkpoint")
        if hi != -neg_lo:
            raise RuntimeError(f"ShardedFusedAdam: ranks disagree on the step count ({-neg_lo}..{hi})")

    def abort(self):
        """Raise the abort word of every rank: peers waiting inside a step return at once (and report it in check())."""
        dev = self.params.device
        with _lib.on_device(dev):
            _lib.check(_lib.lib().ghr_sharded_abort(self.world, self._f_ptrs, _lib.raw_stream(dev.index)),
                       "ghr_sharded_abort")

    def _barrier_on_stream(self):
        # a one-element NCCL all-reduce: completes on this stream only when every rank's stream has reached it
        self.dist.all_reduce(self._token, group=self.group)

    @torch.no_grad()
    def step(self):
        """Call where the reference calls optimizer.step() after the batch's backward passes; NO gradient all-reduce
        before it.  On return (stream order) every replica holds the new parameters; clear the bucket afterwards
        (zero_grads of the render step)."""
        self.steps += 1
        table = (_lib.ShardGroup * len(self.param_groups))()
        for e, g in zip(table, self.param_groups):
            e.offset, e.count, e.lr, e.step = int(g["offset"]), int(g["count"]), float(g["lr"]), self.steps
        dev = self.params.device
        if self.sync == "flags":
            with _lib.on_device(dev):
                try:
                    _lib.check(_lib.lib().ghr_adam_step_sharded_synced(
                        self.world, self.rank, self._g_ptrs, self._p_ptrs, self.exp_avg.data_ptr(),
                        self.exp_avg_sq.data_ptr(), self.lo, self.hi, len(self.param_groups), table, float(self.betas[0]),
                        float(self.betas[1]), float(self.eps), self._f_ptrs, self.steps, _lib.raw_stream(dev.index)),
                        "ghr_adam_step_sharded_synced")
                except Exception:
                    self.abort()  # the peers are (or will be) waiting for this rank inside their kernels
                    raise
            return
        with _lib.on_device(dev):
            self._barrier_on_stream()  # every rank's bucket is complete
            _lib.check(_lib.lib().ghr_adam_step_sharded(
                self.world, self.rank, self._g_ptrs, self._p_ptrs, self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                self.lo, self.hi, len(self.param_groups), table, float(self.betas[0]), float(self.betas[1]),
                float(self.eps), _lib.raw_stream(dev.index or 0)), "ghr_adam_step_sharded")
            self._barrier_on_stream()  # every replica is written, every bucket has been read
