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
