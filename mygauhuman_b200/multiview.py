"""View-sharded data parallelism for multi-view batches (not in the reference, which is
single-GPU: utils/general_utils.py:140 hard-codes cuda:0).

Every rank holds the replicated Gaussian parameters and renders its own cameras; the only exchange on
the path is ONE sum all-reduce of the parameter gradients per step (NCCL over NVLink on GPUs, gloo in
the CPU tests).  Semantics = gradient accumulation over the whole batch of views.
"""
import torch
import torch.distributed as dist


def shard_views(n_views, rank, world):
    """Indices of the views rank `rank` renders: rank, rank+world, ... (round-robin keeps heavy/light views mixed)."""
    return list(range(rank, n_views, world))


class GradientAllReducer:
    """Sums a fixed list of gradient tensors over all ranks through one flat fp32 bucket
    (59 floats per Gaussian at SH degree 3: xyz 3 + SH 48 + opacity 1 + scale 3 + rotation 4)."""

    def __init__(self, group=None):
        self.group = group
        self.flat = None

    def __call__(self, grads):
        if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(self.group) == 1:
            return grads
        n = sum(g.numel() for g in grads)
        if self.flat is None or self.flat.numel() != n or self.flat.device != grads[0].device:
            self.flat = torch.empty(n, dtype=grads[0].dtype, device=grads[0].device)
        torch.cat([g.reshape(-1) for g in grads], out=self.flat)
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
        out, o = [], 0
        for g in grads:
            v = self.flat[o:o + g.numel()].view_as(g)
            g.copy_(v)
            out.append(g)
            o += g.numel()
        return out
