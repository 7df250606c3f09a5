"""Worker of tests/test_sharded_adam_gpu.py (one process per GPU under torchrun): ShardedFusedAdam -- reduce-scatter +
Adam + all-gather over peer memory -- against NCCL all-reduce + FusedAdam, three steps, lr change in between."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from mygauhuman_b200.optim import FusedAdam, ShardedFusedAdam, flat_layout, flatten_leaves, shard_bounds  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    for sync in ("flags", "nccl"):
        run(rank, world, dev, sync)
    if rank == 0:
        print("SHARDED_ADAM_OK", world, flush=True)
    dist.barrier()
    dist.destroy_process_group()


def run(rank, world, dev, sync):
    P = 5001  # odd sizes: padded groups, ragged last shard
    shapes = {"xyz": (P, 3), "log_scale": (P, 3), "rotation": (P, 4), "features": (P, 16, 3), "opacity_logit": (P, 1)}
    lrs = {"xyz": 1.6e-4, "log_scale": 5e-3, "rotation": 1e-3, "features": 2.5e-3, "opacity_logit": 5e-2}
    g0 = torch.Generator().manual_seed(0)                      # same parameters on every rank
    leaves = {k: torch.randn(*s, generator=g0).to(dev) for k, s in shapes.items()}
    flat, views = flatten_leaves(leaves)
    total, table = flat_layout(leaves)
    assert flat.numel() == total and total % 4 == 0
    bucket = torch.zeros_like(flat)
    groups = [dict(name=k, offset=table[k][0], count=table[k][1], lr=lrs[k]) for k in shapes]
    opt = ShardedFusedAdam(flat, bucket, groups, eps=1e-15, sync=sync)
    lo, hi = shard_bounds(total, world, rank)
    assert (opt.lo, opt.hi) == (lo, hi) and opt.exp_avg.numel() >= hi - lo
    # the unfused path on copies: all-reduce the bucket, FusedAdam over the separate tensors
    ref = {k: torch.nn.Parameter(v.clone()) for k, v in leaves.items()}
    ref_opt = FusedAdam([{"params": [ref[k]], "lr": lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
    gr = torch.Generator().manual_seed(100 + rank)             # different gradients per rank (its views)
    for it in range(3):
        bucket.zero_()
        for k in shapes:
            o, n = table[k]
            g = (torch.randn(n, generator=gr) * 10.0 ** float(torch.randint(-5, 1, (1,), generator=gr))).to(dev)
            bucket[o:o + n] = g
        summed = bucket.clone()
        dist.all_reduce(summed)
        for k in shapes:
            o, n = table[k]
            ref[k].grad = summed[o:o + n].view(shapes[k]).clone()
        if it == 2:
            opt.param_groups[0]["lr"] = ref_opt.param_groups[0]["lr"] = 1.2e-4
        ref_opt.step()
        opt.step()
        torch.cuda.synchronize()
        worst = 0.0
        for k in shapes:
            a, b = views[k].detach(), ref[k].detach()
            if world == 2:   # a + b is the same in either order: bit-exact against the NCCL sum
                assert torch.equal(a, b), (k, it, float((a - b).abs().max()))
            worst = max(worst, float((a - b).abs().max() / b.abs().max()))
        assert worst <= 1e-6, worst
        # replicas are bit-identical across ranks
        mine = flat.clone()
        other = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(other, mine)
        assert all(torch.equal(o, other[0]) for o in other)


if __name__ == "__main__":
    main()
