"""Optimiser step of view-sharded data parallelism, per rank, under torchrun (one process per GPU):

  fused     ShardedFusedAdam.step(): barrier + ONE kernel (reduce-scatter of the ranks' gradient buckets over NVLink peer
            loads + Adam on the owned slice + all-gather of the new parameters by peer stores) + barrier
  unfused   NCCL all-reduce of the bucket + FusedAdam.step() over the full replica (what the bench's step would do)
  torch     NCCL all-reduce + torch.optim.Adam (foreach): the reference's optimiser with a collective in front

CUDA events on the launching stream, max over ranks, median of the iterations; P Gaussians x 59 floats (+ padding).
    python -m torch.distributed.run --nproc-per-node N tools/sharded_adam_bench.py > profiles/rNN_sharded_adam_nN.json
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mygauhuman_b200.optim import FusedAdam, ShardedFusedAdam, flat_layout, flatten_leaves  # noqa: E402

SHAPES = {"xyz": (3,), "log_scale": (3,), "rotation": (4,), "features": (16, 3), "opacity_logit": (1,)}
LRS = {"xyz": 1.6e-4, "log_scale": 5e-3, "rotation": 1e-3, "features": 2.5e-3, "opacity_logit": 5e-2}


def timed(fn, dev, iters=30, warm=5):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(float(t.item()))
    return float(np.median(ts))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    out = []
    for P in (50_000, 300_000):
        g0 = torch.Generator().manual_seed(0)
        leaves = {k: torch.randn(P, *s, generator=g0).to(dev) for k, s in SHAPES.items()}
        flat, views = flatten_leaves(leaves)
        total, table = flat_layout(leaves)
        bucket = torch.randn(total, generator=torch.Generator().manual_seed(1 + rank)).to(dev) * 1e-3
        groups = [dict(name=k, offset=table[k][0], count=table[k][1], lr=LRS[k]) for k in SHAPES]
        fused = ShardedFusedAdam(flat, bucket, groups, eps=1e-15, sync="flags")
        fused_nccl = ShardedFusedAdam(flat, bucket, groups, eps=1e-15, sync="nccl")
        ref = {k: torch.nn.Parameter(v.clone()) for k, v in leaves.items()}
        for k in ref:
            o, n = table[k]
            ref[k].grad = bucket[o:o + n].view(ref[k].shape)       # gradients alias the bucket, as in the render step
        ours = FusedAdam([{"params": [ref[k]], "lr": LRS[k]} for k in SHAPES], lr=0.0, eps=1e-15)
        tref = {k: torch.nn.Parameter(v.clone()) for k, v in leaves.items()}
        for k in tref:
            o, n = table[k]
            tref[k].grad = bucket[o:o + n].view(tref[k].shape)
        tadam = torch.optim.Adam([{"params": [tref[k]], "lr": LRS[k]} for k in SHAPES], lr=0.0, eps=1e-15)

        def unfused():
            dist.all_reduce(bucket)
            ours.step()

        def torch_path():
            dist.all_reduce(bucket)
            tadam.step()

        res = {"P": P, "floats": total, "world": world,
               "fused_ms": round(timed(fused.step, dev), 4),
               "fused_nccl_barriers_ms": round(timed(fused_nccl.step, dev), 4),
               "unfused_allreduce_plus_fused_adam_ms": round(timed(unfused, dev), 4),
               "allreduce_only_ms": round(timed(lambda: dist.all_reduce(bucket), dev), 4),
               "allreduce_plus_torch_adam_ms": round(timed(torch_path, dev), 4),
               "nvlink_bytes_per_rank": int(2 * 4 * total * (world - 1) / world)}
        out.append(res)
        del fused, fused_nccl
    if rank == 0:
        sys.stderr.flush()
        print(json.dumps({"sharded_adam": out, "gpu": torch.cuda.get_device_name(0)}), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
