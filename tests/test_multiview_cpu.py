"""World-size-2 gloo test of the N>1 path's host logic (view sharding + gradient all-reduce): two CPU ranks each run the
oracle's LBS + rasteriser step on their shard of the views; the all-reduced gradients must equal the single-process sum
over all views.  (The CUDA kernels themselves are covered by the -m gpu tests; the collective is the only multi-rank piece.)"""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _view_grads(view_idx, P=300, W=48, H=48):
    """Gradients of a linear loss of one view w.r.t. (xyz, features) through oracle LBS + C raster oracle."""
    sys.path.insert(0, ROOT)
    import oracle as O
    from oracle import lbs_oracle as LO
    from mygauhuman_b200 import synthetic as S
    smpl = S.make_smpl_model(0)
    big = S.big_pose_params()
    tv = S.big_pose_vertices(smpl, big)
    g = S.make_gaussians(P, tv.numpy(), seed=0, sh_degree=3, scale=0.03)
    cam = S.make_camera(W, H, angle=0.7 * view_idx, radius=3.0, seed=view_idx)
    pose = S.make_pose_params(view_idx)
    pose["Th"] = pose["Th"] * 0.1
    xyz = g["xyz"].clone().requires_grad_(True)
    feat = g["features"].clone().requires_grad_(True)
    _, world, _, T, _, _, _ = LO.coarse_deform_c2source(smpl, xyz, pose, big, tv, normals=g["normals"])
    cov = LO.covariance(torch.exp(g["log_scale"]), 1.0, g["rotation"], T)
    col = LO.view_colors(3, feat, world, cam["camera_center"])
    a = lambda t: np.ascontiguousarray(t.detach().numpy(), dtype=np.float32)
    opa = torch.sigmoid(g["opacity_logit"])
    vm, pm, cp = a(cam["world_view_transform"]), a(cam["full_proj_transform"]), a(cam["camera_center"])
    st = O.raster_forward(np.zeros(3, np.float32), a(world), a(col), a(opa), None, None, 1.0, a(cov), vm, pm,
                          cam["tanfovx"], cam["tanfovy"], H, W, None, 3, cp)
    rng = np.random.default_rng(view_idx)
    gc = rng.normal(size=(3, H, W)).astype(np.float32)
    z = np.zeros((1, H, W), np.float32)
    gr = O.raster_backward(st, np.zeros(3, np.float32), a(world), a(col), None, None, 1.0, a(cov), vm, pm, cam["tanfovx"],
                           cam["tanfovy"], gc, z, z, None, 3, cp)
    # (the covariance depends on the pose and skinning weights only, not on xyz / SH)
    torch.autograd.backward([world, col], [torch.from_numpy(gr["dL_dmeans3D"]), torch.from_numpy(gr["dL_dcolors"])])
    return [xyz.grad.clone(), feat.grad.clone()]


def _worker(rank, world, port, n_views, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    from mygauhuman_b200.multiview import GradientAllReducer, shard_views
    mine = shard_views(n_views, rank, world)
    acc = None
    for v in mine:
        g = _view_grads(v)
        acc = g if acc is None else [a + b for a, b in zip(acc, g)]
    out = GradientAllReducer()(acc)
    if rank == 0:
        q.put([t.numpy() for t in out])
    dist.barrier()
    dist.destroy_process_group()


def test_view_sharding_allreduce_equals_batch_accumulation():
    from mygauhuman_b200.multiview import shard_views
    n_views, world = 4, 2
    shards = [shard_views(n_views, r, world) for r in range(world)]
    assert sorted(sum(shards, [])) == list(range(n_views))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_views, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    want = None
    for v in range(n_views):
        g = _view_grads(v)
        want = g if want is None else [a + b for a, b in zip(want, g)]
    for a, b in zip(got, want):
        assert np.abs(b.numpy()).max() > 0
        np.testing.assert_allclose(a, b.numpy(), rtol=1e-5, atol=1e-7)


def test_flat_layout_and_shard_bounds_partition_the_buffer():
    """Host logic of the sharded optimiser step (mygauhuman_b200.optim): the flat layout pads every leaf to 16 bytes and
    the ranks' slices tile the buffer without gaps, on 4-float boundaries."""
    import torch
    from mygauhuman_b200.optim import flat_layout, shard_bounds
    leaves = {"xyz": torch.zeros(5001, 3), "rotation": torch.zeros(5001, 4), "opacity": torch.zeros(5001, 1)}
    total, table = flat_layout(leaves)
    assert total % 4 == 0 and table["xyz"] == (0, 15003) and table["rotation"][0] == 15004
    assert table["opacity"][0] == 15004 + 20004 and total == 15004 + 20004 + 5004
    for world in (1, 2, 3, 4, 8):
        edges = [shard_bounds(total, world, r) for r in range(world)]
        assert edges[0][0] == 0 and edges[-1][1] == total
        assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
        assert all(lo % 4 == 0 and hi % 4 == 0 and hi >= lo for lo, hi in edges)
