"""Probe: nearest-vertex kernel timing + grid statistics on the bench workload."""
import os, sys, struct
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from mygauhuman_b200 import synthetic as S
from mygauhuman_b200.lbs import SMPLDeformer

smpl = S.make_smpl_model(0); big = S.big_pose_params(); tv = S.big_pose_vertices(smpl, big)
g = S.make_gaussians(50000, tv.numpy(), seed=0)
d = SMPLDeformer(smpl, "cuda")
tvc, q = tv.cuda(), g["xyz"].cuda()
vid = d.nearest_vertex(tvc, q)
torch.cuda.synchronize()
ws = d._grid.cpu().numpy()
minx, miny, minz, inv_cell, cell, nx, ny, nz = struct.unpack("fffffiii", ws[64:96].tobytes())
print("grid", nx, ny, nz, "cell", cell, "cells", nx * ny * nz)
# candidate statistics (host)
c = np.floor((tv.numpy() - [minx, miny, minz]) * inv_cell).astype(int).clip(0, [nx - 1, ny - 1, nz - 1])
cid = (c[:, 2] * ny + c[:, 1]) * nx + c[:, 0]
cnt = np.bincount(cid, minlength=nx * ny * nz).reshape(nz, ny, nx)
print("occupied cells", (cnt > 0).sum(), "max per cell", cnt.max())
pad = np.pad(cnt, 1)
blk = sum(pad[1 + dz:1 + dz + nz, 1 + dy:1 + dy + ny, 1 + dx:1 + dx + nx] for dz in (-1, 0, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1))
qc = np.floor((g["xyz"].numpy() - [minx, miny, minz]) * inv_cell).astype(int).clip(0, [nx - 1, ny - 1, nz - 1])
cand = blk[qc[:, 2], qc[:, 1], qc[:, 0]]
print("candidates per query: mean", cand.mean(), "max", cand.max())
for name in ("warm",):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    for _ in range(5): d.nearest_vertex(tvc, q)
    torch.cuda.synchronize(); e0.record()
    for _ in range(50): d.nearest_vertex(tvc, q)
    e1.record(); torch.cuda.synchronize()
    print("nearest_vertex avg ms", e0.elapsed_time(e1) / 50)
ref = torch.cdist(q.double(), tvc.double()).argmin(1)
print("agree", float((ref == vid.long()).float().mean()))
