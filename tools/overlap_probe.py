"""Probe: how much do K graphed views overlap when replayed on K streams?  (tuning aid, not shipped API)
K independent GraphedRenderStep instances (own deformer workspaces, own gradient buckets), view i on stream i % K."""
import json, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench


class A:  # bench args
    workload, views, batch_views, mode, sync_free = "zju_512", 64, 8, "graph", False


dev = torch.device("cuda")
wl = bench.build_workload(A, 0, 1)
views = wl["views"]
vec_dev = [bench.view_pack(v).to(dev) for v in views]
targets = wl["targets"].to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
out = {}
for K in (1, 2, 3, 4):
    arms = [bench.OursArm(wl, dev, False, graph=True) for _ in range(K)]
    for a in arms:
        a.capture(vec_dev, targets)
        a.use_graph = True
    streams = [torch.cuda.Stream() for _ in range(K)]
    cur = torch.cuda.current_stream()

    def group(i0):
        ev0 = torch.cuda.Event(); ev0.record()
        for k in range(K):
            streams[k].wait_event(ev0)
            with torch.cuda.stream(streams[k]):
                arms[k].graphed(vec_dev[(i0 + k) % len(views)], arms[k].graphed.loss_inputs[0])
            e = torch.cuda.Event(); e.record(streams[k]); cur.wait_event(e)

    for a in arms:
        a.graphed.zero_grads()
    for i in range(0, 24, K):
        group(i)
    torch.cuda.synchronize()
    ts = []
    n_groups = 96 // K
    for gi in range(n_groups):
        flush.fill_(gi & 0xFF)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); group(gi * K); e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    per_view = float(np.sum(ts)) / (n_groups * K)
    out[f"streams_{K}"] = {"ms_per_view": round(per_view, 4), "views_per_s": round(1e3 / per_view, 1)}
    del arms
print(json.dumps(out))
