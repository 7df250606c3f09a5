"""Probe: do K independent blend passes overlap when issued on K streams?  (tuning aid, not shipped API)"""
import os, sys, json
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mygauhuman_b200 import synthetic as S
from tests import common as Cm

DEV = torch.device("cuda")
P, W, H, K = 50_000, 512, 512, 7
smpl = S.make_smpl_model(0)
tv = S.big_pose_vertices(smpl, S.big_pose_params())
tc = Cm.to_torch(Cm.raster_case(P, W, H, seed=0, mode="precomp", cloud=tv.numpy()), "cuda")
fws = [Cm.ours_forward(tc) for _ in range(K)]
streams = [torch.cuda.Stream() for _ in range(K)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)


def run(multi, what):
    ts = []
    for it in range(25):
        flush.fill_(it & 0xFF)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        cur = torch.cuda.current_stream()
        if multi:
            ev0 = torch.cuda.Event(); ev0.record()
            for k in range(K):
                streams[k].wait_event(ev0)
                with torch.cuda.stream(streams[k]):
                    what(k)
                e = torch.cuda.Event(); e.record(streams[k]); cur.wait_event(e)
        else:
            for k in range(K):
                what(k)
        b.record()
        torch.cuda.synchronize()
        if it >= 5:
            ts.append(a.elapsed_time(b))
    return round(float(np.median(ts)), 4)


out = {}
out["bwd_sequential_ms"] = run(False, lambda k: Cm.ours_backward(tc, fws[k]))
out["bwd_7_streams_ms"] = run(True, lambda k: Cm.ours_backward(tc, fws[k]))
out["fwd_sequential_ms"] = run(False, lambda k: Cm.ours_forward(tc))
out["fwd_7_streams_ms"] = run(True, lambda k: Cm.ours_forward(tc))
print(json.dumps(out, indent=1))
