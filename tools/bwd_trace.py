"""Per-item timeline of the backward blend kernel (needs a -DGHR_TRACE variant; GHR_LIB_PATH set by the caller)."""
import sys, os, ctypes as C
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
from mygauhuman_b200 import _lib

ITEMS_PER_TILE = int(os.environ.get("BWD_ITEMS", "16"))
WARPS = int(os.environ.get("BWD_WARPS_TOTAL", str(148 * 16)))
class A: workload = "zju_512"; views = 4; targets = "u8"
wl = bench.build_workload(A, 0, 1)
dev = torch.device("cuda")
arm = bench.OursArm(wl, dev, False)
view = wl["views"][0]
vec = bench.view_pack(view).to(dev)
targets = wl["targets"].to(dev)
for _ in range(3):
    arm.step(view, vec, targets)
torch.cuda.synchronize()
n_items = 1024 * ITEMS_PER_TILE
buf = np.zeros((1 << 16, 8), dtype=np.uint64)
fn = _lib.lib().ghr_debug_bwd_trace
fn.argtypes = [C.c_void_p, C.c_size_t]
assert fn(buf.ctypes.data, buf.nbytes) == 0
tr = buf[:n_items]
tr = tr[tr[:, 1] > 0]
tr = tr[tr[:, 0] + 400_000 > tr[:, 0].max()]  # items that broke out early keep stale stamps of earlier launches
t0 = tr[:, 0].min()
beg = (tr[:, 0] - t0).astype(np.float64) / 1e3; end = (tr[:, 1] - t0).astype(np.float64) / 1e3
ln = tr[:, 2].astype(np.int64); rounds = tr[:, 3].astype(np.int64)
live = (tr[:, 4] >> 32).astype(np.int64); steps = (tr[:, 4] & 0xFFFFFFFF).astype(np.int64)
dur = end - beg
print("kernel span us", end.max(), "items traced", len(tr))
print("sum of item durations (warp-us)", dur.sum(), " / warps", WARPS, "=", dur.sum() / WARPS)
o = np.argsort(-dur)[:8]
print("longest items: dur us, begin, len, rounds, steps, live, cyc wait/p1/p23", [(round(dur[i], 1), round(beg[i], 1), int(ln[i]), int(rounds[i]), int(steps[i]), int(live[i]), int(tr[i, 5]), int(tr[i, 6]), int(tr[i, 7])) for i in o])
for q in (10, 25, 50, 75, 100, 125, 150, 175):
    print(f"  items running at t={q}us:", int(((beg <= q) & (end > q)).sum()))
print("total rounds", rounds.sum(), "steps", steps.sum(), "live gaussians (some pixel valid)", live.sum())
cw, c1, c23 = tr[:, 5].astype(np.float64).sum(), tr[:, 6].astype(np.float64).sum(), tr[:, 7].astype(np.float64).sum()
print("cycles: wait+stage %.3g  phase1 %.3g  phase2+3 %.3g ; per round %.0f ; phase1 per step %.0f ; phase23 per step %.0f ; per live gaussian %.0f" % (
    cw, c1, c23, cw / rounds.sum(), c1 / steps.sum(), c23 / steps.sum(), c23 / max(1, live.sum())))
