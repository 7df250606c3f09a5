"""Per-item timeline of the forward blend kernel (needs variants/libghr_trace.so; GHR_LIB_PATH set by the caller)."""
import sys, os, ctypes as C
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
from mygauhuman_b200 import _lib
from mygauhuman_b200.diff_gaussian_rasterization import _C

class A: workload = "zju_512"; views = 4; targets = "u8"
wl = bench.build_workload(A, 0, 1)
dev = torch.device("cuda")
arm = bench.OursArm(wl, dev, False)
view = wl["views"][0]
vec = bench.view_pack(view).to(dev)
camd, pose = bench.unpack(vec)
L = arm.leaves
with torch.no_grad():
    d = arm.deformer.deform_gaussians(L["xyz"], pose, arm.big, arm.t_vertices, scaling=torch.exp(L["log_scale"]),
                                      rotation=L["rotation"], features=L["features"], sh_degree=3, campos=camd["campos"])
    e = torch.Tensor([]); cam = view["cam"]
    for _ in range(3):
        out = _C.rasterize_gaussians(arm.bg, d["means3D"], d["colors"], torch.sigmoid(L["opacity_logit"]), e, e, 1.0, d["cov3D"],
                                     camd["view"], camd["proj"], cam["tanfovx"], cam["tanfovy"], wl["H"], wl["W"], e, 3,
                                     camd["campos"], False, False)
    torch.cuda.synchronize()
n_items = 1024 * 16
buf = np.zeros((1 << 16, 8), dtype=np.uint64)
fn = _lib.lib().ghr_debug_fwd_trace
fn.argtypes = [C.c_void_p, C.c_size_t]
assert fn(buf.ctypes.data, buf.nbytes) == 0
tr = buf[:n_items]
t0 = tr[:, 0].min()
beg = (tr[:, 0] - t0).astype(np.float64) / 1e3; end = (tr[:, 1] - t0).astype(np.float64) / 1e3
ln = (tr[:, 2] & 0xFFFFFFFF).astype(np.int64); sm = (tr[:, 2] >> 32).astype(np.int64); rounds = tr[:, 3].astype(np.int64)
dur = end - beg
print("kernel span us", end.max(), "items", n_items, "non-empty", int((ln > 0).sum()))
print("sum of item durations (warp-us)", dur.sum(), " / warps", 148 * 12, "=", dur.sum() / (148 * 12))
o = np.argsort(-dur)[:10]
print("longest items: dur us, begin, len, rounds", [(round(dur[i], 1), round(beg[i], 1), int(ln[i]), int(rounds[i])) for i in o])
print("empty items: mean dur us", dur[ln == 0].mean(), "count", int((ln == 0).sum()), "last empty end", end[ln == 0].max())
print("non-empty: begin max", beg[ln > 0].max(), "end max", end[ln > 0].max())
print("us per round (non-empty, rounds>5):", np.median(dur[rounds > 5] / rounds[rounds > 5]))
for q in (10, 20, 40, 60, 80):
    print(f"  items running at t={q}us:", int(((beg <= q) & (end > q)).sum()), "of which non-empty", int(((beg <= q) & (end > q) & (ln > 0)).sum()))
print("total rounds", rounds.sum())
surv = (tr[:, 4] >> 32).astype(np.int64); steps = (tr[:, 4] & 0xFFFFFFFF).astype(np.int64)
print("survivors total", surv.sum(), "steps", steps.sum(), "survivors/entry", surv.sum() / max(1, (rounds * 64).sum()))
cw, cc, ce = tr[:, 5].astype(np.float64).sum(), tr[:, 6].astype(np.float64).sum(), tr[:, 7].astype(np.float64).sum()
print("cycles: wait+stage %.3g  cull %.3g  eval %.3g   per round: %.0f %.0f %.0f ; eval cycles per step %.0f" % (cw, cc, ce, cw / rounds.sum(), cc / rounds.sum(), ce / rounds.sum(), ce / steps.sum()))
