"""Tile-list statistics of bench.py's workload (view 0): list lengths, per-4x4-block cull survivors, contributors."""
import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
from mygauhuman_b200.diff_gaussian_rasterization import _C

class A: workload = sys.argv[1] if len(sys.argv) > 1 else "zju_512"; views = 4; targets = "u8"
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
    e = torch.Tensor([])
    cam = view["cam"]
    out = _C.rasterize_gaussians(arm.bg, d["means3D"], d["colors"], torch.sigmoid(L["opacity_logit"]), e, e, 1.0, d["cov3D"],
                                 camd["view"], camd["proj"], cam["tanfovx"], cam["tanfovy"], wl["H"], wl["W"], e, 3,
                                 camd["campos"], False, False)
n, color, depth, alpha, radii, geom, binning, img = out
sv = _C.state_views(wl["P"], n, wl["W"], wl["H"], geom, binning, img)
ranges = sv["ranges"].cpu().numpy().astype(np.int64)
lens = ranges[:, 1] - ranges[:, 0]
print("N", n, "tiles", len(lens), "non-empty", int((lens > 0).sum()), "max", lens.max(), "mean(non-empty)", lens[lens > 0].mean())
print("top lens", np.sort(lens)[::-1][:12])
print("percentiles of non-empty", np.percentile(lens[lens > 0], [50, 75, 90, 99]))
ncon = sv["n_contrib"].cpu().numpy()
print("n_contrib max", ncon.max(), "mean over covered px", ncon[ncon > 0].mean())
# heaviest tile: survivors per 4x4 block and per 8x4 block, contributors per pixel
t = int(np.argmax(lens))
gx = (wl["W"] + 15) // 16
tx, ty = t % gx, t // gx
ids = sv["point_list"][ranges[t, 0]:ranges[t, 1]].long()
m2 = sv["means2D"][ids].cpu().numpy(); co = sv["conic_opacity"][ids].cpu().numpy()
# alpha >= 1/255 ellipse half extents: power = ln(1/(255*op)); x extent = sqrt(2*|p|*c/(a c - b^2))
a, b, c, op = co[:, 0], co[:, 1], co[:, 2], co[:, 3]
pw = np.log(np.maximum(255.0 * op, 1.0))
det = a * c - b * b
ex = np.sqrt(2 * pw * c / det); ey = np.sqrt(2 * pw * a / det)
print("heaviest tile", t, "len", len(ids), "median extents px", np.median(ex), np.median(ey), "opacity median", np.median(op))
def surv(bx0, by0, w, h):
    ddx = np.maximum(np.maximum(bx0 - m2[:, 0], m2[:, 0] - (bx0 + w - 1)), 0)
    ddy = np.maximum(np.maximum(by0 - m2[:, 1], m2[:, 1] - (by0 + h - 1)), 0)
    return int(((ddx <= ex) & (ddy <= ey)).sum())
s44 = [surv(tx * 16 + 4 * i, ty * 16 + 4 * j, 4, 4) for j in range(4) for i in range(4)]
s84 = [surv(tx * 16 + 8 * i, ty * 16 + 4 * j, 8, 4) for j in range(4) for i in range(2)]
print("survivors per 4x4 block", s44)
print("survivors per 8x4 block", s84)
blk = ncon[ty * 16:ty * 16 + 16, tx * 16:tx * 16 + 16]
print("n_contrib (last contributor pos) in heaviest tile: min/mean/max", blk.min(), blk.mean(), blk.max())
fT = sv["final_T"].cpu().numpy()[ty * 16:ty * 16 + 16, tx * 16:tx * 16 + 16]
print("final_T in heaviest tile min/mean/max", fT.min(), fT.mean(), fT.max())
