import sys, torch, numpy as np, argparse
sys.path.insert(0, '/root/repo')
import bench as B
args = argparse.Namespace(workload="zju_512", views=8, targets="u8")
wl = B.build_workload(args, 0, 1)
dev = torch.device("cuda", 0)
ours = B.OursArm(wl, dev, False, graph=False)
ref = B.ReferenceArm(wl, dev)
print(ref.lbs_source)
view = wl["views"][0]
vec = B.view_pack(view).to(dev)
tg = wl["targets"].to(dev)
camd, pose = B.unpack(vec)
L = ours.leaves
d = ours.deformer.deform_gaussians(L["xyz"], pose, ours.big, ours.t_vertices, normals=L["normals"], scaling=torch.exp(L["log_scale"]), rotation=L["rotation"], features=L["features"], sh_degree=3, campos=camd["campos"])
world, cov, col, opac = ref.deform(pose, camd["campos"])
vid_o = ours.deformer.nearest_vertex(ours.t_vertices, L["xyz"]).long()
_, vid_r = ref.model.knn(ref.t_vertices[None], ref.leaves["xyz"][None])
print("vid mismatch", int((vid_o != vid_r.reshape(-1)).sum()))
def rel(a, b): return float((a-b).abs().max() / b.abs().max())
print("means", rel(d["means3D"], world), "cov", rel(d["cov3D"], cov), "col", rel(d["colors"], col))
bad = (d["means3D"]-world).abs().max(1).values > 1e-4
print("means off >1e-4:", int(bad.sum()))
lo, go = ours.step(view, vec, tg)
lr, gr = ref.step(view, vec, tg)
print("loss", float(lo.detach()), float(lr.detach()))
for n, a, b in zip(("xyz","features","opacity","log_scale","rotation"), go, gr):
    e = (a-b).abs(); s = b.abs().max()
    print(n, "max err/scale", float(e.max()/s), "frac>1e-4", float((e > 1e-4*b.abs()+1e-5*s).float().mean()))
