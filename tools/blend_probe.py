"""Stage times (us, L2 flushed before every call) of the rasteriser on bench.py's workload, for tuning variants:
    GHR_LIB_PATH=$PWD/variants/libghr_TAG.so python tools/blend_probe.py [workload] [views]"""
import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
from mygauhuman_b200 import _lib
from mygauhuman_b200.diff_gaussian_rasterization import _C

class A: workload = sys.argv[1] if len(sys.argv) > 1 else "zju_512"; views = int(sys.argv[2]) if len(sys.argv) > 2 else 4; targets = "u8"
wl = bench.build_workload(A, 0, 1)
dev = torch.device("cuda")
arm = bench.OursArm(wl, dev, False)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
tg = bench.targets_f32(wl["targets"].to(dev))
gc, gd, ga = tg[0:3].contiguous(), tg[3:4].contiguous(), tg[4:5].contiguous()
e = torch.Tensor([])
acc = {}
for vi, view in enumerate(wl["views"]):
    vec = bench.view_pack(view).to(dev)
    camd, pose = bench.unpack(vec)
    L = arm.leaves
    cam = view["cam"]
    with torch.no_grad():
        d = arm.deformer.deform_gaussians(L["xyz"], pose, arm.big, arm.t_vertices, scaling=torch.exp(L["log_scale"]),
                                          rotation=L["rotation"], features=L["features"], sh_degree=3, campos=camd["campos"])
        opa = torch.sigmoid(L["opacity_logit"])
        fa = (arm.bg, d["means3D"], d["colors"], opa, e, e, 1.0, d["cov3D"], camd["view"], camd["proj"], cam["tanfovx"],
              cam["tanfovy"], wl["H"], wl["W"], e, 3, camd["campos"], False, False)
        def run():
            fw = _C.rasterize_gaussians(*fa)
            n, color, depth, alpha, radii, geom, binning, img = fw
            _C.rasterize_gaussians_backward(arm.bg, d["means3D"], radii, d["colors"], e, e, 1.0, d["cov3D"], camd["view"],
                                            camd["proj"], cam["tanfovx"], cam["tanfovy"], gc, gd, ga, e, 3, camd["campos"],
                                            geom, n, binning, img, alpha, False)
            return n
        for _ in range(3):
            n = run()
        torch.cuda.synchronize()
        _lib.profile_enable(True); _lib.profile_read()
        for i in range(10):
            flush.fill_(i)
            run()
        torch.cuda.synchronize()
        prof = _lib.profile_read(); _lib.profile_enable(False)
    for k, (ms, c) in prof.items():
        if c:
            acc.setdefault(k, []).append(1e3 * ms / c)
print(os.path.basename(os.environ.get("GHR_LIB_PATH", "libghr.so")), A.workload, "N(last)", n,
      {k: round(float(np.mean(v)), 1) for k, v in acc.items()}, "total", round(sum(float(np.mean(v)) for v in acc.values()), 1))
