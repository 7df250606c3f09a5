"""preprocess_kernel at 3 M Gaussians / 1080p with and without the SH -> RGB path (colours precomputed instead): what the
192-byte-stride SH row loads cost.  python tools/pre_probe.py"""
import math, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from mygauhuman_b200 import _lib, synthetic as S
from mygauhuman_b200.diff_gaussian_rasterization import _C
dev = torch.device("cuda", 0)
P, W, H = 3_000_000, 1920, 1080
g = torch.Generator(device=dev).manual_seed(7)
xyz = torch.rand(P, 3, device=dev, generator=g) * 10.0 - 5.0; xyz[:, 2] += 6.0
scal = torch.exp(math.log(0.02) + 0.3 * torch.randn(P, 3, device=dev, generator=g))
rot = torch.nn.functional.normalize(torch.randn(P, 4, device=dev, generator=g))
opac = torch.sigmoid(1.5 * torch.randn(P, 1, device=dev, generator=g))
sh = torch.cat([torch.rand(P, 1, 3, device=dev, generator=g) * 2 - 1, 0.05 * torch.randn(P, 15, 3, device=dev, generator=g)], 1).contiguous()
col = torch.rand(P, 3, device=dev, generator=g)
cam = S.make_camera(W, H, angle=math.pi, radius=6.0, target=(0.0, 0.0, 6.0), elev=0.0, seed=0, focal=1100.0)
e = torch.Tensor([])
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def run(colors, shs):
    args = (torch.zeros(3, device=dev), xyz, colors, opac, scal, rot, 1.0, e, cam["world_view_transform"].to(dev),
            cam["full_proj_transform"].to(dev), cam["tanfovx"], cam["tanfovy"], H, W, shs, 3, cam["camera_center"].to(dev), False, False)
    _C.rasterize_gaussians(*args)
    _lib.profile_enable(True); _lib.profile_read()
    for i in range(5):
        flush.fill_(i); _C.rasterize_gaussians(*args)
    torch.cuda.synchronize()
    prof = _lib.profile_read(); _lib.profile_enable(False)
    return {k: round(1e3 * t / c, 1) for k, (t, c) in prof.items() if c}
print("SH path      ", run(e, sh))
print("precomputed  ", run(col, e))
