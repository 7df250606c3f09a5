"""Quick kernel-level timing: ours vs the reference CUDA build, rasteriser fwd+bwd (U1)."""
import sys, os, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from tests import common as Cm
from oracle import ref_cuda
from mygauhuman_b200.diff_gaussian_rasterization import _C

def timeit(fn, iters=50, warm=10):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

def main():
    for (P, W, H, mode) in [(50000, 512, 512, "precomp"), (50000, 512, 512, "sh_scale_rot"), (100000, 1024, 1024, "precomp")]:
        case = Cm.raster_case(P, W, H, seed=0, mode=mode)
        tc = Cm.to_torch(case, "cuda")
        fw = Cm.ours_forward(tc)
        t_f = timeit(lambda: Cm.ours_forward(tc))
        t_b = timeit(lambda: Cm.ours_backward(tc, fw))
        _C.set_sync_free_capacity(int(fw[0] * 1.5))
        t_fs = timeit(lambda: Cm.ours_forward(tc))
        _C.set_sync_free_capacity(None)
        ref = ref_cuda.RefRasterizer()
        def rf():
            return ref.forward(tc["bg"], tc["means3D"], tc["colors"], tc["opacity"], tc["scales"], tc["rotations"], 1.0, tc["cov3D"], tc["view"], tc["proj"], tc["tanfovx"], tc["tanfovy"], H, W, tc["sh"], case["degree"], tc["campos"])
        n, color, depth, alpha, radii = rf()
        def rb():
            return ref.backward(tc["bg"], tc["means3D"], radii, tc["colors"], tc["scales"], tc["rotations"], 1.0, tc["cov3D"], tc["view"], tc["proj"], tc["tanfovx"], tc["tanfovy"], tc["dL_dcolor"], tc["dL_ddepth"], tc["dL_dalpha"], tc["sh"], case["degree"], tc["campos"], alpha)
        t_rf = timeit(rf); t_rb = timeit(rb)
        print(f"P={P} {W}x{H} {mode}: N={fw[0]} ours fwd {t_f:.3f} ms (sync-free {t_fs:.3f}) bwd {t_b:.3f} ms | ref fwd {t_rf:.3f} bwd {t_rb:.3f} ms | speedup fwd+bwd {(t_rf+t_rb)/(t_f+t_b):.2f}x", flush=True)

if __name__ == "__main__":
    main()
