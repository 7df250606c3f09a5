"""ORACLE -- TEST INFRASTRUCTURE ONLY.

Generates tests/golden/raster_*.npz by running the UNMODIFIED reference CUDA
rasteriser (oracle/_ref/libref_gauhuman.so, built by oracle/build_ref.sh from
/root/reference) on a B200 on seeded synthetic inputs.  Run on the GPU box:

    gpurun -- 'python oracle/gen_golden_raster.py gpurun_out/golden'

then copy gpurun_out/golden/*.npz into tests/golden/.  The fixtures pin the CPU
oracle (tests/test_oracle_cpu.py) and the CUDA path (tests/test_golden_gpu.py).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_cuda  # noqa: E402
from tests import common as Cm  # noqa: E402

GOLDEN_CASES = [
    # name, P, W, H, mode, degree, scale, seed
    ("precomp_128x96", 1200, 128, 96, "precomp", 3, 0.01, 101),
    ("sh_scale_rot_100x72", 1000, 100, 72, "sh_scale_rot", 3, 0.012, 102),
    ("sh_cov_deg1_80x64", 600, 80, 64, "sh_cov", 1, 0.03, 103),
    ("big_splats_64x64", 300, 64, 64, "color_scale_rot", 0, 0.25, 104),
]


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    for name, P, W, H, mode, degree, scale, seed in GOLDEN_CASES:
        case = Cm.raster_case(P, W, H, seed=seed, mode=mode, degree=degree, scale=scale)
        tc = Cm.to_torch(case, "cuda")
        ref = ref_cuda.RefRasterizer()
        n, color, depth, alpha, radii = ref.forward(tc["bg"], tc["means3D"], tc["colors"], tc["opacity"], tc["scales"],
                                                    tc["rotations"], tc["scale_modifier"], tc["cov3D"], tc["view"],
                                                    tc["proj"], tc["tanfovx"], tc["tanfovy"], H, W, tc["sh"], degree,
                                                    tc["campos"])
        st = ref.state()
        g = ref.backward(tc["bg"], tc["means3D"], radii, tc["colors"], tc["scales"], tc["rotations"], 1.0, tc["cov3D"],
                         tc["view"], tc["proj"], tc["tanfovx"], tc["tanfovy"], tc["dL_dcolor"], tc["dL_ddepth"],
                         tc["dL_dalpha"], tc["sh"], degree, tc["campos"], alpha)
        out = {"num_rendered": np.int64(n), "radii": radii.cpu().numpy(), "color": color.cpu().numpy(),
               "depth": depth.cpu().numpy(), "alpha": alpha.cpu().numpy()}
        for k in ("tiles_touched", "point_offsets", "point_list_keys", "point_list", "ranges", "n_contrib", "final_T",
                  "depths", "means2D", "conic_opacity", "rgb", "cov3D", "clamped"):
            out[k] = st[k]
        for k, v in g.items():
            out["g_" + k] = v.cpu().numpy()
        for k, v in case.items():
            out["in_" + k] = np.zeros(0, np.float32) if v is None else np.asarray(v)
        path = os.path.join(outdir, f"raster_{name}.npz")
        np.savez_compressed(path, **out)
        print(name, "N =", n, "visible =", int((out["radii"] > 0).sum()), os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
