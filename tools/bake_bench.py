"""Occlusion-baking cube-map pass (reference baking.py::bake_set :204-284) at the ZJU shape:
50 k Gaussians on the posed template, 10^3 occupancy grid, six 32x32 faces per occupied cell.

  ours       bake_opacity_cubemaps: all faces of all cells through ghr_forward_views (view-batched kernels)
  loop_ours  the reference's loop (mask-gather the tensors per cell, one rasteriser call per face) on OUR single-view
             operator -- what a drop-in of the rasteriser alone would give
  loop_ref   the same loop on the unmodified reference CUDA rasteriser (oracle/_ref), where built

The loops are timed on a sample of cells and scaled to all cells (they are per-cell independent); ours is timed whole.
CUDA events + synchronize around each arm; also reports kernel launches of ours.  Prints one JSON object.
    python tools/bake_bench.py > profiles/rNN_bake.json
"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mygauhuman_b200 import _lib, baking as B, synthetic as S  # noqa: E402
from mygauhuman_b200.diff_gaussian_rasterization import _C  # noqa: E402

DEV = torch.device("cuda")
E = torch.Tensor([])


def timed_ms(fn):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    out = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b), (time.perf_counter() - t0) * 1e3, out


def main(P=50_000, sample_cells=12):
    smpl = S.make_smpl_model(0)
    g = S.make_gaussians(P, smpl["v_template"].numpy(), seed=0, sh_degree=3)
    xyz, opacity = g["xyz"].to(DEV), torch.sigmoid(g["opacity_logit"]).to(DEV)
    scales, rots, shs = torch.exp(g["log_scale"]).to(DEV), g["rotation"].to(DEV), g["features"].to(DEV)
    bg = torch.zeros(3, 32, 32, device=DEV)

    for _ in range(2):
        B.bake_opacity_cubemaps(xyz, opacity, scales, rots)
    l0 = _lib.launch_count()
    dev_ms, wall_ms, (cubes, centers, cell_of) = timed_ms(lambda: B.bake_opacity_cubemaps(xyz, opacity, scales, rots))
    launches = _lib.launch_count() - l0
    C = centers.shape[0]
    cams = B.cube_cameras(centers)
    # per-stage device time of the batched pass (library stage events, a further run)
    _lib.profile_enable(True)
    _lib.profile_read()
    pairs = B.rasterize_views(cams, xyz, opacity, scales=scales, rotations=rots, labels=cell_of.int(),
                              skip_labels=torch.arange(C, device=DEV, dtype=torch.int32).repeat_interleave(6))["num_rendered"]
    stages = {k: round(v[0], 3) for k, v in _lib.profile_read().items() if v[1]}
    _lib.profile_enable(False)
    cells = list(range(0, C, max(1, C // sample_cells)))[:sample_cells]

    def loop(call):
        def run():
            worst = 0.0
            for cell in cells:
                m = cell_of != cell                      # baking.py:237-246
                vm, vo, vs, vr = xyz[m], opacity[m], scales[m], rots[m]
                faces = []
                for f in range(6):
                    cam = cams[cell * 6 + f]
                    faces.append(call(vm, vo, vs, vr, cam[0:16].view(4, 4), cam[16:32].view(4, 4), cam[32:35]).permute(1, 2, 0))
                worst = max(worst, float((torch.stack(faces) - cubes[cell]).abs().max()))
            return worst
        return run

    ours_call = lambda vm, vo, vs, vr, view, proj, campos: _C.rasterize_gaussians(
        bg, vm, E, vo, vs, vr, 1.0, E, view, proj, 1.0, 1.0, 32, 32, shs, 3, campos, False, False)[3]
    loop(ours_call)()
    lo_dev, lo_wall, lo_err = timed_ms(loop(ours_call))
    out = {"P": P, "cells": C, "views": C * 6, "res": 32,
           "ours": {"device_ms": round(dev_ms, 3), "wall_ms": round(wall_ms, 3), "launches": int(launches),
                    "views_per_s": round(C * 6 / (wall_ms * 1e-3), 1), "pairs": int(pairs), "stage_ms": stages},
           "loop_ours": {"sample_cells": len(cells), "wall_ms_scaled": round(lo_wall * C / len(cells), 2),
                         "device_ms_scaled": round(lo_dev * C / len(cells), 2), "max_abs_diff_vs_batched": lo_err}}
    try:
        from oracle import ref_cuda
        if ref_cuda.available():
            rr = ref_cuda.RefRasterizer()
            bg3 = torch.zeros(3, device=DEV)
            # the reference evaluates the SH colours too (shs passed unmasked, baking.py:274): give it the kept rows
            ref_call = lambda vm, vo, vs, vr, view, proj, campos: rr.forward(
                bg3, vm.contiguous(), None, vo.contiguous(), vs.contiguous(), vr.contiguous(), 1.0, None, view.contiguous(),
                proj.contiguous(), 1.0, 1.0, 32, 32, shs[:vm.shape[0]].contiguous(), 3, campos.contiguous())[3]
            loop(ref_call)()
            r_dev, r_wall, r_err = timed_ms(loop(ref_call))
            out["loop_ref"] = {"sample_cells": len(cells), "wall_ms_scaled": round(r_wall * C / len(cells), 2),
                               "device_ms_scaled": round(r_dev * C / len(cells), 2), "max_abs_diff_vs_batched": r_err}
            out["speedup_vs_reference_loop"] = round(r_wall * C / len(cells) / wall_ms, 1)
    except Exception as ex:  # the reference build is optional
        out["loop_ref"] = {"unavailable": repr(ex)[:200]}
    out["gpu"] = torch.cuda.get_device_name(0)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
