"""Secondary measurements BASELINE.json's metric names besides the headline (bench.py):

  * sort Mkeys/s sweep: libghr's onesweep (ghr_sort_pairs_u64) vs the reference's
    cub::DeviceRadixSort::SortPairs call (oracle/_ref shim), same keys, N = 2^16 .. 2^26
  * forward-only render FPS at the ZJU 512x512 shape (LBS + cov + SH + rasteriser forward), ours vs reference
  * stress render: 3M Gaussians at 1920x1080, forward only, ours vs the reference CUDA rasteriser
  * distCUDA2 (simple-knn) at 6 890 / 50 000 / 300 000 points, ours vs reference
  * seven-attribute render (reference render(): image + normal + world normal + albedo + occlusion + roughness +
    axis), rasteriser forward + backward at the ZJU shape: forward_multi vs one call per attribute vs the reference CUDA

Every timing: CUDA events on the launching stream around each call, L2 flushed (256 MiB fill) between calls,
median over the iterations.  Prints one JSON object.
    python tools/aux_bench.py [--quick] > profiles/rNN_aux.json
"""
import argparse
import json
import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mygauhuman_b200 import _lib, synthetic as S  # noqa: E402
from mygauhuman_b200 import diff_gaussian_rasterization as dgr  # noqa: E402
from mygauhuman_b200.diff_gaussian_rasterization import _C  # noqa: E402
from mygauhuman_b200.simple_knn._C import distCUDA2  # noqa: E402

DEV = torch.device("cuda")
_flush = None


def timed(fn, iters, warm=3, before=None):
    """median / min ms of fn(); `before` runs outside the event pair (restores inputs)."""
    global _flush
    if _flush is None:
        _flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    for _ in range(warm):
        if before:
            before()
        fn()
    ts = []
    for i in range(iters):
        if before:
            before()
        _flush.fill_(i & 0xFF)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts)), float(np.min(ts))


def ref_or_none():
    try:
        from oracle import ref_cuda
        return ref_cuda if ref_cuda.available() else None
    except Exception:
        return None


# ---------------------------------------------------------------------------------------------------------
def sort_sweep(quick):
    ref = ref_or_none()
    l = _lib.lib()
    rows = []
    tiles_bits = 13  # 1920x1080 -> 8160 tiles: 32 depth bits + 13 tile bits, as the rasteriser would ask
    bits = 32 + tiles_bits
    for lg in ([16, 20, 24] if quick else range(16, 27)):
        N = 1 << lg
        g = torch.Generator().manual_seed(lg)
        tile = torch.randint(0, 1 << tiles_bits, (N,), generator=g, dtype=torch.int64)
        depth = (torch.rand(N, generator=g) * 9.8 + 0.2).view(torch.int32).to(torch.int64)
        keys = ((tile << 32) | depth).to(DEV)
        vals = torch.arange(N, dtype=torch.int32, device=DEV)
        ka, va = keys.clone(), vals.clone()
        kb, vb = torch.empty_like(ka), torch.empty_like(va)
        ws = torch.zeros(max(l.ghr_sort_workspace_bytes(N), 256), dtype=torch.uint8, device=DEV)
        sel = torch.zeros(1, dtype=torch.int32, device=DEV)
        stream = torch.cuda.current_stream().cuda_stream

        def restore():
            ka.copy_(keys)
            va.copy_(vals)

        def ours():
            _lib.check(l.ghr_sort_pairs_u64(N, bits, ka.data_ptr(), kb.data_ptr(), va.data_ptr(), vb.data_ptr(),
                                            ws.data_ptr(), ws.numel(), sel.data_ptr(), stream), "sort")
        iters = 5 if lg >= 24 else 15
        med, best = timed(ours, iters, before=restore)
        row = {"N": N, "bits": bits, "ours_ms": round(med, 4), "ours_Mkeys_s": round(N / med / 1e3, 1)}
        # algorithmic bytes: histogram read 8N + per pass (12 read + 12 write)
        passes = math.ceil(bits / 8)
        row["ours_GBps_alg"] = round(N * (8 + 24 * passes) / (med * 1e-3) / 1e9, 1)
        if ref is not None:
            kc, vc, run = ref.cub_sort_pairs(keys, vals, bits)
            restore()
            ours()
            torch.cuda.synchronize()
            out_k, out_v = (kb, vb) if int(sel.item()) else (ka, va)
            row["bit_exact_vs_cub"] = bool(torch.equal(out_k, kc) and torch.equal(out_v, vc))
            medc, _ = timed(run, iters)
            row["cub_ms"] = round(medc, 4)
            row["cub_Mkeys_s"] = round(N / medc / 1e3, 1)
            row["speedup"] = round(medc / med, 3)
        rows.append(row)
        del keys, vals, ka, va, kb, vb, ws
    return rows


# ---------------------------------------------------------------------------------------------------------
def forward_fps_zju(quick):
    """LBS + covariance + SH + rasteriser forward (no autograd), 50k Gaussians, 512x512, per-view pose."""
    from mygauhuman_b200.lbs import SMPLDeformer
    from oracle import lbs_oracle as LO
    P, W, H = 50_000, 512, 512
    smpl = S.make_smpl_model(0)
    big = S.big_pose_params()
    tv = S.big_pose_vertices(smpl, big)
    g = S.make_gaussians(P, tv.numpy(), seed=0, sh_degree=3)
    L = {k: g[k].to(DEV) for k in g}
    bigd = {k: v.to(DEV) for k, v in big.items()}
    tvd = tv.to(DEV)
    deformer = SMPLDeformer(smpl, DEV)
    views = []
    for i in range(16):
        cam = {k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in S.make_camera(W, H, angle=2 * np.pi * i / 16, seed=i).items()}
        pose = S.make_pose_params(i)
        pose["Th"] = pose["Th"] * 0.2
        views.append((cam, {k: v.to(DEV) for k, v in pose.items()}))
    bg = torch.zeros(3, device=DEV)
    opac = torch.sigmoid(L["opacity_logit"])
    scal = torch.exp(L["log_scale"])
    it = [0]

    def ours():
        cam, pose = views[it[0] % len(views)]
        it[0] += 1
        with torch.no_grad():
            d = deformer.deform_gaussians(L["xyz"], pose, bigd, tvd, scaling=scal, rotation=L["rotation"],
                                          features=L["features"], sh_degree=3, campos=cam["camera_center"])
            rs = dgr.GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], bg, 1.0, cam["world_view_transform"],
                                                   cam["full_proj_transform"], 3, cam["camera_center"], False, False)
            return dgr.GaussianRasterizer(rs)(means3D=d["means3D"], means2D=None, opacities=opac, colors_precomp=d["colors"],
                                              cov3D_precomp=d["cov3D"])
    med, best = timed(ours, 30 if quick else 100, warm=20)
    out = {"P": P, "W": W, "H": H, "ours_ms": round(med, 4), "ours_fps": round(1e3 / med, 1),
           "what": "eager public API, torch.no_grad: SMPLDeformer.deform_gaussians + GaussianRasterizer, one view per call"}
    ref = ref_or_none()
    if ref is not None:
        rr = ref.RefRasterizer()
        smpl_d = {k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in smpl.items()}
        e = torch.Tensor([])

        def theirs():
            cam, pose = views[it[0] % len(views)]
            it[0] += 1
            with torch.no_grad():
                vid = torch.cdist(L["xyz"], tvd).argmin(1)
                o = LO.coarse_deform_c2source(smpl_d, L["xyz"], pose, bigd, tvd, vert_ids=vid)
                means3D, transforms = o[1], o[3]
                cov = LO.covariance(scal, 1.0, L["rotation"], transforms)
                col = LO.view_colors(3, L["features"], means3D, cam["camera_center"])
                return rr.forward(bg, means3D.contiguous(), col.contiguous(), opac, e, e, 1.0, cov.contiguous(), cam["world_view_transform"], cam["full_proj_transform"],
                                  cam["tanfovx"], cam["tanfovy"], H, W, e, 3, cam["camera_center"])
        try:
            medr, _ = timed(theirs, 20 if quick else 50, warm=5)
            out.update(ref_ms=round(medr, 4), ref_fps=round(1e3 / medr, 1), speedup=round(medr / med, 2),
                       ref_what="PyTorch LBS restatement on the GPU + the reference's unmodified CUDA rasteriser forward")
        except Exception as ex:  # oracle signature drift must not hide our own number
            out["ref_error"] = repr(ex)[:200]
    return out


# ---------------------------------------------------------------------------------------------------------
def stress_1080p(quick):
    """3M Gaussians in a 10 m box, 1920x1080, SH degree 3 evaluated by the rasteriser (baking.py-style call)."""
    P, W, H = (1_000_000 if quick else 3_000_000), 1920, 1080
    g = S.make_scene_cloud(P, seed=0)
    cam = S.make_camera(W, H, angle=math.pi, radius=6.0, target=(0.0, 0.0, 6.0), elev=0.0, seed=0, focal=1100.0)
    cam = {k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in cam.items()}
    xyz, sh = g["xyz"].to(DEV), g["features"].to(DEV).contiguous()
    opac, scal, rot = torch.sigmoid(g["opacity_logit"]).to(DEV), torch.exp(g["log_scale"]).to(DEV), g["rotation"].to(DEV)
    bg = torch.zeros(3, device=DEV)
    e = torch.Tensor([])
    args = (bg, xyz, e, opac, scal, rot, 1.0, e, cam["world_view_transform"], cam["full_proj_transform"], cam["tanfovx"],
            cam["tanfovy"], H, W, sh, 3, cam["camera_center"], False, False)
    res = _C.rasterize_gaussians(*args)
    n = int(res[0])
    vis = int((res[4] > 0).sum())
    med, best = timed(lambda: _C.rasterize_gaussians(*args), 10 if quick else 30)
    out = {"P": P, "W": W, "H": H, "num_rendered": n, "visible": vis, "ours_ms": round(med, 3), "ours_fps": round(1e3 / med, 1)}
    _lib.profile_enable(True)
    _lib.profile_read()
    for _ in range(5):
        _C.rasterize_gaussians(*args)
    torch.cuda.synchronize()
    out["stages_us"] = {k: round(1e3 * ms / c, 1) for k, (ms, c) in _lib.profile_read().items() if c}
    _lib.profile_enable(False)
    ref = ref_or_none()
    if ref is not None:
        rr = ref.RefRasterizer()
        rf = lambda: rr.forward(bg, xyz, e, opac, scal, rot, 1.0, e, cam["world_view_transform"], cam["full_proj_transform"],
                                cam["tanfovx"], cam["tanfovy"], H, W, sh, 3, cam["camera_center"])
        n_ref = rf()[0]
        medr, _ = timed(rf, 10 if quick else 30)
        out.update(ref_ms=round(medr, 3), ref_fps=round(1e3 / medr, 1), speedup=round(medr / med, 2),
                   num_rendered_matches=bool(int(n_ref) == n))
    return out


# ---------------------------------------------------------------------------------------------------------
def knn(quick):
    ref = ref_or_none()
    rows = []
    smpl = S.make_smpl_model(0)
    for P in (6890, 50_000, 300_000):
        if P == 6890:
            pts = smpl["v_template"].float()
        else:
            pts = S.make_gaussians(P, smpl["v_template"].numpy(), seed=1)["xyz"]
        pts = pts.to(DEV).contiguous()
        med, _ = timed(lambda: distCUDA2(pts), 10 if quick else 30)
        row = {"P": P, "ours_ms": round(med, 4)}
        if ref is not None:
            want = ref.dist_cuda2(pts)
            got = distCUDA2(pts)
            row["max_rel_err_vs_ref"] = float(((got - want).abs() / want.abs().clamp_min(1e-20)).max())
            medr, _ = timed(lambda: ref.dist_cuda2(pts), 5 if quick else 15)
            row.update(ref_ms=round(medr, 4), speedup=round(medr / med, 2))
        rows.append(row)
    return rows


# ---------------------------------------------------------------------------------------------------------
def render7(quick):
    """Rasteriser forward + backward of the reference's seven attribute passes (gaussian_renderer/__init__.py:203-272),
    50k Gaussians on the posed template, 512x512, colours and covariances precomputed (what render() feeds it)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from tests import common as Cm
    P, W, H, NA = 50_000, 512, 512, 6
    smpl = S.make_smpl_model(0)
    tv = S.big_pose_vertices(smpl, S.big_pose_params())
    tc = Cm.to_torch(Cm.raster_case(P, W, H, seed=0, mode="precomp", cloud=tv.numpy()), "cuda")
    gen = torch.Generator().manual_seed(0)
    attrs = [torch.rand(P, 3, generator=gen).to(DEV).requires_grad_(True) for _ in range(NA)]
    wts = [torch.randn(3, H, W, generator=gen).to(DEV) for _ in range(NA + 1)]
    wd, wa = torch.randn(1, H, W, generator=gen).to(DEV), torch.randn(1, H, W, generator=gen).to(DEV)
    lv = {k: tc[k].clone().requires_grad_(True) for k in ("means3D", "opacity", "colors", "cov3D")}
    rs = dgr.GaussianRasterizationSettings(H, W, tc["tanfovx"], tc["tanfovy"], tc["bg"], 1.0, tc["view"], tc["proj"], 3,
                                           tc["campos"], False, False)
    rast = dgr.GaussianRasterizer(rs)
    geo = dict(means3D=lv["means3D"], means2D=torch.zeros(P, 3, device=DEV, requires_grad=True), opacities=lv["opacity"], cov3D_precomp=lv["cov3D"])

    def clear():
        for t in list(lv.values()) + attrs:
            t.grad = None

    def multi():
        color, radii, depth, alpha, extra = rast.forward_multi(extra_colors=attrs, colors_precomp=lv["colors"], **geo)
        loss = (depth * wd).sum() + (alpha * wa).sum()
        for im, w in zip([color] + extra, wts):
            loss = loss + (im * w).sum()
        loss.backward()

    def separate():
        color, radii, depth, alpha = rast(colors_precomp=lv["colors"], **geo)
        loss = (depth * wd).sum() + (alpha * wa).sum() + (color * wts[0]).sum()
        for a, w in zip(attrs, wts[1:]):
            loss = loss + (rast(colors_precomp=a, **geo)[0] * w).sum()
        loss.backward()

    n = 10 if quick else 40
    med_m, _ = timed(multi, n, warm=5, before=clear)
    med_s, _ = timed(separate, n, warm=5, before=clear)
    e0 = torch.Tensor([])
    det = [a.detach() for a in attrs]

    def direct():  # the drop-in _C entry points without autograd, as the reference arm below is timed
        fw = _C.rasterize_gaussians_multi(tc["bg"], tc["means3D"], tc["colors"], tc["opacity"], e0, e0, 1.0, tc["cov3D"],
                                          tc["view"], tc["proj"], tc["tanfovx"], tc["tanfovy"], H, W, e0, 3, tc["campos"],
                                          False, False, det)
        n_, color, depth, alpha, radii, geom, binning, img, images, first, packed = fw
        _C.rasterize_gaussians_multi_backward(tc["bg"], tc["means3D"], radii, tc["colors"], e0, e0, 1.0, tc["cov3D"],
                                              tc["view"], tc["proj"], tc["tanfovx"], tc["tanfovy"], wts[0], wd, wa, e0, 3,
                                              tc["campos"], geom, n_, binning, img, alpha, False, first, packed, det,
                                              wts[1:])
    med_d, _ = timed(direct, n, warm=5)
    _lib.profile_enable(True)
    _lib.profile_read()
    for i in range(10):
        clear()
        _flush.fill_(i)
        multi()
    stages = {k: round(1e3 * ms / max(c, 1), 1) for k, (ms, c) in _lib.profile_read().items() if c}
    _lib.profile_enable(False)
    out = {"P": P, "W": W, "H": H, "attributes": NA + 1, "forward_multi_ms": round(med_m, 4),
           "forward_multi_direct_ms": round(med_d, 4),
           "direct_what": "_C.rasterize_gaussians_multi + _multi_backward called directly (no autograd, no loss kernels)",
           "forward_multi_stage_us": stages,
           "one_call_per_attribute_ms": round(med_s, 4), "multi_vs_separate": round(med_s / med_m, 2),
           "what": "rasteriser forward + backward through torch autograd (linear loss on every image), L2 flushed"}
    ref = ref_or_none()
    if ref is not None:
        e = torch.Tensor([])
        rrs = [ref.RefRasterizer() for _ in range(NA + 1)]
        zero1 = torch.zeros(1, H, W, device=DEV)
        cols = [tc["colors"]] + [a.detach() for a in attrs]

        def theirs():
            for k, (rr, c) in enumerate(zip(rrs, cols)):
                n_, color, depth, alpha, radii = rr.forward(tc["bg"], tc["means3D"], c, tc["opacity"], e, e, 1.0, tc["cov3D"],
                                                            tc["view"], tc["proj"], tc["tanfovx"], tc["tanfovy"], H, W, e, 3,
                                                            tc["campos"])
                rr.backward(tc["bg"], tc["means3D"], radii, c, e, e, 1.0, tc["cov3D"], tc["view"], tc["proj"], tc["tanfovx"],
                            tc["tanfovy"], wts[k], wd if k == 0 else zero1, wa if k == 0 else zero1, e, 3, tc["campos"], alpha)
        try:
            med_r, _ = timed(theirs, 5 if quick else 20, warm=3)
            out.update(ref_ms=round(med_r, 4), speedup_vs_ref=round(med_r / med_m, 2),
                       direct_speedup_vs_ref=round(med_r / med_d, 2),
                       ref_what="seven forward+backward calls of the reference's unmodified CUDA rasteriser (no autograd overhead)")
        except Exception as ex:
            out["ref_error"] = repr(ex)[:200]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    parts = {"sort_sweep": sort_sweep, "forward_fps_zju_512": forward_fps_zju, "stress_1080p_forward": stress_1080p,
             "dist_cuda2": knn, "render7_zju_512": render7}
    out = {"device": torch.cuda.get_device_name(0), "ref_cuda_available": ref_or_none() is not None}
    for k, f in parts.items():
        if a.only and k not in a.only.split(","):
            continue
        try:
            out[k] = f(a.quick)
        except Exception as ex:
            import traceback
            out[k] = {"error": repr(ex)[:300], "trace": traceback.format_exc()[-600:]}
        torch.cuda.empty_cache()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
