#!/usr/bin/env python
"""Benchmark of the GauHuman render hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference|reference-cpu]
                    [--workload zju_512|monocap_1024|multiview_300k|stress_1080p]

One "step" = one pass of the hot path over one camera view: (LBS deformation +
covariance + SH when --lbs is on) -> rasteriser forward -> loss ->
rasteriser backward, through the public operator API (GaussianRasterizer).
Prints ONE JSON line on rank 0 (contract in DESIGN.md "Measurement").

Timing: every step is bracketed by its own CUDA-event pair on the launching
stream; between steps (outside the events) a 256 MiB buffer is overwritten to
flush the 126 MB L2.  The K-step region is bracketed by barrier + synchronize on
both sides; ms_per_step is the max over ranks.

`--impl reference` times the UNMODIFIED reference CUDA rasteriser
(oracle/_ref/libref_gauhuman.so, built from /root/reference by
oracle/build_ref.sh) on the same GPU, same inputs, same Python-level protocol.
`--impl reference-cpu` times the CPU oracle port on the host cores.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch

METRIC = "lbs_raster_fwd_bwd_iters_per_s"
UNIT = "iters/s"

WORKLOADS = {
    # name: (P, W, H, views, description)
    "zju_512": dict(P=50_000, W=512, H=512, desc="ZJU-MoCap-refine shape: 50k Gaussians, 512x512, SH deg 3, fwd+bwd with alpha mask"),
    "monocap_1024": dict(P=100_000, W=1024, H=1024, desc="MonoCap shape: 100k Gaussians, 1024x1024, fwd+bwd"),
    "multiview_300k": dict(P=300_000, W=1024, H=1024, desc="multi-view batch: 64 cameras x 300k Gaussians, view-sharded"),
    "stress_1080p": dict(P=3_000_000, W=1920, H=1080, desc="stress render: 3M Gaussians 1920x1080"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-cpu"])
    ap.add_argument("--workload", default="zju_512", choices=list(WORKLOADS))
    ap.add_argument("--views", type=int, default=64, help="distinct cameras cycled through (sharded across ranks)")
    ap.add_argument("--cpu-iters", type=int, default=20, help="CPU-oracle iterations for cpu_baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--host-profile", action="store_true", help="cProfile of 50 un-synchronised steps to stderr")
    ap.add_argument("--mode", default="graph", choices=["graph", "eager"],
                    help="ours: replay the step as one CUDA graph (mygauhuman_b200.graphed.GraphedRenderStep, default) "
                         "or issue it kernel by kernel through the eager operator API")
    ap.add_argument("--batch-views", type=int, default=8,
                    help="views each rank renders per optimisation step (multi-view batch): gradients are accumulated over "
                         "them and summed over the ranks ONCE per step (NCCL).  1 = an all-reduce after every view")
    ap.add_argument("--lanes", type=int, default=8,
                    help="graph mode: views of a batch replayed concurrently on this many CUDA streams "
                         "(MultiLaneRenderStep); reduced to a divisor of --batch-views; 1 = one view at a time")
    ap.add_argument("--sync-free", action="store_true",
                    help="ours: size the binning buffer from the previous step instead of a blocking read")
    return ap.parse_args()


# ---------------------------------------------------------------------------
# distributed plumbing
# ---------------------------------------------------------------------------

def init_dist(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout to the one JSON line
        dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo", rank=rank, world_size=world,
                                device_id=torch.device("cuda", local) if torch.cuda.is_available() else None)
    return world, rank, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


# ---------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ---------------------------------------------------------------------------

class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 7:
                    continue
                try:
                    sm.append(float(f[0])); mx.append(float(f[1]))
                except ValueError:
                    continue
                for n, v in zip(names, f[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


# ---------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------

PACK = dict(view=(0, 16), proj=(16, 32), campos=(32, 35), poses=(35, 107), shapes=(107, 117), R=(117, 126), Th=(126, 129))
PACK_FLOATS = 129


def build_workload(args, rank, world):
    """Seeded synthetic inputs (mygauhuman_b200/synthetic.py).  Host tensors.
    Canonical Gaussians sit on the big-pose SMPL-topology surface; every view has its own camera AND its own
    body pose (as a ZJU-MoCap frame does)."""
    from mygauhuman_b200 import synthetic as S
    wl = WORKLOADS[args.workload]
    P, W, H = wl["P"], wl["W"], wl["H"]
    smpl = S.make_smpl_model(0)
    big = S.big_pose_params()
    t_vertices = S.big_pose_vertices(smpl, big)
    g = S.make_gaussians(P, t_vertices.numpy(), seed=0, sh_degree=3)
    n_views = max(args.views, world)
    views = []
    for i in range(n_views):
        cam = S.make_camera(W, H, angle=2 * np.pi * i / n_views, radius=3.0, seed=i)
        pose = S.make_pose_params(i)
        pose["Th"] = pose["Th"] * 0.2  # keep the body inside the frustum of the ring cameras
        views.append(dict(cam=cam, pose=pose))
    views = views[rank::world]  # view sharding: rank r renders views r, r+world, ...
    leaves = {"xyz": g["xyz"], "log_scale": g["log_scale"], "rotation": g["rotation"], "features": g["features"],
              "opacity_logit": g["opacity_logit"], "normals": g["normals"]}
    gen = torch.Generator().manual_seed(1234 + rank)
    targets = torch.randn(5, H, W, generator=gen)  # linear loss weights on colour (3), depth, alpha
    return dict(P=P, W=W, H=H, smpl=smpl, big=big, t_vertices=t_vertices, leaves=leaves, views=views, targets=targets,
                bg=torch.zeros(3))


def view_pack(view):
    """[view 16 | proj 16 | campos 3 | poses 72 | shapes 10 | R 9 | Th 3] as one pinned host vector: the
    per-step input besides the target planes."""
    cam, pose = view["cam"], view["pose"]
    v = torch.cat([cam["world_view_transform"].reshape(-1), cam["full_proj_transform"].reshape(-1),
                   cam["camera_center"].reshape(-1), pose["poses"].reshape(-1), pose["shapes"].reshape(-1),
                   pose["R"].reshape(-1), pose["Th"].reshape(-1)]).float()
    assert v.numel() == PACK_FLOATS
    return v.pin_memory() if torch.cuda.is_available() else v


def unpack(vec):
    g = lambda k: vec[PACK[k][0]:PACK[k][1]]
    cam = dict(view=g("view").view(4, 4), proj=g("proj").view(4, 4), campos=g("campos"))
    pose = dict(poses=g("poses").view(1, 72), shapes=g("shapes").view(1, 10), R=g("R").view(3, 3), Th=g("Th").view(1, 3))
    return cam, pose


def linear_loss(color, depth, alpha, targets):
    """sum(colour*t[0:3]) + sum(depth*t[3]) + sum(alpha*t[4]): every output plane gets a dense random gradient."""
    return (torch.cat([color, depth, alpha]) * targets).sum()


# ---------------------------------------------------------------------------
# the two GPU arms
# ---------------------------------------------------------------------------

class OursArm:
    """The product: fused LBS+covariance+SH (mygauhuman_b200.lbs) -> rasteriser
    (mygauhuman_b200.diff_gaussian_rasterization), all on libghr.so through the C ABI."""
    name = "ours"

    def __init__(self, wl, dev, sync_free, graph=False, lanes=1):
        from mygauhuman_b200 import _lib
        from mygauhuman_b200 import diff_gaussian_rasterization as dgr
        from mygauhuman_b200.lbs import SMPLDeformer
        self._lib, self.dgr, self.wl, self.dev = _lib, dgr, wl, dev
        _lib.lib()  # fail loudly if the CUDA library is missing
        self.deformer = SMPLDeformer(wl["smpl"], dev)
        self.leaves = {k: v.to(dev).requires_grad_(True) for k, v in wl["leaves"].items()}
        self.big = {k: v.to(dev) for k, v in wl["big"].items()}
        self.t_vertices = wl["t_vertices"].to(dev)
        self.bg = wl["bg"].to(dev)
        self.sync_free = sync_free
        self.last_n = None
        self.graphed, self.use_graph, self.multi = None, False, None
        if graph:
            from mygauhuman_b200.graphed import GraphedRenderStep
            cam = wl["views"][0]["cam"]
            assert all(v["cam"]["tanfovx"] == cam["tanfovx"] and v["cam"]["tanfovy"] == cam["tanfovy"] for v in wl["views"])
            self.multi = None
            if lanes > 1:
                from mygauhuman_b200.graphed import MultiLaneRenderStep
                deformers = [self.deformer] + [SMPLDeformer(wl["smpl"], dev) for _ in range(lanes - 1)]
                self.multi = MultiLaneRenderStep(deformers, self.leaves, self.big, self.t_vertices, wl["H"], wl["W"],
                                                 cam["tanfovx"], cam["tanfovy"], self.bg, linear_loss, sh_degree=3)
                self.graphed = self.multi.lanes[0]  # single-view calls (warm-up, verification) go through lane 0
            else:
                self.graphed = GraphedRenderStep(self.deformer, self.leaves, self.big, self.t_vertices, wl["H"], wl["W"],
                                                 cam["tanfovx"], cam["tanfovy"], self.bg, linear_loss, sh_degree=3,
                                                 accumulate=True)

    def capture(self, vecs, targets_dev):
        if self.multi is not None:
            self.multi.capture(vecs, [targets_dev])
        else:
            self.graphed.capture(vecs, [targets_dev])

    def step(self, view, vec_dev, targets_dev):
        dgr, L = self.dgr, self.leaves
        if self.graphed is not None and self.use_graph:
            loss = self.graphed(vec_dev, targets_dev)
            return loss, [L[k].grad for k in ("xyz", "features", "opacity_logit", "log_scale", "rotation")]
        camd, pose = unpack(vec_dev)
        cam = view["cam"]
        d = self.deformer.deform_gaussians(L["xyz"], pose, self.big, self.t_vertices, normals=L["normals"],
                                           scaling=torch.exp(L["log_scale"]), rotation=L["rotation"],
                                           scaling_modifier=1.0, features=L["features"], sh_degree=3,
                                           campos=camd["campos"])
        if self.sync_free:
            dgr._C.set_sync_free_capacity(None if self.last_n is None else int(self.last_n * 1.3) + 4096)
        rs = dgr.GaussianRasterizationSettings(
            image_height=self.wl["H"], image_width=self.wl["W"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
            bg=self.bg, scale_modifier=1.0, viewmatrix=camd["view"], projmatrix=camd["proj"], sh_degree=3,
            campos=camd["campos"], prefiltered=False, debug=False)
        means2D = torch.zeros_like(d["means3D"], requires_grad=True)
        color, radii, depth, alpha = dgr.GaussianRasterizer(rs)(
            means3D=d["means3D"], means2D=means2D, opacities=torch.sigmoid(L["opacity_logit"]),
            colors_precomp=d["colors"], cov3D_precomp=d["cov3D"])
        loss = linear_loss(color, depth, alpha, targets_dev)
        for t in L.values():
            t.grad = None
        loss.backward()
        return loss, [L[k].grad for k in ("xyz", "features", "opacity_logit", "log_scale", "rotation")]

    def measure_NV(self, view, vec_dev):
        """(num_rendered, visible Gaussians) of one view, for the algorithmic-byte formulas."""
        camd, pose = unpack(vec_dev)
        cam, L, e = view["cam"], self.leaves, torch.Tensor([])
        with torch.no_grad():
            d = self.deformer.deform_gaussians(L["xyz"], pose, self.big, self.t_vertices,
                                               scaling=torch.exp(L["log_scale"]), rotation=L["rotation"],
                                               features=L["features"], sh_degree=3, campos=camd["campos"])
            out = self.dgr._C.rasterize_gaussians(self.bg, d["means3D"], d["colors"], torch.sigmoid(L["opacity_logit"]),
                                                  e, e, 1.0, d["cov3D"], camd["view"], camd["proj"], cam["tanfovx"],
                                                  cam["tanfovy"], self.wl["H"], self.wl["W"], e, 3, camd["campos"],
                                                  False, False)
        return int(out[0]), int((out[4] > 0).sum())


class _RefRasterize(torch.autograd.Function):
    """Same protocol as the reference's _RasterizeGaussians (dgr __init__.py:44-158) around the reference's own
    native forward/backward (through oracle/ref_shim.cu instead of its torch binding); precomputed colour +
    covariance inputs, which is how the reference's render() calls it."""

    @staticmethod
    def forward(ctx, ref, means3D, means2D, colors, opacities, cov3D, bg, view, proj, campos, tfx, tfy, H, W):
        n, color, depth, alpha, radii = ref.forward(bg, means3D, colors, opacities, None, None, 1.0, cov3D, view, proj,
                                                    tfx, tfy, H, W, None, 3, campos)
        ctx.ref, ctx.cfg = ref, (bg, view, proj, campos, tfx, tfy)
        ctx.save_for_backward(means3D, colors, cov3D, radii, alpha)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, gc, gr, gd, ga):
        means3D, colors, cov3D, radii, alpha = ctx.saved_tensors
        bg, view, proj, campos, tfx, tfy = ctx.cfg
        g = ctx.ref.backward(bg, means3D, radii, colors, None, None, 1.0, cov3D, view, proj, tfx, tfy, gc, gd, ga, None,
                             3, campos, alpha)
        return (None, g["dL_dmeans3D"], g["dL_dmeans2D"], g["dL_dcolors"], g["dL_dopacity"], g["dL_dcov3D"]) + (None,) * 8


class ReferenceArm:
    """The reference's path on the same GPU: its PyTorch LBS / covariance / eval_sh (the restatement in
    oracle/lbs_oracle.py run on CUDA tensors -- the reference module itself cannot travel to the GPU box; knn_cuda
    is replaced by torch.cdist + argmin as SURVEY 8c plans) followed by the UNMODIFIED reference CUDA rasteriser
    built for sm_100 (oracle/_ref).  None of libghr's kernels are on this path."""
    name = "reference"

    def __init__(self, wl, dev):
        from oracle import lbs_oracle, ref_cuda
        if not ref_cuda.available():
            raise FileNotFoundError(ref_cuda.LIB_PATH)
        self.LO = lbs_oracle
        self.ref = ref_cuda.RefRasterizer()
        self.wl, self.dev = wl, dev
        self.smpl = {k: v.to(dev) for k, v in wl["smpl"].items()}
        self.leaves = {k: v.to(dev).requires_grad_(True) for k, v in wl["leaves"].items()}
        self.big = {k: v.to(dev) for k, v in wl["big"].items()}
        self.t_vertices = wl["t_vertices"].to(dev)
        self.bg = wl["bg"].to(dev)

    def step(self, view, vec_dev, targets_dev):
        L, LO = self.leaves, self.LO
        camd, pose = unpack(vec_dev)
        cam = view["cam"]
        with torch.no_grad():
            vid = torch.cdist(L["xyz"], self.t_vertices).argmin(1)  # stand-in for knn_cuda KNN(k=1)
        _, world, _, T, _, wn, _ = LO.coarse_deform_c2source(self.smpl, L["xyz"], pose, self.big, self.t_vertices,
                                                             normals=L["normals"] / L["normals"].norm(dim=1, keepdim=True),
                                                             vert_ids=vid)
        cov = LO.covariance(torch.exp(L["log_scale"]), 1.0, L["rotation"], T)
        col = LO.view_colors(3, L["features"], world, camd["campos"])
        means2D = torch.zeros_like(world, requires_grad=True)
        color, radii, depth, alpha = _RefRasterize.apply(
            self.ref, world, means2D, col, torch.sigmoid(L["opacity_logit"]), cov, self.bg, camd["view"].contiguous(),
            camd["proj"].contiguous(), camd["campos"].contiguous(), cam["tanfovx"], cam["tanfovy"], self.wl["H"],
            self.wl["W"])
        loss = linear_loss(color, depth, alpha, targets_dev)
        for t in L.values():
            t.grad = None
        loss.backward()
        return loss, [L[k].grad for k in ("xyz", "features", "opacity_logit", "log_scale", "rotation")]


# ---------------------------------------------------------------------------
# CPU baseline (the oracle port on the host cores) -- reported, not the target
# ---------------------------------------------------------------------------

def cpu_baseline(wl, view, iters):
    """One full step (LBS + covariance + SH -> raster fwd -> raster bwd -> LBS bwd) on the host cores:
    the reference's PyTorch LBS math (oracle/lbs_oracle.py, torch CPU, autograd for its backward) and the C
    restatement of the rasteriser (oracle/raster_oracle.c, OpenMP)."""
    import oracle as O
    from oracle import lbs_oracle as LO
    cam, pose = view["cam"], view["pose"]
    a = lambda t: np.ascontiguousarray(t.detach().cpu().numpy(), dtype=np.float32)
    vm, pm, campos = a(cam["world_view_transform"]), a(cam["full_proj_transform"]), a(cam["camera_center"])
    tg = a(wl["targets"])
    bg = a(wl["bg"])
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ts = []
    for _ in range(max(1, iters)):
        L = {k: v.clone().requires_grad_(True) for k, v in wl["leaves"].items()}
        t0 = time.perf_counter()
        with torch.no_grad():
            vid = torch.cdist(L["xyz"], wl["t_vertices"]).argmin(1)
        _, world, _, T, _, wn, _ = LO.coarse_deform_c2source(wl["smpl"], L["xyz"], pose, wl["big"], wl["t_vertices"],
                                                             normals=L["normals"], vert_ids=vid)
        cov = LO.covariance(torch.exp(L["log_scale"]), 1.0, L["rotation"], T)
        col = LO.view_colors(3, L["features"], world, cam["camera_center"])
        opa = torch.sigmoid(L["opacity_logit"])
        st = O.raster_forward(bg, a(world), a(col), a(opa), None, None, 1.0, a(cov), vm, pm, cam["tanfovx"],
                              cam["tanfovy"], wl["H"], wl["W"], None, 3, campos)
        g = O.raster_backward(st, bg, a(world), a(col), None, None, 1.0, a(cov), vm, pm, cam["tanfovx"], cam["tanfovy"],
                              tg[0:3], tg[3:4], tg[4:5], None, 3, campos)
        torch.autograd.backward([world, cov, col, opa],
                                [torch.from_numpy(g["dL_dmeans3D"]), torch.from_numpy(g["dL_dcov3D"]),
                                 torch.from_numpy(g["dL_dcolors"]), torch.from_numpy(g["dL_dopacity"])])
        ts.append(time.perf_counter() - t0)
    best = min(ts)
    return {"value": 1.0 / best, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{len(ts)} full steps of the workload (view 0): torch-CPU LBS/cov/SH restatement with autograd + "
                      f"oracle/raster_oracle.c (C, OpenMP) fwd+bwd; best of {len(ts)}; {sum(ts):.1f} s CPU wall"}


# ---------------------------------------------------------------------------
# roofline of the dominant kernel
# ---------------------------------------------------------------------------

def stage_bytes(stage, P, V, N, X, tiles, passes):
    """ALGORITHMIC bytes per launch (DESIGN.md 'Kernels')."""
    return {
        "preprocess": 48 * P + 28 * V,  # precomputed colour + covariance inputs
        "lbs_fwd": 453 * P,             # fused LBS + covariance + SH deg 3 (ghr_lbs_deform.cu header)
        "lbs_bwd": (453 - 97 + 96 + 12 + 12 + 12 + 16 + 192) * P,
        "emit_keys": 8 * P + 12 * V + 12 * N,
        "sort": N * (8 + 24 * passes),
        "tile_ranges": 8 * N + 8 * tiles,
        "blend_fwd": 44 * N + 28 * X,
        "blend_bwd": 44 * N + 28 * X + 36 * V,
        "preprocess_bwd": 132 * P,
    }.get(stage)


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def load_traffic(stage):
    """dram bytes per launch of the stage's kernel from the committed ncu --set full summary, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(stage)
    except Exception:
        return None


# ---------------------------------------------------------------------------
# main
# ---------------------------------------------------------------------------

def run_reference_cpu(args, world, rank):
    """--impl reference-cpu: the CPU oracle port, rank 0 only."""
    if rank != 0:
        return
    wl = build_workload(args, 0, 1)
    res = cpu_baseline(wl, wl["views"][0], args.steps if args.steps <= 10 else 3)
    line = {"impl": "reference-cpu", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": 0,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / res["value"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "desc": WORKLOADS[args.workload]["desc"]}, "cpu_baseline": res,
            "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line goes to the process's real stdout; everything else any library prints goes to stderr."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    args = parse()
    # NCCL (version banner), nvidia-smi children and the like write to fd 1 behind Python's back: point fd 1 at stderr for
    # the whole run and keep the original for the result line
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    world, rank, local = init_dist(args)
    if args.impl == "reference-cpu":
        run_reference_cpu(args, world, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device. The hot path has no CPU fallback "
                         "(use --impl reference-cpu for the CPU oracle baseline).")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    wl = build_workload(args, rank, world)
    P, W, H = wl["P"], wl["W"], wl["H"]
    X, tiles = W * H, ((W + 15) // 16) * ((H + 15) // 16)
    if args.impl == "ours":
        lanes = math.gcd(max(1, args.lanes), max(1, args.batch_views)) if args.mode == "graph" else 1
        arm = OursArm(wl, dev, args.sync_free, graph=(args.mode == "graph"), lanes=lanes)
    else:
        try:
            arm = ReferenceArm(wl, dev)
        except (FileNotFoundError, OSError) as e:
            if rank == 0:
                emit({"impl": "reference", "unavailable": f"oracle/_ref not built: {e}"})
            return
    views = wl["views"]
    vec_host = [view_pack(v) for v in views]
    vec_dev = [v.to(dev) for v in vec_host]
    targets_host = wl["targets"].pin_memory()
    targets_dev = targets_host.to(dev)
    vec_stage = torch.empty(PACK_FLOATS, device=dev)
    targets_stage = torch.empty_like(targets_dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    if world > 1:
        import torch.distributed as dist
    from mygauhuman_b200.multiview import GradientAllReducer
    reducer = GradientAllReducer()

    ar_err = None

    B = max(1, args.batch_views)

    def graph_on():
        return args.impl == "ours" and arm.use_graph

    def begin_batch():
        if graph_on():
            arm.graphed.zero_grads()  # the graphed step accumulates the batch's gradients in its flat bucket

    def end_batch(grads):
        # view-sharded data parallelism: ONE sum of the Gaussian-parameter gradients over the ranks per batch
        if graph_on():
            arm.graphed.reduce_grads()
        elif world > 1:
            reducer(grads)

    def one_view(i, e2e):
        k = i % len(views)
        if graph_on():
            # the graphed step owns the static input buffers: it copies from pinned host memory itself (e2e), or from
            # the resident device copies (the target planes already sit in its static buffer)
            if e2e:
                loss, grads = arm.step(views[k], vec_host[k], targets_host)
            else:
                loss, grads = arm.step(views[k], vec_dev[k], arm.graphed.loss_inputs[0])
        elif e2e:
            vec_stage.copy_(vec_host[k], non_blocking=True)
            targets_stage.copy_(targets_host, non_blocking=True)
            loss, grads = arm.step(views[k], vec_stage, targets_stage)
        else:
            loss, grads = arm.step(views[k], vec_dev[k], targets_dev)
        return loss, grads

    def one_step(i, e2e, want_grads=False):
        """A batch of one view (warm-up, verification)."""
        begin_batch()
        loss, grads = one_view(i, e2e)
        end_batch(grads)
        if e2e:
            return float(loss.item())
        return grads if want_grads else None

    def timed_lanes(K, e2e):
        """K steps of B views, the views of a batch replayed `lanes` at a time on as many streams (MultiLaneRenderStep).
        One event pair per GROUP of concurrent views on the stream the lanes fork from and join into (L2 flushed before
        every group, outside the pair); the first group's pair of a batch includes the bucket reset, the last one's the
        fold of the lanes' buckets and the all-reduce.  e2e: every view's inputs come from pinned host memory (the next
        group's copies are issued right after this group's replays, on the lanes' copy streams) and every view's loss
        goes back to pinned memory, consumed one group later -- all inside the timed region."""
        ml = arm.multi
        Kl = ml.K
        groups = K * (B // Kl)
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(groups)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(groups)]
        nv = len(views)
        pending, host_losses = [], 0.0
        torch.cuda.synchronize(); barrier(world); torch.cuda.synchronize()
        if e2e:
            for j in range(Kl):
                ml.lanes[j].prefetch(vec_host[j % nv], targets_host)
        i = 0
        for gi in range(groups):
            flush.fill_(gi & 0xFF)  # L2 flush, outside the event pair
            starts[gi].record()
            if (gi * Kl) % B == 0:
                ml.zero_grads()
            if e2e:
                nxt = [(i + Kl + j) % nv for j in range(Kl)]
                ml.run_group([None] * Kl, prefetched=True, next_vecs=[vec_host[k] for k in nxt],
                             next_loss_inputs=[(targets_host,)] * Kl)
            else:
                ml.run_group([vec_dev[(i + j) % nv] for j in range(Kl)])
            if ((gi + 1) * Kl) % B == 0:
                ml.reduce_grads()
            if e2e:
                new_pending = [ml.lanes[j].loss_to_host() for j in range(Kl)]
                for slot, ev in pending:  # the previous group's losses: their copies finished long ago
                    ev.synchronize()
                    host_losses += float(slot[0])
                pending = new_pending
            ends[gi].record()
            i += Kl
        for slot, ev in pending:
            ev.synchronize()
            host_losses += float(slot[0])
        torch.cuda.synchronize(); barrier(world); torch.cuda.synchronize()
        ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def timed(K, e2e):
        """K steps of B views.  Every view has its own event pair (L2 flushed before it, outside the pair); the first
        view's pair includes the bucket reset, the last one's the all-reduce.  e2e: inputs from pinned host memory and
        the view's loss read back.  With the graphed step the e2e loop is pipelined the way a training loop would be:
        the inputs of view i+1 travel on a copy stream while view i computes, and the loss of view i is copied to
        pinned memory asynchronously and consumed one view later -- every copy still happens inside the timed region."""
        if graph_on() and arm.multi is not None:
            return timed_lanes(K, e2e)
        n = K * B
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
        piped = e2e and graph_on()
        pending, host_losses = None, 0.0
        torch.cuda.synchronize(); barrier(world); torch.cuda.synchronize()
        if piped:
            arm.graphed.prefetch(vec_host[0], targets_host)
        for i in range(n):
            flush.fill_(i & 0xFF)  # L2 flush, outside the event pair
            starts[i].record()
            if i % B == 0:
                begin_batch()
            if piped:
                loss = arm.graphed.run_prefetched()
                grads = None
                arm.graphed.prefetch(vec_host[(i + 1) % len(views)], targets_host)
            else:
                loss, grads = one_view(i, e2e)
            if i % B == B - 1:
                end_batch(grads)
            if piped:
                nxt = arm.graphed.loss_to_host()
                if pending is not None:  # the previous view's loss: its copy finished long ago
                    pending[1].synchronize()
                    host_losses += float(pending[0][0])
                pending = nxt
            elif e2e:
                host_losses += float(loss.item())  # the device->host read of the view's result
            ends[i].record()
        if pending is not None:
            pending[1].synchronize()
            host_losses += float(pending[0][0])
        torch.cuda.synchronize(); barrier(world); torch.cuda.synchronize()
        ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    K, Wm = args.steps, max(args.warmup, 3)
    for i in range(Wm):
        one_step(i, False)
        one_step(i, True)
    torch.cuda.synchronize()
    graphed = args.impl == "ours" and args.mode == "graph"
    if graphed:
        # capacity sized from every view this rank cycles through, then one capture; warm the replay path too
        arm.capture(vec_dev, targets_dev)
        arm.use_graph = True
        for i in range(Wm):
            one_step(i, False)
            one_step(i, True)
        torch.cuda.synchronize()

    if args.host_profile and rank == 0:
        import cProfile, pstats
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pr = cProfile.Profile()
        pr.enable()
        for i in range(50):
            one_step(i, False)
        pr.disable()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        print(f"[host-profile] 50 steps: host issue {1e3 * (t1 - t0) / 50:.3f} ms/step, drained after "
              f"{1e3 * (t2 - t1):.3f} ms", file=sys.stderr)
        pstats.Stats(pr, stream=sys.stderr).sort_stats("tottime").print_stats(35)

    sampler = ClockSampler(torch.cuda.current_device() if "CUDA_VISIBLE_DEVICES" not in os.environ else local)
    if rank == 0:
        sampler.start()
    launches0 = arm._lib.launch_count() if args.impl == "ours" else 0
    ms_dev = timed(K, False)
    launches = (arm._lib.launch_count() - launches0) if args.impl == "ours" else None
    ms_e2e = timed(K, True)
    eager = None
    if graphed:
        launches = K * B * arm.graphed.kernels_per_replay  # libghr kernels inside the K*B replays
        checker = arm.multi if arm.multi is not None else arm.graphed
        if not checker.check():  # a view overflowed the captured binning capacity: re-captured, measure again
            ms_dev, ms_e2e = timed(K, False), timed(K, True)
            assert checker.check()
        reduced = [g.clone() for g in one_step(0, False, want_grads=True)] if world > 1 else None
        arm.use_graph = False  # the same step issued kernel by kernel, for comparison
        if world > 1:  # the in-graph all-reduce must give what the bucketed NCCL all-reduce of the eager step gives
            ref = one_step(0, False, want_grads=True)
            ar_err = max(float((a - b).abs().max() / b.abs().max().clamp_min(1e-30)) for a, b in zip(reduced, ref))
        for i in range(3):
            one_step(i, False)
        eager = (timed(K, False), timed(K, True))
    clocks = sampler.stop() if rank == 0 else None

    # per-stage device time (ours): a third identical pass with the library's stage events enabled
    stages, roof = None, None
    if args.impl == "ours":
        arm._lib.profile_enable(True)
        arm._lib.profile_read()
        timed(K, False)
        prof = arm._lib.profile_read()
        arm._lib.profile_enable(False)
        # N, V per view (measured, averaged over the views this rank cycles through)
        nv = [arm.measure_NV(views[i % len(views)], vec_dev[i % len(views)]) for i in range(min(K, len(views)))]
        N = float(np.mean([a for a, _ in nv])); V = float(np.mean([b for _, b in nv]))
        from mygauhuman_b200 import _lib
        passes = _lib.state_layout(P, 1, W, H).sort_passes
        peak, peak_src = load_peaks()
        stages = {}
        for name, (ms, cnt) in prof.items():
            if cnt == 0:
                continue
            b = stage_bytes(name, P, V, N, X, tiles, passes)
            us = 1e3 * ms / cnt
            stages[name] = {"us": round(us, 2), "launches_timed": int(cnt),
                            "alg_bytes": None if b is None else int(b),
                            "GBps": None if b is None else round(b / (us * 1e-6) / 1e9, 1)}
        dom = max((s for s in stages if stages[s]["GBps"] is not None), key=lambda s: stages[s]["us"])
        roof = {"kernel": dom, "bound": "hbm", "achieved": stages[dom]["GBps"], "peak": peak, "unit": "GB/s",
                "frac": round(stages[dom]["GBps"] / peak, 4), "traffic": load_traffic(dom), "peak_source": peak_src,
                "note": "blend kernels are fp32-issue/MUFU bound, not HBM bound (DESIGN.md); frac is reported as the contract asks",
                "N_pairs": N, "V_visible": V}

    if rank != 0:
        return
    cpu = None
    if args.impl == "ours" and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(wl, views[0], args.cpu_iters)

    views_per_step = world * B  # B views per rank per step
    n_lanes = arm.multi.K if (args.impl == "ours" and getattr(arm, "multi", None) is not None) else 1
    value = views_per_step * K / (ms_dev * 1e-3)
    e2e_val = views_per_step * K / (ms_e2e * 1e-3)
    h2d = PACK_FLOATS * 4 + 5 * X * 4
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": round(ms_dev / K, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "impl": args.impl,
        "config": {"workload": args.workload, "desc": WORKLOADS[args.workload]["desc"], "P": P, "W": W, "H": H,
                   "sh_degree": 3, "views_cycled": len(views) * world, "parallelism": f"view-sharded dp{world}",
                   "l2": ("flushed before every view (256 MiB fill outside the per-view event pair)" if n_lanes == 1 else
                          f"flushed before every group of {n_lanes} concurrent views (256 MiB fill outside the per-group event pair)"),
                   "lanes": n_lanes,
                   "sync_free": bool(args.sync_free),
                   "mode": (args.mode if args.impl == "ours" else "eager"),
                   "batch_views_per_rank": B,
                   "all_reduce": (None if world == 1 else "one NCCL sum of the flat gradient bucket per step (batch)")},
        "e2e": {"value": round(e2e_val, 2), "unit": UNIT, "ms_per_step": round(ms_e2e / K, 4),
                "h2d_bytes_per_step": h2d * B, "d2h_bytes_per_step": 4 * B,
                "what": "public API step (GraphedRenderStep replay, or in eager mode SMPLDeformer.deform_gaussians -> "
                        "GaussianRasterizer -> loss -> backward); "
                        "camera + SMPL pose + 5 target planes copied from pinned host memory and the loss copied back to "
                        "pinned memory for every view inside the timed region (graph mode: input copies of the next view(s) on "
                        "copy streams during the current one(s), losses consumed one view / group later; with lanes > 1 "
                        "the views of a batch are replayed `lanes` at a time on as many streams (MultiLaneRenderStep); "
                        "eager/reference: blocking loss.item()); "
                        "Gaussian parameters stay resident (model state)"},
        "gpu_launches": launches, "clocks": clocks,
    }
    if eager is not None:
        line["eager"] = {"value": round(views_per_step * K / (eager[0] * 1e-3), 2), "ms_per_step": round(eager[0] / K, 4),
                         "e2e": round(views_per_step * K / (eager[1] * 1e-3), 2),
                         "what": "same step through the eager operator API (one Python call per kernel); host-bound"}
        if ar_err is not None:
            line["eager"]["graph_vs_eager_reduced_grad_rel_err"] = ar_err
    if roof is not None:
        line["roofline"] = roof
        line["stages"] = stages
    if cpu is not None:
        line["cpu_baseline"] = cpu
    if args.impl == "reference":
        line["cpu_baseline"] = {"kind": "reference", "cores": 0, "value": line["value"], "unit": UNIT,
                                "sample": "the reference's own implementation of this path is CUDA: unmodified "
                                          "rasteriser sources built for sm_100 (oracle/_ref), timed on the same GPU; "
                                          "its CPU port is `--impl reference-cpu`"}
    emit(line)


if __name__ == "__main__":
    main()
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
