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
    ap.add_argument("--eager-chain", default="operators", choices=["operators", "fused"],
                    help="--mode eager: the reference-shaped two-operator chain (deform_gaussians -> GaussianRasterizer) or "
                         "the fused per-Gaussian kernels the graph configuration runs (for profiling them one by one)")
    ap.add_argument("--no-optimizer", action="store_true",
                    help="leave the optimiser step out of the measured step (round-1 behaviour: gradient all-reduce only)")
    ap.add_argument("--lr-scale", type=float, default=1e-3,
                    help="the reference's learning rates (arguments/__init__.py) are multiplied by this: the targets are "
                         "random, so full-size Adam steps would walk the synthetic scene away from the named config")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the secondary measurements of the line (u1, stages_large, parity vs the reference arm, config 4)")
    ap.add_argument("--targets", default="u8", choices=["u8", "f32"],
                    help="dtype of the five target planes a view ships from the host (e2e): uint8 (an image dataset's own "
                         "dtype, converted to fp32 inside the loss) or fp32 (round-1 behaviour, 4x the bytes)")
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
    if args.targets == "u8":
        targets = torch.randint(0, 256, (5, H, W), generator=gen, dtype=torch.uint8)
    else:
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


def targets_f32(t):
    return t.to(torch.float32) * (1.0 / 127.5) - 1.0 if t.dtype == torch.uint8 else t


def linear_loss(color, depth, alpha, targets):
    """sum(colour*t[0:3]) + sum(depth*t[3]) + sum(alpha*t[4]): every output plane gets a dense random gradient.
    uint8 targets (what an image dataset holds) are mapped to [-1, 1] here, on the device, inside the step."""
    if targets.dtype == torch.uint8:
        targets = targets.to(torch.float32) * (1.0 / 127.5) - 1.0
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
        # the parameters live in ONE flat buffer with the gradient bucket's layout (leaves = views of it): a parameter is
        # then addressed by the same offset on every rank, which is what the fused sharded optimiser step needs
        from mygauhuman_b200.optim import flatten_leaves
        self.flat_params, self.leaves = flatten_leaves({k: v.to(dev) for k, v in wl["leaves"].items()})
        self.big = {k: v.to(dev) for k, v in wl["big"].items()}
        self.t_vertices = wl["t_vertices"].to(dev)
        self.bg = wl["bg"].to(dev)
        self.sync_free = sync_free
        self.last_n = None
        self.graphed, self.use_graph, self.multi = None, False, None
        self.fused_eager = False
        if graph:
            from mygauhuman_b200.graphed import GraphedRenderStep
            cam = wl["views"][0]["cam"]
            assert all(v["cam"]["tanfovx"] == cam["tanfovx"] and v["cam"]["tanfovy"] == cam["tanfovy"] for v in wl["views"])
            self.multi = None
            if lanes >= 1:  # one lane too: the nearest-vertex search then runs once per batch, not once per view
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
        if self.fused_eager:  # the kernels of the timed (graph) configuration, issued one by one (stage profile)
            from mygauhuman_b200.fused import render_deformed
            rs = dgr.GaussianRasterizationSettings(
                image_height=self.wl["H"], image_width=self.wl["W"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                bg=self.bg, scale_modifier=1.0, viewmatrix=camd["view"], projmatrix=camd["proj"], sh_degree=3,
                campos=camd["campos"], prefiltered=False, debug=False)
            if self.sync_free:
                dgr._C.set_sync_free_capacity(None if self.last_n is None else int(self.last_n * 1.3) + 4096)
            color, radii, depth, alpha = render_deformed(
                self.deformer, L["xyz"], pose, self.big, self.t_vertices, rs, log_scaling=L["log_scale"],
                rotation=L["rotation"], features=L["features"], opacity_logit=L["opacity_logit"])
            loss = linear_loss(color, depth, alpha, targets_dev)
            for t in L.values():
                t.grad = None
            loss.backward()
            return loss, [L[k].grad for k in ("xyz", "features", "opacity_logit", "log_scale", "rotation")]
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
    """The reference's path on the same GPU, none of libghr's kernels on it:
    * LBS / covariance / SH: the reference's OWN modules -- GaussianModel.coarse_deform_c2source (scene/gaussian_model.py
      :768-872), GaussianModel.get_covariance (:208-209) and utils/sh_utils.eval_sh (:135-190) exactly as render()
      (gaussian_renderer/__init__.py:94-96,124-125,181,193-195) calls them -- imported unmodified from baseline/_ref
      (oracle/install_ref_py.sh + oracle/ref_py.py; knn_cuda, absent from the image, is an exact torch.cdist + topk).
      Only when baseline/_ref is missing does the arm fall back to the restatement oracle/lbs_oracle.py, and says so.
    * rasteriser: the UNMODIFIED reference CUDA sources built for sm_100 (oracle/_ref)."""
    name = "reference"

    def __init__(self, wl, dev):
        from oracle import lbs_oracle, ref_cuda, ref_py
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
        self.model, self.eval_sh = None, None
        if ref_py.available():
            ns = ref_py.load()
            self.model = ref_py.make_model(self.smpl, self.leaves, dev)
            self.eval_sh = ns.sh_utils.eval_sh
        self.lbs_source = ("reference modules from baseline/_ref: scene/gaussian_model.py coarse_deform_c2source + "
                           "get_covariance, utils/sh_utils.py eval_sh" if self.model is not None else
                           "RESTATEMENT oracle/lbs_oracle.py (baseline/_ref not installed)")

    def deform(self, pose, campos):
        """-> (world means, covariance, colours, opacity) the way render() computes them."""
        L = self.leaves
        if self.model is not None:
            pc = self.model
            _, means3D, _, transforms, _, _ = pc.coarse_deform_c2source(
                pc.get_xyz[None], pose, self.big, self.t_vertices[None], normals=pc.get_normal[None])
            means3D = means3D.squeeze()
            cov = pc.get_covariance(1.0, transforms.squeeze())
            dir_pp = means3D - campos.repeat(pc.get_features.shape[0], 1)
            dir_pp_normalized = dir_pp / dir_pp.norm(dim=1, keepdim=True)
            shs_view = pc.get_features.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
            col = torch.clamp_min(self.eval_sh(pc.active_sh_degree, shs_view, dir_pp_normalized) + 0.5, 0.0)
            return means3D, cov, col, pc.get_opacity
        LO = self.LO
        with torch.no_grad():
            from oracle.ref_py import BruteForceKNN  # exact nearest vertex (GEMM candidates, exact re-rank)
            vid = BruteForceKNN(1)(self.t_vertices[None], L["xyz"][None])[1].reshape(-1)
        _, world, _, T, _, wn, _ = LO.coarse_deform_c2source(self.smpl, L["xyz"], pose, self.big, self.t_vertices,
                                                             normals=L["normals"] / L["normals"].norm(dim=1, keepdim=True),
                                                             vert_ids=vid)
        return (world, LO.covariance(torch.exp(L["log_scale"]), 1.0, L["rotation"], T),
                LO.view_colors(3, L["features"], world, campos), torch.sigmoid(L["opacity_logit"]))

    def step(self, view, vec_dev, targets_dev):
        L = self.leaves
        camd, pose = unpack(vec_dev)
        cam = view["cam"]
        world, cov, col, opac = self.deform(pose, camd["campos"])
        means2D = torch.zeros_like(world, requires_grad=True)
        color, radii, depth, alpha = _RefRasterize.apply(
            self.ref, world.contiguous(), means2D, col.contiguous(), opac.contiguous(), cov.contiguous(), self.bg,
            camd["view"].contiguous(), camd["proj"].contiguous(), camd["campos"].contiguous(), cam["tanfovx"],
            cam["tanfovy"], self.wl["H"], self.wl["W"])
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
    tg = a(targets_f32(wl["targets"]))
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
        # fused chain (ghr_fused.cu): read xyz 12 + vid 4 + weight row 96 + scale 12 + quat 16 + logit 4 + SH 192; write
        # radii / tiles_touched / point_offsets 12, and per visible Gaussian the 48-byte record + clamp flags
        "deform_project": 348 * P + 49 * V,
        # read the same inputs + radii 4, per visible Gaussian the 48-byte accumulator row; write dL/dmean2D 12 and add
        # (atomics) xyz 12 + scale 12 + quat 16 + logit 4 + SH 192 per visible Gaussian
        "project_deform_bwd": 352 * P + 48 * V + 236 * V,
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
# optimiser of the measured step
# ---------------------------------------------------------------------------

# arguments/__init__.py OptimizationParams of the reference: position_lr_init * spatial_lr_scale(=1), feature_lr,
# opacity_lr, scaling_lr, rotation_lr; `normals` follows the feature rate
REFERENCE_LRS = {"xyz": 1.6e-4, "log_scale": 5e-3, "rotation": 1e-3, "features": 2.5e-3, "opacity_logit": 5e-2,
                 "normals": 2.5e-3}


def make_optimizer(args, arm, world, bucket):
    """The optimiser step that ends a batch (reference: torch.optim.Adam(l, lr=0.0, eps=1e-15), gaussian_model.py:283,
    stepped at train.py:418).  ours, 1 GPU: FusedAdam (one launch).  ours, N GPUs: ShardedFusedAdam(sync="flags") --
    gradient reduce-scatter + Adam + parameter all-gather in ONE kernel over NVLink peer memory, its two cross-rank
    barriers inside the kernel; it REPLACES the NCCL all-reduce of the bucket.  reference arm: NCCL all-reduce + the
    reference's own optimiser (torch.optim.Adam, foreach)."""
    lrs = {k: REFERENCE_LRS[k] * args.lr_scale for k in arm.leaves}
    if args.impl != "ours":
        return torch.optim.Adam([{"params": [t], "lr": lrs[k], "name": k} for k, t in arm.leaves.items()], lr=0.0, eps=1e-15), "torch.optim.Adam (foreach), the reference's"
    from mygauhuman_b200.optim import FusedAdam, ShardedFusedAdam, flat_layout
    if world > 1 and bucket is not None:
        total, table = flat_layout(arm.leaves)
        assert total == arm.flat_params.numel() == bucket.numel()
        groups = [dict(name=k, offset=table[k][0], count=table[k][1], lr=lrs[k]) for k in arm.leaves]
        return ShardedFusedAdam(arm.flat_params, bucket, groups, eps=1e-15, sync="flags"), \
            "ShardedFusedAdam(sync=flags): reduce-scatter + Adam + all-gather in one kernel over NVLink peer memory (ghr_adam_step_sharded_synced), no NCCL call"
    return FusedAdam([{"params": [t], "lr": lrs[k], "name": k} for k, t in arm.leaves.items()], lr=0.0, eps=1e-15), \
        "FusedAdam (ghr_adam_step, one launch)" + (" after one NCCL all-reduce of the bucket" if world > 1 else "")


def verify_sharded_step(arm, opt, world, rank, run_batch):
    """The fused compute+collective kernel against what it replaces, inside the run: one batch of views, then
    ShardedFusedAdam.step() on the live buffers vs NCCL all-reduce of a copy of the bucket + FusedAdam on a copy of the
    parameters (fresh moments on both sides: this is the optimiser's first step).  Bit-exact at 2 ranks (a + b commutes);
    at 4 / 8 see the rule below.  Returns the report."""
    import torch.distributed as dist
    from mygauhuman_b200.optim import FusedAdam, flat_layout
    assert opt.steps == 0
    run_batch()  # fills the bucket with this rank's gradients
    torch.cuda.synchronize()
    total, table = flat_layout(arm.leaves)
    summed = opt.grads.clone()
    dist.all_reduce(summed)
    ref = {k: torch.nn.Parameter(t.detach().clone()) for k, t in arm.leaves.items()}
    for k, t in ref.items():
        o, n = table[k]
        t.grad = summed[o:o + n].view_as(t).clone()
    lr_of = {g["name"]: g["lr"] for g in opt.param_groups}
    FusedAdam([{"params": [ref[k]], "lr": lr_of[k]} for k in ref], lr=0.0, eps=1e-15).step()
    opt.step()
    torch.cuda.synchronize()
    opt.check()
    # More than 2 ranks: NCCL's summation order differs from the kernel's fixed rank order, so the summed gradients
    # differ in the last bits.  This is the optimiser's FIRST step, whose update is lr * g / (|g| + eps) = +-lr whatever
    # the magnitude of g: an element whose gradient cancels to ~0 over the ranks can land on either sign.  The bound is
    # therefore per element |dp| <= 2 lr of its group, and such elements must be rare; everything else must agree to
    # 1e-6 of the step.
    worst, exact, frac, steps_off = 0.0, True, 0.0, 0.0
    for k, t in arm.leaves.items():
        a, b = t.detach(), ref[k].detach()
        exact = exact and bool(torch.equal(a, b))
        worst = max(worst, float((a - b).abs().max() / b.abs().max().clamp_min(1e-30)))
        d = (a - b).abs() / max(lr_of[k], 1e-30)          # difference in units of this group's Adam step
        steps_off = max(steps_off, float(d.max()))
        frac = max(frac, float((d > 1e-6).float().mean()))
    w = torch.tensor([worst, 0.0 if exact else 1.0, frac, steps_off], device=summed.device, dtype=torch.float64)
    dist.all_reduce(w, op=dist.ReduceOp.MAX)
    worst, exact, frac, steps_off = float(w[0]), float(w[1]) == 0.0, float(w[2]), float(w[3])
    if world == 2:
        assert exact, f"sharded step differs from all-reduce + FusedAdam at 2 ranks (worst rel {worst:.3e})"
    assert steps_off <= 2.0 * (1 + 1e-5) and frac <= 1e-3, \
        f"sharded step differs from all-reduce + FusedAdam: {steps_off:.3e} Adam steps on {frac:.3e} of a tensor"
    # replicas stay bit-identical
    mine = arm.flat_params.detach().clone()
    other = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(other, mine)
    assert all(torch.equal(o, other[0]) for o in other), "parameter replicas differ across ranks"
    return {"vs": "NCCL all-reduce + FusedAdam on copies, first optimiser step", "worst_rel": worst, "bit_exact": exact,
            "worst_in_adam_steps": steps_off, "fraction_of_elements_over_1e-6_step": frac, "replicas_bit_identical": True}


# ---------------------------------------------------------------------------
# secondary measurements carried by the line (ours arm, rank 0, N = 1)
# ---------------------------------------------------------------------------

def _event_times(fn, flush, iters, warm=3, sync_each=True):
    """Device ms of fn() per call: CUDA events on the current stream, L2 flushed before each call (outside the pair)."""
    for _ in range(warm):
        fn()
    ts = []
    for i in range(iters):
        flush.fill_(i & 0xFF)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def measure_u1(arm, view, vec_dev, targets_dev, flush, iters=40):
    """U1 (north_star's ">= 2x the reference diff-gaussian-rasterization fwd+bwd iterations/s"): ONE
    rasteriser call, forward + backward, on the deformed Gaussians of view 0 (precomputed colour + covariance, as
    render() calls it), through the raw _C entry points.  ours: replayed as a CUDA graph (device time; the eager figure
    beside it is host-bound) and eagerly; reference: oracle/_ref, the unmodified CUDA sources, eagerly with its blocking
    num_rendered read, which is its real cost.  L2 flushed before every call."""
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    camd, pose = unpack(vec_dev)
    cam, L, e = view["cam"], arm.leaves, torch.Tensor([])
    H, W = arm.wl["H"], arm.wl["W"]
    with torch.no_grad():
        d = arm.deformer.deform_gaussians(L["xyz"], pose, arm.big, arm.t_vertices, scaling=torch.exp(L["log_scale"]),
                                          rotation=L["rotation"], features=L["features"], sh_degree=3, campos=camd["campos"])
        means, col, cov = d["means3D"].contiguous(), d["colors"].contiguous(), d["cov3D"].contiguous()
        opa = torch.sigmoid(L["opacity_logit"]).contiguous()
    view_m, proj_m, campos = camd["view"].contiguous(), camd["proj"].contiguous(), camd["campos"].contiguous()
    tg = targets_f32(targets_dev)
    gc, gd, ga = tg[0:3].contiguous(), tg[3:4].contiguous(), tg[4:5].contiguous()
    fargs = (arm.bg, means, col, opa, e, e, 1.0, cov, view_m, proj_m, cam["tanfovx"], cam["tanfovy"], H, W, e, 3, campos,
             False, False)

    def ours_fwd():
        return _C.rasterize_gaussians(*fargs)

    def ours_bwd(fw):
        n, color, depth, alpha, radii, geom, binning, img = fw
        return _C.rasterize_gaussians_backward(arm.bg, means, radii, col, e, e, 1.0, cov, view_m, proj_m, cam["tanfovx"],
                                               cam["tanfovy"], gc, gd, ga, e, 3, campos, geom, n, binning, img, alpha, False)

    fw = ours_fwd()
    n = int(fw[0])
    out = {"N_pairs": n, "V_visible": int((fw[4] > 0).sum())}
    out["ours_eager_fwd_us"] = round(1e3 * _event_times(ours_fwd, flush, iters), 1)
    out["ours_eager_fwd_bwd_us"] = round(1e3 * _event_times(lambda: ours_bwd(ours_fwd()), flush, iters), 1)
    # graph replay: the same two calls captured once (sync-free binning capacity), what GraphedRenderStep does per view
    cap = _C._round_cap(int(n * 1.3) + 4096)
    graphs = {}
    try:
        _C.set_sync_free_capacity(cap)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                ours_bwd(ours_fwd())
        torch.cuda.current_stream().wait_stream(side)
        for name, body in (("fwd", ours_fwd), ("fwd_bwd", lambda: ours_bwd(ours_fwd()))):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                keep = body()
            graphs[name] = (g, keep)
    finally:
        _C.set_sync_free_capacity(None)
    out["ours_fwd_us"] = round(1e3 * _event_times(graphs["fwd"][0].replay, flush, iters), 1)
    out["ours_fwd_bwd_us"] = round(1e3 * _event_times(graphs["fwd_bwd"][0].replay, flush, iters), 1)
    out["ours_fwd_bwd_iters_per_s"] = round(1e6 / out["ours_fwd_bwd_us"], 1)
    try:
        from oracle import ref_cuda
        have_ref = ref_cuda.available()
    except Exception:
        have_ref = False
    if have_ref:
        rr = ref_cuda.RefRasterizer()

        def ref_fwd():
            return rr.forward(arm.bg, means, col, opa, None, None, 1.0, cov, view_m, proj_m, cam["tanfovx"], cam["tanfovy"],
                              H, W, None, 3, campos)

        def ref_fwd_bwd():
            rn, rcolor, rdepth, ralpha, rradii = ref_fwd()
            return rr.backward(arm.bg, means, rradii, col, None, None, 1.0, cov, view_m, proj_m, cam["tanfovx"],
                               cam["tanfovy"], gc, gd, ga, None, 3, campos, ralpha)

        out["reference_fwd_us"] = round(1e3 * _event_times(ref_fwd, flush, iters), 1)
        out["reference_fwd_bwd_us"] = round(1e3 * _event_times(ref_fwd_bwd, flush, iters), 1)
        out["reference_fwd_bwd_iters_per_s"] = round(1e6 / out["reference_fwd_bwd_us"], 1)
        out["speedup_fwd_bwd"] = round(out["reference_fwd_bwd_us"] / out["ours_fwd_bwd_us"], 2)
        out["speedup_fwd"] = round(out["reference_fwd_us"] / out["ours_fwd_us"], 2)
        out["reference"] = "oracle/_ref: unmodified diff-gaussian-rasterization sources built for sm_100, same inputs, same run"
    else:
        out["reference"] = "oracle/_ref not built"
    out["what"] = ("one GaussianRasterizer-equivalent call (raw _C.rasterize_gaussians + _backward), precomputed colour + "
                   "covariance, view 0; ours_* = CUDA-graph replay, ours_eager_* = same calls issued from Python; "
                   "median of %d calls, L2 flushed before each" % iters)
    return out


def grad_agreement(got, want, rel=1e-4):
    """Per-element agreement of two gradient sets (dict name -> tensor): an element passes when
    |a - b| <= rel * |b| + 0.1 * rel * max|b| (north_star: parameter gradients within 1e-4 rel; the second term is the
    floor below which fp32 accumulation order decides).  -> (all pass, report dict)."""
    rep, ok = {}, True
    for k in want:
        a, b = got[k].detach().double().reshape(-1), want[k].detach().double().reshape(-1)
        scale = float(b.abs().max().clamp_min(1e-30))
        err = (a - b).abs()
        bad = int((err > rel * b.abs() + 0.1 * rel * scale).sum())
        rep[k] = {"max_abs_err_over_scale": float(err.max() / scale), "elements_outside": bad, "elements": a.numel(),
                  "rel_l2": float((a - b).norm() / b.norm().clamp_min(1e-30))}
        ok = ok and bad == 0
    return ok, rep


def batch_parity(arm, views, vec_dev, targets_dev, B, ref_arm):
    """The batch gradient the TIMED configuration produces (graph replay, `lanes` concurrent views adding into one bucket
    with fp32 atomics) against (1) the same B views through the eager operator API one at a time, summed in fp64, and
    (2) the reference arm (its own LBS modules + its CUDA rasteriser) on the same views, summed in fp64."""
    names = ("xyz", "features", "opacity_logit", "log_scale", "rotation")
    ml = arm.multi
    nv = len(views)
    if ml is not None:
        ml.zero_grads()
        for g0 in range(0, B, ml.K):
            ml.run_group([vec_dev[(g0 + j) % nv] for j in range(min(ml.K, B - g0))])
        got = {k: ml.flat_views[k].clone() for k in names}
    else:
        arm.graphed.zero_grads()
        for i in range(B):
            arm.graphed(vec_dev[i % nv], arm.graphed.loss_inputs[0])
        got = {k: arm.graphed.flat_views[k].clone() for k in names}
    torch.cuda.synchronize()
    out = {"views": B, "criterion": "|a-b| <= 1e-4 |b| + 1e-5 max|b| per element"}
    was = arm.use_graph
    arm.use_graph = False
    want = None
    for i in range(B):
        _, grads = arm.step(views[i % nv], vec_dev[i % nv], targets_dev)
        gs = {k: g.detach().double() for k, g in zip(names, grads)}
        want = gs if want is None else {k: want[k] + gs[k] for k in want}
    arm.use_graph = was
    if ml is not None:
        ml._alias_grads()
    else:
        for k, t in arm.leaves.items():
            t.grad = arm.graphed.flat_views[k]
    ok, rep = grad_agreement(got, want)
    out["graph_lanes_vs_eager_sum"] = {"ok": ok, "per_tensor": rep}
    assert ok, f"batch gradient of the timed configuration differs from the eager per-view sum: {rep}"
    if ref_arm is not None:
        for k in names:  # same parameter values on both arms
            ref_arm.leaves[k].data.copy_(arm.leaves[k].detach())
        ref_arm.leaves["normals"].data.copy_(arm.leaves["normals"].detach())
        rwant = None
        for i in range(B):
            _, grads = ref_arm.step(views[i % nv], vec_dev[i % nv], targets_dev)
            gs = {k: g.detach().double() for k, g in zip(names, grads)}
            rwant = gs if rwant is None else {k: rwant[k] + gs[k] for k in rwant}
        _, rep_r = grad_agreement(got, rwant)
        # Two complete fp32 pipelines: the rasteriser inputs of the two arms differ in the last bits (4e-7 relative, other
        # LBS code), so a handful of discrete decisions inside the rasteriser (alpha < 1/255, T < 1e-4, ceil of the
        # radius) fall differently and the gradients of the Gaussians involved differ visibly -- one flipped (Gaussian,
        # pixel) pair moves that Gaussian's gradient by a whole pixel's contribution, up to a few % of the tensor's largest
        # entry.  Hence an aggregate rule: at least 99 % of the elements inside the per-element criterion and the whole
        # tensor within 1e-2 in relative L2; the largest outlier is reported, not gated.
        # (Same inputs into both rasterisers: tests/test_parity_at_scale_gpu.py, every element inside.)
        ok_r = all(v["elements_outside"] <= 0.01 * v["elements"] and v["rel_l2"] <= 1e-2 for v in rep_r.values())
        out["graph_lanes_vs_reference_arm"] = {"ok": ok_r, "rule": ">= 99 % of elements inside the criterion and relative L2 "
                                               "error of each tensor <= 1e-2 (discrete rasteriser decisions flip on "
                                               "last-bit input differences)", "per_tensor": rep_r, "lbs": ref_arm.lbs_source}
    return out


def measure_large(dev, flush, peak):
    """The streaming kernels where HBM IS the bound (BASELINE.json config 5): the 3M-Gaussian 1920x1080 forward
    (SH + scale/rotation path, as baking.py calls the rasteriser) stage by stage, and the onesweep sort of 2^24
    45-bit depth|tile keys.  GB/s = algorithmic bytes (SURVEY 8d formulas, with the radix passes actually run) / time."""
    from mygauhuman_b200 import _lib
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    out = {}
    # One frame owns the GPU here: leave the shared-GPU blend configuration the multi-lane step switched on (half the CTAs
    # per SM, so that two views' blend kernels fit an SM together) for the duration of this measurement.
    was_shared = _lib.lib().ghr_shared_gpu()
    _lib.lib().ghr_set_shared_gpu(0)
    try:
        return _measure_large(dev, flush, peak, _lib, _C, out)
    finally:
        _lib.lib().ghr_set_shared_gpu(was_shared)


def _measure_large(dev, flush, peak, _lib, _C, out):
    P, W, H = 3_000_000, 1920, 1080
    g = torch.Generator(device=dev).manual_seed(7)
    xyz = torch.rand(P, 3, device=dev, generator=g) * 10.0 - 5.0
    xyz[:, 2] += 6.0
    scal = torch.exp(math.log(0.02) + 0.3 * torch.randn(P, 3, device=dev, generator=g))
    rot = torch.nn.functional.normalize(torch.randn(P, 4, device=dev, generator=g))
    opac = torch.sigmoid(1.5 * torch.randn(P, 1, device=dev, generator=g))
    sh = torch.cat([torch.rand(P, 1, 3, device=dev, generator=g) * 2 - 1, 0.05 * torch.randn(P, 15, 3, device=dev, generator=g)], 1).contiguous()
    from mygauhuman_b200 import synthetic as S
    cam = S.make_camera(W, H, angle=math.pi, radius=6.0, target=(0.0, 0.0, 6.0), elev=0.0, seed=0, focal=1100.0)
    e = torch.Tensor([])
    args = (torch.zeros(3, device=dev), xyz, e, opac, scal, rot, 1.0, e, cam["world_view_transform"].to(dev),
            cam["full_proj_transform"].to(dev), cam["tanfovx"], cam["tanfovy"], H, W, sh, 3, cam["camera_center"].to(dev),
            False, False)
    res = _C.rasterize_gaussians(*args)
    N, V = int(res[0]), int((res[4] > 0).sum())
    ms = _event_times(lambda: _C.rasterize_gaussians(*args), flush, 10)
    passes = _lib.state_layout(P, 64, W, H).sort_passes
    tiles, X = ((W + 15) // 16) * ((H + 15) // 16), W * H
    _lib.profile_enable(True)
    _lib.profile_read()
    for i in range(5):
        flush.fill_(i)
        _C.rasterize_gaussians(*args)
    torch.cuda.synchronize()
    prof = _lib.profile_read()
    _lib.profile_enable(False)
    alg = {"preprocess": 244 * P + 67 * V, "emit_keys": 8 * P + 12 * V + 12 * N, "sort": N * (8 + 24 * passes),
           "tile_ranges": 8 * N + 8 * tiles, "blend_fwd": 44 * N + 28 * X}
    stages = {}
    for name, (t, c) in prof.items():
        if c and name in alg:
            us = 1e3 * t / c
            stages[name] = {"us": round(us, 1), "alg_bytes": int(alg[name]), "GBps": round(alg[name] / us / 1e3, 1),
                            "frac_of_hbm_peak": round(alg[name] / us / 1e3 / peak, 3)}
    out["stress_1080p_forward"] = {"P": P, "W": W, "H": H, "N_pairs": N, "V_visible": V, "ms": round(ms, 3),
                                   "fps": round(1e3 / ms, 1), "radix_passes": passes, "stages": stages}
    del res, xyz, scal, rot, opac, sh
    # sort alone: 2^24 keys of 32 depth bits + 13 tile bits (8 160 tiles at 1080p)
    l = _lib.lib()
    Nk, bits = 1 << 24, 45
    tile = torch.randint(0, 8160, (Nk,), device=dev, generator=g, dtype=torch.int64)
    depth = (torch.rand(Nk, device=dev, generator=g) * 9.8 + 0.2).view(torch.int32).to(torch.int64)
    keys = (tile << 32) | depth
    vals = torch.arange(Nk, dtype=torch.int32, device=dev)
    ka, kb, va, vb = keys.clone(), torch.empty_like(keys), vals.clone(), torch.empty_like(vals)
    ws = torch.empty(l.ghr_sort_workspace_bytes(Nk), dtype=torch.uint8, device=dev)
    sel = torch.zeros(1, dtype=torch.int32, device=dev)
    stream = _lib.raw_stream(dev.index)

    def run():
        _lib.check(l.ghr_sort_pairs_u64(Nk, bits, ka.data_ptr(), kb.data_ptr(), va.data_ptr(), vb.data_ptr(), ws.data_ptr(),
                                        ws.numel(), sel.data_ptr(), stream), "ghr_sort_pairs_u64")
    ts = []
    for i in range(8):
        ka.copy_(keys); va.copy_(vals)
        flush.fill_(i)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ms = float(np.median(ts[2:]))
    sorted_keys = kb if int(sel.item()) else ka
    assert bool((sorted_keys[1:] >= sorted_keys[:-1]).all()), "sort_2p24: output not sorted"
    npass = (bits + 8) // 9
    bytes_alg = Nk * (8 + 24 * npass)
    out["sort_2p24"] = {"keys": Nk, "bits": bits, "radix_passes": npass, "ms": round(ms, 3),
                        "Gkeys_per_s": round(Nk / ms / 1e6, 2), "alg_bytes": bytes_alg,
                        "GBps": round(bytes_alg / ms / 1e6, 1), "frac_of_hbm_peak": round(bytes_alg / ms / 1e6 / peak, 3)}
    return out


def load_issue_counts():
    """Warp-level instructions executed per launch (ncu smsp__inst_executed.sum) of the blend kernels at the zju_512
    shape, from the committed capture named in the file, or {}."""
    try:
        with open(os.path.join(ROOT, "profiles", "issue.json")) as f:
            return json.load(f)
    except Exception:
        return {}


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


def run_workload(args, workload, world, rank, local, dev, K, Wm, headline):
    """Measure one workload; returns the line (rank 0) or None.  headline=False: a secondary workload carried under an
    extra key (config 4 in multi-GPU runs): device-timed value + e2e only."""
    if world > 1:
        import torch.distributed as dist
    wargs = argparse.Namespace(**vars(args))
    wargs.workload = workload
    wl = build_workload(wargs, rank, world)
    P, W, H = wl["P"], wl["W"], wl["H"]
    X, tiles = W * H, ((W + 15) // 16) * ((H + 15) // 16)
    ours = args.impl == "ours"
    if ours:
        lanes = math.gcd(max(1, args.lanes), max(1, args.batch_views)) if args.mode == "graph" else 1
        arm = OursArm(wl, dev, args.sync_free, graph=(args.mode == "graph"), lanes=lanes)
        arm.fused_eager = args.mode == "eager" and args.eager_chain == "fused"
    else:
        arm = ReferenceArm(wl, dev)
    views = wl["views"]
    vec_host = [view_pack(v) for v in views]
    vec_dev = [v.to(dev) for v in vec_host]
    targets_host = wl["targets"].pin_memory()
    targets_dev = targets_host.to(dev)
    vec_stage = torch.empty(PACK_FLOATS, device=dev)
    targets_stage = torch.empty_like(targets_dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    from mygauhuman_b200.multiview import GradientAllReducer
    reducer = GradientAllReducer()
    B = max(1, args.batch_views)
    opt, opt_desc, sharded = None, None, False

    def graph_on():
        return ours and arm.use_graph

    def begin_batch():
        if graph_on():
            (arm.multi or arm.graphed).zero_grads()  # clears the bucket (and, with lanes, refreshes the nearest-vertex ids)

    def end_batch(grads):
        """End of a multi-view batch: the gradients of the batch are summed over the ranks and the optimiser steps.
        With ShardedFusedAdam the sum IS the optimiser kernel; otherwise one NCCL all-reduce precedes it."""
        if graph_on():
            if not sharded:
                arm.graphed.reduce_grads()
        elif world > 1:
            reducer(grads)
        if opt is not None:
            opt.step()

    def one_view(i, e2e):
        k = i % len(views)
        if graph_on():
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
        """A batch of one view (warm-up)."""
        begin_batch()
        loss, grads = one_view(i, e2e)
        end_batch(grads)
        if e2e:
            return float(loss.item())
        return grads if want_grads else None

    flush_ms = [0.0]

    def finish(starts, ends, t0, n_flushes):
        """Sum of the event pairs (max over ranks) and the wall clock of the whole loop minus the L2 flushes."""
        torch.cuda.synchronize(); barrier(world); torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3 - n_flushes * flush_ms[0]
        ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
        if world > 1:
            t = torch.tensor([ms, wall], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, wall = float(t[0]), float(t[1])
        return ms, wall

    def timed_lanes(K, e2e):
        """K steps of B views, the views of a batch replayed `lanes` at a time on as many streams (MultiLaneRenderStep).
        One event pair per GROUP of concurrent views on the stream the lanes fork from and join into (L2 flushed before
        every group, outside the pair); the first group's pair of a batch includes the bucket reset, the last one's the
        gradient reduction + optimiser step.  e2e: every view's inputs come from pinned host memory (the next group's
        copies are issued right after this group's replays, on the lanes' copy streams) and every view's loss goes back
        to pinned memory, consumed one group later -- all inside the timed region."""
        ml = arm.multi
        Kl = ml.K
        groups = K * (B // Kl)
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(groups)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(groups)]
        nv = len(views)
        pending, host_losses = [], 0.0
        torch.cuda.synchronize(); barrier(world); torch.cuda.synchronize()
        t0 = time.perf_counter()
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
                end_batch(None)
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
        return finish(starts, ends, t0, groups)

    def timed(K, e2e):
        """K steps of B views.  Every view has its own event pair (L2 flushed before it, outside the pair); the first
        view's pair includes the bucket reset, the last one's the gradient reduction + optimiser step.  e2e: inputs
        from pinned host memory and the view's loss read back.  With the graphed step the e2e loop is pipelined the way
        a training loop would be: the inputs of view i+1 travel on a copy stream while view i computes, and the loss of
        view i is copied to pinned memory asynchronously and consumed one view later -- every copy still happens inside
        the timed region.  Returns (sum of event pairs, wall clock of the loop minus the flushes), max over ranks."""
        if graph_on() and arm.multi is not None:
            return timed_lanes(K, e2e)
        n = K * B
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
        piped = e2e and graph_on()
        pending, host_losses = None, 0.0
        torch.cuda.synchronize(); barrier(world); torch.cuda.synchronize()
        t0 = time.perf_counter()
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
        return finish(starts, ends, t0, n)

    # the L2 flush alone (subtracted from the wall-clock cross-check)
    fl = []
    for i in range(6):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); flush.fill_(i); b.record()
        torch.cuda.synchronize()
        fl.append(a.elapsed_time(b))
    flush_ms[0] = float(np.median(fl[2:]))

    for i in range(Wm):
        one_step(i, False)
        one_step(i, True)
    torch.cuda.synchronize()
    graphed = ours and args.mode == "graph"
    if graphed:
        # capacity sized from every view this rank cycles through, then one capture; warm the replay path too
        arm.capture(vec_dev, targets_dev)
        arm.use_graph = True
    bucket = (arm.multi or arm.graphed).flat if graphed else None
    sharded_check = None
    if not args.no_optimizer:
        opt, opt_desc = make_optimizer(args, arm, world, bucket)
        sharded = ours and world > 1 and bucket is not None
        if sharded:
            def first_batch():
                begin_batch()
                if arm.multi is not None:
                    for g0 in range(0, B, arm.multi.K):
                        arm.multi.run_group([vec_dev[(g0 + j) % len(views)] for j in range(arm.multi.K)])
                else:
                    for i in range(B):
                        one_view(i, False)
            sharded_check = verify_sharded_step(arm, opt, world, rank, first_batch)
    for i in range(Wm):
        one_step(i, False)
        one_step(i, True)
    torch.cuda.synchronize()

    if args.host_profile and rank == 0 and headline:
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
    launches0 = arm._lib.launch_count() if ours else 0
    ms_dev, wall_dev = timed(K, False)
    launches = (arm._lib.launch_count() - launches0) if ours else None
    ms_e2e, wall_e2e = timed(K, True)
    eager = None
    if graphed:
        per_batch = (1 if opt is not None else 0)  # optimiser kernel (graph replays count their own kernels)
        launches = K * B * arm.graphed.kernels_per_replay + K * per_batch
        if arm.multi is not None:
            launches += K * 1  # nearest-vertex search once per batch (zero_grads)
            if arm.multi.batch_pose:
                launches += K * (B // arm.multi.K)  # the batched pose stage, once per group of concurrent views
        checker = arm.multi if arm.multi is not None else arm.graphed
        if not checker.check():  # a view overflowed the captured binning capacity: re-captured, measure again
            (ms_dev, wall_dev), (ms_e2e, wall_e2e) = timed(K, False), timed(K, True)
            assert checker.check()
        if sharded:
            opt.check()
    clocks = sampler.stop() if rank == 0 else None
    if graphed and headline:
        arm.use_graph = False  # the same step issued kernel by kernel, for comparison
        was_opt, was_sharded = opt, sharded
        if sharded:
            opt, sharded = None, False  # the sharded step needs the flat bucket the graph path fills
        for i in range(3):
            one_step(i, False)
        eager = (timed(K, False)[0], timed(K, True)[0])
        opt, sharded = was_opt, was_sharded
        arm.use_graph = True
        for k, t in arm.leaves.items():
            t.grad = (arm.multi or arm.graphed).flat_views[k]

    # per-stage device time (ours): a further eager pass with the library's stage events enabled
    stages, roof, NV = None, None, None
    if ours and headline:
        arm.use_graph = False
        keep_opt, opt = opt, None
        keep_chain = arm.fused_eager
        arm.fused_eager = graphed or keep_chain  # the graph configuration runs the fused per-Gaussian kernels: profile those
        arm._lib.profile_enable(True)
        arm._lib.profile_read()
        timed(K, False)
        prof = arm._lib.profile_read()
        arm._lib.profile_enable(False)
        arm.fused_eager = keep_chain
        opt = keep_opt
        arm.use_graph = graphed
        if graphed:
            for k, t in arm.leaves.items():
                t.grad = (arm.multi or arm.graphed).flat_views[k]
        # N, V per view (measured, averaged over the views this rank cycles through)
        nv = [arm.measure_NV(views[i % len(views)], vec_dev[i % len(views)]) for i in range(min(K, len(views)))]
        N = float(np.mean([a for a, _ in nv])); V = float(np.mean([b for _, b in nv]))
        NV = (N, V)
        from mygauhuman_b200 import _lib
        passes = _lib.state_layout(P, 64, W, H).sort_passes
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
        # The dominant kernel is bound by INSTRUCTION ISSUE along the heaviest tiles' lists, not by HBM (its DRAM traffic
        # is below the algorithmic bytes).  achieved = warp instructions per second: ncu's smsp__inst_executed.sum of
        # the committed capture (profiles/issue.json) / the kernel's duration measured live here; peak = SMs x 4
        # schedulers x SM clock (one warp instruction per scheduler per cycle).  The HBM figure stays beside it.
        issue = load_issue_counts().get(workload, {}).get(dom)
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        mhz = (clocks or {}).get("sm_mhz") or (clocks or {}).get("sm_max_mhz") or 1965.0
        peak_issue = sms * 4 * mhz * 1e6 / 1e9  # G warp-instructions / s
        hbm = {"achieved": stages[dom]["GBps"], "peak": peak, "unit": "GB/s", "frac": round(stages[dom]["GBps"] / peak, 4),
               "alg_bytes": stages[dom]["alg_bytes"], "traffic": load_traffic(dom), "peak_source": peak_src}
        if issue:
            ach = issue["inst_executed"] / (stages[dom]["us"] * 1e-6) / 1e9
            roof = {"kernel": dom, "bound": "issue", "achieved": round(ach, 1), "peak": round(peak_issue, 1),
                    "unit": "Gwarp-inst/s", "frac": round(ach / peak_issue, 4), "traffic": load_traffic(dom),
                    "inst_executed_per_launch": issue["inst_executed"], "inst_source": issue.get("source"),
                    "peak_source": f"{sms} SMs x 4 schedulers x {mhz:.0f} MHz (clock sampled during this run)",
                    "us": stages[dom]["us"], "hbm": hbm, "N_pairs": N, "V_visible": V}
        else:
            roof = dict(hbm, kernel=dom, bound="hbm", N_pairs=N, V_visible=V,
                        note="no committed instruction count for this kernel (profiles/issue.json): HBM figure only; the "
                             "blend kernels are issue-bound, not HBM-bound (DESIGN.md 5.4)")

    extras = {}
    if ours and headline and world == 1 and not args.no_extras:
        arm.use_graph = graphed
        extras["u1"] = measure_u1(arm, views[0], vec_dev[0], targets_dev, flush)
        ref_arm = None
        try:
            ref_arm = ReferenceArm(wl, dev)
        except Exception as ex:  # oracle/_ref absent: the check against the reference arm is skipped, and said so
            extras["parity_note"] = f"reference arm unavailable: {ex}"
        if graphed:
            extras["parity"] = batch_parity(arm, views, vec_dev, targets_dev, B, ref_arm)
        del ref_arm
        torch.cuda.empty_cache()
        extras["stages_large"] = measure_large(dev, flush, load_peaks()[0])

    if rank != 0:
        return None
    cpu = None
    if ours and world == 1 and headline and not args.no_cpu_baseline:
        cpu = cpu_baseline(wl, views[0], args.cpu_iters)

    views_per_step = world * B  # B views per rank per step
    n_lanes = arm.multi.K if (ours and getattr(arm, "multi", None) is not None) else 1
    value = views_per_step * K / (ms_dev * 1e-3)
    e2e_val = views_per_step * K / (ms_e2e * 1e-3)
    tb = wl["targets"].element_size()
    h2d = PACK_FLOATS * 4 + 5 * X * tb
    if opt is None:
        comm = None if world == 1 else "one NCCL sum all-reduce of the flat gradient bucket per step (batch)"
    elif sharded:
        comm = ("none through NCCL: the gradient sum, Adam and the parameter broadcast are ONE kernel over NVLink peer "
                "memory (ghr_adam_step_sharded_synced, flag barriers inside the kernel)")
    else:
        comm = None if world == 1 else "one NCCL sum all-reduce of the gradients per step (batch), then the optimiser"
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": round(ms_dev / K, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "impl": args.impl,
        "config": {"workload": workload, "desc": WORKLOADS[workload]["desc"], "P": P, "W": W, "H": H,
                   "sh_degree": 3, "views_cycled": len(views) * world, "parallelism": f"view-sharded dp{world}",
                   "l2": ("flushed before every view (256 MiB fill outside the per-view event pair)" if n_lanes == 1 else
                          f"flushed before every group of {n_lanes} concurrent views (256 MiB fill outside the per-group event pair)"),
                   "lanes": n_lanes,
                   "sync_free": bool(args.sync_free),
                   "mode": (args.mode if ours else "eager"),
                   "batch_views_per_rank": B,
                   "step": f"{B} views per rank (LBS + cov + SH -> raster fwd -> loss -> raster bwd -> LBS bwd), gradient sum "
                           "over the ranks" + ("" if opt is None else ", optimiser step"),
                   "optimizer": opt_desc, "lr_scale": (None if opt is None else args.lr_scale),
                   "targets_dtype": str(wl["targets"].dtype).replace("torch.", ""),
                   "comm": comm, "all_reduce": comm},
        "e2e": {"value": round(e2e_val, 2), "unit": UNIT, "ms_per_step": round(ms_e2e / K, 4),
                "h2d_bytes_per_step": h2d * B, "d2h_bytes_per_step": 4 * B,
                "wall_clock_value": round(views_per_step * K / (wall_e2e * 1e-3), 2),
                "wall_clock_ms_per_step": round(wall_e2e / K, 4),
                "what": "public API step (GraphedRenderStep / MultiLaneRenderStep replay, or in eager mode "
                        "SMPLDeformer.deform_gaussians -> GaussianRasterizer -> loss -> backward); "
                        "camera + SMPL pose + 5 target planes copied from pinned host memory and the loss copied back to "
                        "pinned memory for every view inside the timed region (graph mode: input copies of the next view(s) on "
                        "copy streams during the current one(s), losses consumed one view / group later; "
                        "eager/reference: blocking loss.item()); Gaussian parameters stay resident (model state).  value = sum "
                        "of the per-group CUDA-event pairs; wall_clock_* = host clock around the whole K-step loop minus the "
                        f"L2 flushes ({flush_ms[0] * 1e3:.0f} us each), the cross-check that no copy hides between pairs"},
        "wall_clock_value": round(views_per_step * K / (wall_dev * 1e-3), 2),
        "gpu_launches": launches, "clocks": clocks,
    }
    if sharded_check is not None:
        line["sharded_step_check"] = sharded_check
    if eager is not None:
        line["eager"] = {"value": round(views_per_step * K / (eager[0] * 1e-3), 2), "ms_per_step": round(eager[0] / K, 4),
                         "e2e": round(views_per_step * K / (eager[1] * 1e-3), 2),
                         "what": "same step through the eager operator API (one Python call per kernel); host-bound"}
    if roof is not None:
        line["roofline"] = roof
        line["stages"] = stages
    line.update(extras)
    if cpu is not None:
        line["cpu_baseline"] = cpu
    if args.impl == "reference":
        line["cpu_baseline"] = {"kind": "reference", "cores": 0, "value": line["value"], "unit": UNIT,
                                "sample": "the reference's own implementation of this path is CUDA + PyTorch-on-CUDA, timed on "
                                          "the same GPU: LBS / covariance / SH = " + arm.lbs_source + "; rasteriser = its "
                                          "unmodified CUDA sources built for sm_100 (oracle/_ref); its CPU port is "
                                          "`--impl reference-cpu`"}
    del arm
    torch.cuda.empty_cache()
    return line


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
    K, Wm = args.steps, max(args.warmup, 3)
    try:
        line = run_workload(args, args.workload, world, rank, local, dev, K, Wm, headline=True)
    except (FileNotFoundError, OSError) as e:
        if args.impl != "reference":
            raise
        if rank == 0:
            emit({"impl": "reference", "unavailable": f"oracle/_ref not built: {e}"})
        return
    if world > 1 and args.workload == "zju_512" and not args.no_extras:
        # BASELINE.json config 4 (64 cameras x 300k Gaussians, view-sharded) rides along in multi-GPU runs
        K4 = max(2, min(K, 10))
        sub = run_workload(args, "multiview_300k", world, rank, local, dev, K4, 3, headline=False)
        if rank == 0 and sub is not None:
            line["multiview_300k"] = {k: sub[k] for k in ("value", "unit", "ms_per_step", "steps", "config", "e2e",
                                                          "wall_clock_value", "gpu_launches") if k in sub}
            if "sharded_step_check" in sub:
                line["multiview_300k"]["sharded_step_check"] = sub["sharded_step_check"]
    if (world == 1 and args.impl == "ours" and args.mode == "graph" and args.lanes > 1 and args.batch_views > 1
            and not args.no_extras):
        # Like for like with the reference arm's execution (ONE view at a time, L2 flushed before every view): the same
        # step on one lane.  The headline's 8 concurrent lanes are this repo's way of filling the GPU, not part of the
        # workload; this number takes them out of the comparison.
        a1 = argparse.Namespace(**vars(args))
        a1.lanes = 1
        sub = run_workload(a1, args.workload, world, rank, local, dev, max(10, min(K, 40)), 3, headline=False)
        if sub is not None:
            line["one_lane"] = {k: sub[k] for k in ("value", "unit", "ms_per_step", "steps", "e2e", "gpu_launches") if k in sub}
            line["one_lane"]["what"] = ("--lanes 1: one CUDA-graph replay per view, views of a batch one after the other, L2 "
                                        "flushed before every view -- the reference arm's execution pattern")
    if rank == 0:
        emit(line)


if __name__ == "__main__":
    main()
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
