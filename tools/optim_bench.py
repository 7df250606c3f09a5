"""Optimiser-side kernels against the HBM roofline:

  * FusedAdam.step (ghr_adam_step: every parameter group of the reference's training_setup in ONE launch) vs
    torch.optim.Adam (foreach, the reference's configuration) and torch.optim.Adam(fused=True), at 50 k / 300 k / 3 M
    Gaussians x 59 floats; algorithmic bytes = 28 per parameter (read p, g, m, v; write p, m, v), +4 with zero_grads;
  * add_densification_stats (ghr_densify_stats) vs the reference's three indexed ATen statements.

CUDA events on the launching stream around each call, L2 flushed (256 MiB fill) before every timed call, median.
    python tools/optim_bench.py > profiles/rNN_optim.json
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mygauhuman_b200.optim import FusedAdam, add_densification_stats  # noqa: E402

DEV = torch.device("cuda")
SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}
LR = {"xyz": 0.00016, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}
_flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)


def timed(fn, iters=30, warm=5):
    for _ in range(warm):
        fn()
    ts = []
    for i in range(iters):
        _flush.fill_(i & 0xFF)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def peak():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6500.0, "fallback (B200_PROFILING.md)"


def adam(P):
    g = torch.Generator().manual_seed(0)
    mk = lambda: {n: torch.nn.Parameter(torch.randn((P,) + s, generator=g).to(DEV)) for n, s in SHAPES.items()}
    groups = lambda ps: [{"params": [ps[n]], "lr": LR[n], "name": n} for n in LR]
    arms = {}
    for name, ctor in (("ours", lambda ps: FusedAdam(groups(ps), lr=0.0, eps=1e-15)),
                       ("torch_foreach", lambda ps: torch.optim.Adam(groups(ps), lr=0.0, eps=1e-15)),
                       ("torch_fused", lambda ps: torch.optim.Adam(groups(ps), lr=0.0, eps=1e-15, fused=True))):
        ps = mk()
        for p in ps.values():
            p.grad = torch.randn_like(p)
        opt = ctor(ps)
        arms[name] = timed(opt.step)
        if name == "ours":
            arms["ours_zero_grads"] = timed(lambda: opt.step(zero_grads=True))
        del opt, ps
    n = P * 59
    pk, src = peak()
    gbs = lambda ms, b: round(n * b / (ms * 1e-3) / 1e9, 1)
    return {"P": P, "params": n, "alg_bytes": n * 28,
            "ours_ms": round(arms["ours"], 4), "ours_GBps": gbs(arms["ours"], 28),
            "ours_frac_of_hbm_peak": round(gbs(arms["ours"], 28) / pk, 3),
            "ours_zero_grads_ms": round(arms["ours_zero_grads"], 4), "ours_zero_grads_GBps": gbs(arms["ours_zero_grads"], 32),
            "torch_foreach_ms": round(arms["torch_foreach"], 4), "torch_fused_ms": round(arms["torch_fused"], 4),
            "peak_GBps": pk, "peak_source": src}


def stats(P):
    g = torch.Generator().manual_seed(1)
    radii = torch.randint(0, 40, (P,), generator=g, dtype=torch.int32).to(DEV)
    radii[torch.rand(P, generator=g).to(DEV) < 0.3] = 0
    vgrad = (torch.randn(P, 3, generator=g) * 1e-3).to(DEV)
    mr, acc, dn = torch.zeros(P, device=DEV), torch.zeros(P, 1, device=DEV), torch.zeros(P, 1, device=DEV)

    def reference():  # train.py:401-404, gaussian_model.py:764-766
        vis = radii > 0
        mr[vis] = torch.max(mr[vis], radii[vis])
        acc[vis] += torch.norm(vgrad[vis, :2], dim=-1, keepdim=True)
        dn[vis] += 1

    t_ours = timed(lambda: add_densification_stats(radii, vgrad, mr, acc, dn))
    t_ref = timed(reference)
    return {"P": P, "ours_ms": round(t_ours, 4), "reference_statements_ms": round(t_ref, 4), "alg_bytes": 40 * P,
            "ours_GBps": round(40 * P / (t_ours * 1e-3) / 1e9, 1)}


if __name__ == "__main__":
    out = {"adam": [adam(P) for P in (50_000, 300_000, 3_000_000)],
           "densify_stats": [stats(P) for P in (50_000, 300_000, 3_000_000)],
           "gpu": torch.cuda.get_device_name(0)}
    print(json.dumps(out))
