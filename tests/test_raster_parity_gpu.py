"""-m gpu parity tests of the rasteriser: our CUDA path (through the drop-in
_C / C ABI) against (1) the UNMODIFIED reference CUDA build (oracle/_ref,
bit-exact integer state), (2) the CPU oracle, on the same seeded inputs.

Bars (BASELINE.json north_star): radii / tiles_touched / point_offsets /
num_rendered / sorted keys+ids / tile ranges / n_contrib bit-exact;
color / depth / alpha <= 1e-5 abs; gradients <= 1e-4 rel.
"""
import numpy as np
import pytest
import torch

from tests import common as Cm

pytestmark = pytest.mark.gpu

ABS_IMG = 1e-5
REL_GRAD = 1e-4

CASES = [
    # (P, W, H, mode, degree, scale)
    (2000, 128, 96, "precomp", 3, 0.01),
    (3000, 200, 152, "sh_scale_rot", 3, 0.01),     # baking.py path; H,W not multiples of 16
    (1500, 96, 80, "sh_cov", 2, 0.02),
    (1500, 96, 80, "color_scale_rot", 0, 0.02),
    (20000, 512, 512, "precomp", 3, 0.01),
    (800, 64, 64, "precomp", 3, 0.3),              # screen-filling splats: cooperative key emission
]


def _state(tc, fw):
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    n, color, depth, alpha, radii, geom, binning, img = fw
    return _C.state_views(tc["P"], n, tc["W"], tc["H"], geom, binning, img)


def _np(t):
    return t.detach().cpu().numpy()


def _grad_close(name, a, b, rel=REL_GRAD, stragglers_per_million=0):
    """|a - b| <= rel |b| + 0.1 rel max|b| per element.  stragglers_per_million > 0 (the comparisons of >= 10^6 elements
    against the reference CUDA build, where BOTH sides sum with fp32 atomics in a run-dependent order): that many
    elements per million may sit up to 10x above the floor term (1e-4 of the tensor's scale instead of 1e-5), never
    further -- seen once: 1 of 2 400 000 rotation gradients at 1.17e-5 of the scale."""
    a = np.asarray(a, np.float64).reshape(-1); b = np.asarray(b, np.float64).reshape(-1)
    assert a.shape == b.shape, name
    scale = max(np.abs(b).max(), 1e-30) if b.size else 1.0
    err = np.abs(a - b)
    bad = err > rel * np.abs(b) + rel * 1e-1 * scale
    far = err > rel * np.abs(b) + rel * scale
    allowed = int(stragglers_per_million * a.size // 1_000_000)
    assert bad.sum() <= allowed and not far.any(), \
        f"{name}: {bad.sum()} of {a.size} elements off, max err {err.max():.3e} (scale {scale:.3e})"


def _ref_forward(tc):
    from oracle import ref_cuda
    ref = ref_cuda.RefRasterizer()
    n, color, depth, alpha, radii = ref.forward(tc["bg"], tc["means3D"], tc["colors"], tc["opacity"], tc["scales"],
                                                tc["rotations"], tc["scale_modifier"], tc["cov3D"], tc["view"],
                                                tc["proj"], tc["tanfovx"], tc["tanfovy"], tc["H"], tc["W"], tc["sh"],
                                                tc["degree"], tc["campos"])
    return ref, (n, color, depth, alpha, radii)


@pytest.mark.parametrize("P,W,H,mode,degree,scale", CASES)
def test_forward_backward_vs_reference_cuda(P, W, H, mode, degree, scale):
    from oracle import ref_cuda
    assert ref_cuda.available(), "oracle/_ref/libref_gauhuman.so is missing (built by __graft_entry__.build(); ships with the snapshot)"
    case = Cm.raster_case(P, W, H, seed=P + W, mode=mode, degree=degree, scale=scale)
    tc = Cm.to_torch(case, "cuda")
    fw = Cm.ours_forward(tc)
    ref, (rn, rcolor, rdepth, ralpha, rradii) = _ref_forward(tc)
    n, color, depth, alpha, radii = fw[:5]
    # ---- integers: bit-exact ----
    assert n == rn
    assert torch.equal(radii, rradii)
    st = _state(tc, fw)
    rs = ref.state()
    assert np.array_equal(_np(st["tiles_touched"]).view(np.uint32), rs["tiles_touched"])
    assert np.array_equal(_np(st["point_offsets"]).view(np.uint32), rs["point_offsets"])
    assert np.array_equal(_np(st["point_list_keys"]).view(np.uint64), rs["point_list_keys"])
    assert np.array_equal(_np(st["point_list"]).view(np.uint32), rs["point_list"])
    assert np.array_equal(_np(st["ranges"]).view(np.uint32), rs["ranges"])
    assert np.array_equal(_np(st["n_contrib"]).view(np.uint32), rs["n_contrib"])
    # ---- geometry floats feed the integers: also exact ----
    vis = _np(radii) > 0
    assert np.array_equal(_np(st["depths"])[vis], rs["depths"][vis])
    assert np.array_equal(_np(st["means2D"])[vis], rs["means2D"][vis])
    assert np.array_equal(_np(st["conic_opacity"])[vis], rs["conic_opacity"][vis])
    # ---- images ----
    assert (color - rcolor).abs().max().item() <= ABS_IMG
    assert (depth - rdepth).abs().max().item() <= ABS_IMG * max(1.0, rdepth.abs().max().item())
    assert (alpha - ralpha).abs().max().item() <= ABS_IMG
    assert np.abs(_np(st["final_T"]) - rs["final_T"]).max() <= ABS_IMG
    if mode.startswith("sh"):
        assert np.abs(_np(st["rgb"])[vis] - rs["rgb"][vis]).max() <= ABS_IMG
    # ---- backward ----
    g = Cm.ours_backward(tc, fw)
    rg = ref.backward(tc["bg"], tc["means3D"], rradii, tc["colors"], tc["scales"], tc["rotations"],
                      tc["scale_modifier"], tc["cov3D"], tc["view"], tc["proj"], tc["tanfovx"], tc["tanfovy"],
                      tc["dL_dcolor"], tc["dL_ddepth"], tc["dL_dalpha"], tc["sh"], tc["degree"], tc["campos"], ralpha)
    names = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations"]
    for name, ours in zip(names, g):
        assert tuple(ours.shape) == tuple(rg[name].shape), name
        _grad_close(name, _np(ours), _np(rg[name]))


@pytest.mark.parametrize("P,W,H,mode,degree,scale", CASES[:4])
def test_forward_backward_vs_cpu_oracle(P, W, H, mode, degree, scale):
    case = Cm.raster_case(P, W, H, seed=P + W, mode=mode, degree=degree, scale=scale)
    tc = Cm.to_torch(case, "cuda")
    fw = Cm.ours_forward(tc)
    st_o = Cm.oracle_forward(case)
    n, color, depth, alpha, radii = fw[:5]
    st = _state(tc, fw)
    assert n == st_o.num_rendered
    assert np.array_equal(_np(radii), st_o.radii)
    assert np.array_equal(_np(st["tiles_touched"]).view(np.uint32), st_o.tiles_touched)
    assert np.array_equal(_np(st["point_list_keys"]).view(np.uint64), st_o.point_list_keys)
    assert np.array_equal(_np(st["point_list"]).view(np.uint32), st_o.point_list)
    assert np.array_equal(_np(st["ranges"]).view(np.uint32), st_o.ranges)
    # glibc expf != libdevice expf: images to 1e-5, n_contrib may flip on threshold pixels
    assert np.abs(_np(color) - st_o.color).max() <= ABS_IMG
    assert np.abs(_np(alpha) - st_o.alpha).max() <= ABS_IMG
    assert np.abs(_np(depth) - st_o.depth).max() <= ABS_IMG * max(1.0, np.abs(st_o.depth).max())
    mism = (_np(st["n_contrib"]).view(np.uint32) != st_o.n_contrib).mean()
    assert mism <= 2e-3, f"n_contrib differs on {mism:.2%} of pixels"
    if mism == 0:
        g = Cm.ours_backward(tc, fw)
        go = Cm.oracle_backward(case, st_o)
        names = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
                 "dL_drotations"]
        for name, ours in zip(names, g):
            _grad_close(name, _np(ours), go[name])


def test_empty_and_culled():
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    case = Cm.raster_case(64, 48, 40, seed=3)
    tc = Cm.to_torch(case, "cuda")
    # P == 0 (rasterize_points.cu:83: outputs stay zero, buffers empty)
    tc0 = dict(tc)
    for k in ("means3D", "colors", "cov3D", "opacity"):
        tc0[k] = tc[k][:0]
    tc0["P"] = 0
    n, color, depth, alpha, radii, geom, binning, img = Cm.ours_forward(tc0)
    assert n == 0 and color.abs().max().item() == 0 and radii.numel() == 0 and geom.numel() == 0
    # every Gaussian behind the camera: background only, num_rendered == 0
    tcb = dict(tc)
    tcb["means3D"] = tc["means3D"] + torch.tensor([0.0, 0.0, 100.0], device="cuda")
    fw = Cm.ours_forward(tcb)
    assert fw[0] == 0 and int(fw[4].abs().max()) == 0
    bg = tc["bg"].view(3, 1, 1)
    assert (fw[1] - bg).abs().max().item() == 0 and fw[3].abs().max().item() == 0
    g = Cm.ours_backward(tcb, fw)
    assert all(float(t.abs().max()) == 0 for t in g if t.numel())


def test_mark_visible_matches_reference():
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    from oracle import ref_cuda
    import oracle as O
    case = Cm.raster_case(5000, 64, 64, seed=9, radius=0.05)  # camera inside the body
    tc = Cm.to_torch(case, "cuda")
    ours = _C.mark_visible(tc["means3D"], tc["view"], tc["proj"])
    assert np.array_equal(_np(ours), O.mark_visible(case["means3D"], case["view"], case["proj"]))
    if ref_cuda.available():
        assert torch.equal(ours, ref_cuda.mark_visible(tc["means3D"], tc["view"], tc["proj"]))
    assert 0 < int(ours.sum()) < 5000


def test_sync_free_capacity_and_overflow():
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    case = Cm.raster_case(3000, 128, 96, seed=5)
    tc = Cm.to_torch(case, "cuda")
    exact = Cm.ours_forward(tc)
    try:
        _C.set_sync_free_capacity(4 * exact[0])
        fw = Cm.ours_forward(tc)
        assert fw[0] == (4 * exact[0] + 63) // 64 * 64  # capacities are rounded up to 64 pairs
        assert _C.read_status(fw[5]) == (exact[0], 0)
        for a, b in zip(exact[1:5], fw[1:5]):
            assert torch.equal(a, b)
        g0 = Cm.ours_backward(tc, exact); g1 = Cm.ours_backward(tc, fw)
        for a, b in zip(g0, g1):
            _grad_close("sync_free", _np(b), _np(a), rel=1e-5)
        _C.set_sync_free_capacity(max(exact[0] // 2, 1))
        fw = Cm.ours_forward(tc)
        n, err = _C.read_status(fw[5])
        assert n == exact[0] and err == 4  # GHR_ERR_CAPACITY latched, nothing rendered out of bounds
    finally:
        _C.set_sync_free_capacity(None)


def test_optimistic_launch_recovers_from_a_small_guess():
    """Default mode sizes the binning buffer from the previous call with the same (P, W, H).  A following call that
    needs far more pairs must still be exact (the render stage is re-run with an exact-size buffer)."""
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    small = Cm.raster_case(3000, 128, 96, seed=5, scale=0.004)
    big = Cm.raster_case(3000, 128, 96, seed=5, scale=0.1)
    ts, tb = Cm.to_torch(small, "cuda"), Cm.to_torch(big, "cuda")
    _C._cap_hint.clear()
    ref_big = Cm.ours_forward(tb)          # first call: blocking read, exact size
    _C._cap_hint.clear()
    n_small = Cm.ours_forward(ts)[0]       # seeds the guess
    assert ref_big[0] > max(_C._cap_hint.values())  # the seeded guess is too small for the big case
    fw = Cm.ours_forward(tb)               # guess too small -> redo
    assert fw[0] == ref_big[0]
    for a, b in zip(ref_big[1:5], fw[1:5]):
        assert torch.equal(a, b)
    g0 = Cm.ours_backward(tb, ref_big); g1 = Cm.ours_backward(tb, fw)
    for a, b in zip(g0, g1):
        _grad_close("optimistic", _np(b), _np(a), rel=1e-4)  # two runs of the same kernels: fp32 atomic order differs
    fw2 = Cm.ours_forward(tb)              # guess now large enough: single pass
    assert fw2[0] == ref_big[0] and torch.equal(fw2[1], ref_big[1])


def test_backward_after_many_outstanding_forwards_and_a_redo_in_debug_mode():
    """the reference's render() keeps 7 forwards outstanding before the first backward, so backward must find
    the capacity a binning buffer was laid out for from the buffer itself (ghr_binning_capacity), not from a side table
    that other forwards, address reuse or saved-tensor hooks can invalidate; and debug=True must not raise after a
    successful redo of an undersized optimistic launch (the CAPACITY latch is cleared by the re-run)."""
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    cases = [Cm.to_torch(Cm.raster_case(3000, 128, 96, seed=5 + k, scale=s), "cuda")
             for k, s in enumerate((0.004, 0.1, 0.01, 0.05, 0.004, 0.1, 0.02, 0.03))]
    _C._cap_hint.clear()
    exact = []
    for tc in cases:                       # references: one forward + backward at a time, exact-size buffers
        _C._cap_hint.clear()
        fw = Cm.ours_forward(tc)
        exact.append((fw, Cm.ours_backward(tc, fw)))
    _C._cap_hint.clear()
    fws = [Cm.ours_forward(tc) for tc in cases]   # 8 outstanding forwards; guesses too small AND too large occur
    caps = [_C._capacity_of(fw[6], tc["P"], tc["W"], tc["H"]) for tc, fw in zip(cases, fws)]
    assert any(c > fw[0] + 64 for c, fw in zip(caps, fws)), "some buffer must be larger than its num_rendered"
    for (tc, fw, (efw, eg)) in reversed(list(zip(cases, fws, exact))):
        assert fw[0] == efw[0] and torch.equal(fw[1], efw[1])
        moved = (fw[0], *fw[1:5], fw[5].clone(), fw[6].clone(), fw[7].clone())  # saved tensors relocated by a hook
        for a, b in zip(Cm.ours_backward(tc, moved), eg):
            _grad_close("outstanding", _np(a), _np(b), rel=1e-4)
    # wrong-size buffer: rejected, not mis-read
    tc, fw = cases[1], fws[1]
    bad = (fw[0], *fw[1:5], fw[5], torch.cat([fw[6], fw[6][:64]]), fw[7])
    with pytest.raises(Exception):
        Cm.ours_backward(tc, bad)
    # debug=True after a redo
    _C._cap_hint.clear()
    Cm.ours_forward(cases[0])                       # seeds a small guess
    e = Cm.empty
    tb = cases[1]
    out = _C.rasterize_gaussians(tb["bg"], tb["means3D"], tb["colors"], tb["opacity"], e(), e(), 1.0, tb["cov3D"],
                                 tb["view"], tb["proj"], tb["tanfovx"], tb["tanfovy"], tb["H"], tb["W"], e(), 3,
                                 tb["campos"], False, True)
    assert out[0] == exact[1][0][0] and torch.equal(out[1], exact[1][0][1])
    assert _C.read_status(out[5]) == (out[0], 0)


def test_autograd_module_matches_raw_calls():
    """GaussianRasterizer.forward + autograd == raw _C calls (drop-in surface of render())."""
    from mygauhuman_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    case = Cm.raster_case(2000, 128, 96, seed=11)
    tc = Cm.to_torch(case, "cuda")
    rs = GaussianRasterizationSettings(image_height=tc["H"], image_width=tc["W"], tanfovx=tc["tanfovx"],
                                       tanfovy=tc["tanfovy"], bg=tc["bg"], scale_modifier=1.0, viewmatrix=tc["view"],
                                       projmatrix=tc["proj"], sh_degree=3, campos=tc["campos"], prefiltered=False,
                                       debug=False)
    means3D = tc["means3D"].clone().requires_grad_(True)
    means2D = torch.zeros_like(means3D, requires_grad=True)
    colors = tc["colors"].clone().requires_grad_(True)
    opac = tc["opacity"].clone().requires_grad_(True)
    cov = tc["cov3D"].clone().requires_grad_(True)
    color, radii, depth, alpha = GaussianRasterizer(rs)(means3D=means3D, means2D=means2D, opacities=opac,
                                                        colors_precomp=colors, cov3D_precomp=cov)
    loss = (color * tc["dL_dcolor"]).sum() + (depth * tc["dL_ddepth"]).sum() + (alpha * tc["dL_dalpha"]).sum()
    loss.backward()
    fw = Cm.ours_forward(tc)
    g = Cm.ours_backward(tc, fw)
    assert torch.equal(color, fw[1]) and torch.equal(radii, fw[4])
    _grad_close("means2D", _np(means2D.grad), _np(g[0]), rel=1e-5)
    _grad_close("colors", _np(colors.grad), _np(g[1]), rel=1e-5)
    _grad_close("opacity", _np(opac.grad), _np(g[2]), rel=1e-5)
    _grad_close("means3D", _np(means3D.grad), _np(g[3]), rel=1e-5)
    _grad_close("cov3D", _np(cov.grad), _np(g[4]), rel=1e-5)
    with pytest.raises(Exception):
        GaussianRasterizer(rs)(means3D=means3D, means2D=means2D, opacities=opac, cov3D_precomp=cov)
    with pytest.raises(Exception):
        GaussianRasterizer(rs)(means3D=means3D, means2D=means2D, opacities=opac, colors_precomp=colors)


def test_throughput_regime_kernels_pass_the_same_parity_tests():
    """Frames with more than 2^22 (Gaussian, tile) pairs take the other blend kernel family (ghr_blend_*_wide.cu,
    8x4-pixel blocks, no block masks).  The regime is chosen once per process, so the forward/backward parity tests of
    this file are re-run in a child process with GHR_BLEND_REGIME=throughput."""
    import os
    import subprocess
    import sys
    if os.environ.get("GHR_BLEND_REGIME") or os.environ.get("GHR_SHARED_GPU"):
        pytest.skip("already inside a forced-regime run")
    env = dict(os.environ, GHR_BLEND_REGIME="throughput")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", "tests/test_raster_parity_gpu.py", "-x", "-q", "-m", "gpu", "-W", "ignore",
                        "-p", "no:cacheprovider"], cwd=root, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


def test_shared_gpu_blend_configuration_passes_the_same_parity_tests():
    """ghr_set_shared_gpu(1) (views of a batch on several streams, MultiLaneRenderStep) switches the blend kernels to
    128-entry staging rounds.  The setting is process-wide, so the parity tests of this file are re-run in a child
    process with GHR_SHARED_GPU=1."""
    import os
    import subprocess
    import sys
    if os.environ.get("GHR_BLEND_REGIME") or os.environ.get("GHR_SHARED_GPU"):
        pytest.skip("already inside a forced-regime run")
    env = dict(os.environ, GHR_SHARED_GPU="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", "tests/test_raster_parity_gpu.py", "-x", "-q", "-m", "gpu", "-W", "ignore",
                        "-p", "no:cacheprovider"], cwd=root, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
