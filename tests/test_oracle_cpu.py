"""CPU tests of the oracle itself (test infrastructure): internal invariants of
the rasteriser restatement, a finite-difference check of its backward, the
brute-force distCUDA2 against a KD-tree, and -- once generated on a B200 by the
unmodified reference CUDA sources -- the committed golden fixtures."""
import glob
import os

import numpy as np
import pytest

import oracle as O
from tests import common as Cm

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def case_and_state():
    case = Cm.raster_case(3000, 200, 152, seed=4, mode="sh_scale_rot")
    return case, Cm.oracle_forward(case)


def test_binning_invariants(case_and_state):
    case, st = case_and_state
    N = st.num_rendered
    assert N == int(st.tiles_touched.sum()) == int(st.point_offsets[-1])
    assert np.array_equal(np.cumsum(st.tiles_touched, dtype=np.uint64).astype(np.uint32), st.point_offsets)
    assert ((st.radii > 0) == (st.tiles_touched > 0)).all()
    keys = st.point_list_keys
    assert (np.diff(keys.astype(np.uint64)) >= 0).all() if N > 1 else True     # sortedness
    # stability: equal keys keep ascending emission order == ascending Gaussian id
    same = keys[1:] == keys[:-1]
    assert (st.point_list[1:][same] > st.point_list[:-1][same]).all()
    # sorted is a permutation of emitted
    assert np.array_equal(np.sort(st.keys_unsorted), keys)
    assert np.array_equal(np.sort(st.vals_unsorted), np.sort(st.point_list))
    # ranges partition [0,N) by tile
    tile = (keys >> np.uint64(32)).astype(np.int64)
    for t in np.unique(tile):
        a, b = st.ranges[t]
        assert (tile[a:b] == t).all() and (a == 0 or tile[a - 1] != t) and (b == N or tile[b] != t)
    untouched = np.setdiff1d(np.arange(st.ranges.shape[0]), np.unique(tile))
    assert (st.ranges[untouched] == 0).all()
    # depth half of the key is the float bit pattern of the view-space depth
    d = (keys & np.uint64(0xFFFFFFFF)).astype(np.uint32).view(np.float32)
    assert np.array_equal(d, st.depths[st.point_list]) and (d > 0.2).all()


def test_blend_invariants(case_and_state):
    case, st = case_and_state
    assert (st.final_T <= 1).all() and (st.final_T >= 0).all()
    assert (st.alpha >= 0).all() and (st.alpha <= 1 + 1e-5).all()
    # alpha = sum(alpha_i T_i) = 1 - T exactly in real arithmetic
    assert np.abs(st.alpha[0] + st.final_T - 1).max() < 1e-5
    lens = (st.ranges[:, 1] - st.ranges[:, 0]).reshape(-1)
    gx = (case["W"] + 15) // 16
    H, W = case["H"], case["W"]
    yy, xx = np.mgrid[0:H, 0:W]
    assert (st.n_contrib <= lens[(yy // 16) * gx + xx // 16]).all()
    # background shows through exactly T*bg where nothing contributed
    empty = st.n_contrib == 0
    for c in range(3):
        assert np.array_equal(st.color[c][empty], np.full(empty.sum(), case["bg"][c], np.float32))


def test_linearity_in_colour():
    """Blend is linear in the per-Gaussian colours: render(a*c1 + b*c2) = a*render(c1) + b*render(c2) (bg = 0)."""
    case = Cm.raster_case(1500, 96, 80, seed=8, mode="precomp", bg=(0, 0, 0))
    rng = np.random.default_rng(0)
    c1 = case["colors"]; c2 = rng.uniform(0, 1, c1.shape).astype(np.float32)
    r = []
    for c in (c1, c2, 0.25 * c1 + 0.5 * c2):
        cc = dict(case); cc["colors"] = c.astype(np.float32)
        r.append(Cm.oracle_forward(cc).color)
    assert np.abs(0.25 * r[0] + 0.5 * r[1] - r[2]).max() < 2e-6


def test_backward_matches_finite_differences():
    case = Cm.raster_case(600, 64, 48, seed=2, mode="precomp")
    st = Cm.oracle_forward(case)
    g = Cm.oracle_backward(case, st)

    def loss(c):
        s = Cm.oracle_forward(c)
        return float((s.color.astype(np.float64) * c["dL_dcolor"]).sum() + (s.depth.astype(np.float64) * c["dL_ddepth"]).sum()
                     + (s.alpha.astype(np.float64) * c["dL_dalpha"]).sum())
    vis = np.nonzero(st.radii > 0)[0]
    top = vis[np.argsort(-np.abs(g["dL_dopacity"][vis, 0]))[:4]]
    for i in top:
        for name, key, eps, comps in (("dL_dopacity", "opacity", 1e-3, [0]), ("dL_dcolors", "colors", 1e-2, [0, 1, 2])):
            for k in comps:
                cp = dict(case); cm = dict(case)
                a = case[key].copy(); a[i, k] += eps; cp[key] = a
                b = case[key].copy(); b[i, k] -= eps; cm[key] = b
                fd = (loss(cp) - loss(cm)) / (2 * eps)
                assert abs(fd - g[name][i, k]) <= 2e-2 * max(abs(fd), 1e-2), (name, i, k, fd, g[name][i, k])


def test_knn_brute_force_matches_kdtree():
    from scipy.spatial import cKDTree
    pts = Cm.smpl()["v_template"].numpy()
    ours = O.dist2_knn3(pts)
    d, _ = cKDTree(pts.astype(np.float64)).query(pts.astype(np.float64), k=4)
    np.testing.assert_allclose(ours, (d[:, 1:] ** 2).mean(1), rtol=2e-6)
    # fewer than 4 points: unfilled slots stay FLT_MAX as in simple_knn.cu:156,166-168
    assert np.all(O.dist2_knn3(pts[:3]) > 1e37) and np.all(np.isinf(O.dist2_knn3(pts[:2])))


def test_higher_msb():
    # rasterizer_impl.cu:35-50; SURVEY 8 constants
    assert [O.higher_msb(n) for n in (1, 2, 1024, 4096, 8160)] == [1, 2, 11, 13, 13]


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "raster_*.npz"))) or [None])
def test_oracle_against_reference_cuda_golden(path):
    """Golden vectors produced on a B200 by the UNMODIFIED reference CUDA sources
    (oracle/gen_golden_raster.py).  Integers bit-exact; floats to tolerance."""
    if path is None:
        pytest.skip("golden fixtures not generated yet")
    z = np.load(path)
    case = {k[3:]: z[k] for k in z.files if k.startswith("in_")}
    for k in ("P", "W", "H", "degree"):
        case[k] = int(case[k])
    for k in ("tanfovx", "tanfovy", "scale_modifier"):
        case[k] = float(case[k])
    for k in ("colors", "cov3D", "scales", "rotations", "sh"):
        if k not in case or case[k].size == 0:
            case[k] = None
    st = Cm.oracle_forward(case)
    assert st.num_rendered == int(z["num_rendered"])
    for name in ("radii", "tiles_touched", "point_offsets", "point_list_keys", "point_list", "ranges"):
        assert np.array_equal(st[name], z[name]), name
    vis = st.radii > 0
    for name in ("depths", "means2D", "conic_opacity"):
        assert np.array_equal(st[name][vis], z[name][vis]), name
    assert np.abs(st.color - z["color"]).max() <= 1e-5
    assert np.abs(st.alpha - z["alpha"]).max() <= 1e-5
    assert np.abs(st.depth - z["depth"]).max() <= 1e-5 * max(1.0, np.abs(z["depth"]).max())
    assert (st.n_contrib != z["n_contrib"]).mean() <= 2e-3
    g = Cm.oracle_backward(case, st)
    for name in ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
                 "dL_drotations"):
        ref = z["g_" + name].astype(np.float64)
        if not ref.size:
            assert g[name].size == 0, name
            continue
        scale = max(np.abs(ref).max(), 1e-30)
        # the golden gradients carry the reference's fp32 atomic-order noise and libdevice expf
        assert np.abs(g[name] - ref).max() <= 2e-3 * scale, name
