"""-m gpu parity of the onesweep sort (vs torch stable sort and the reference's
CUB call) and distCUDA2 (vs brute-force oracle and the reference simple-knn)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def ghr_sort(keys, vals, bits):
    from mygauhuman_b200._lib import check, lib
    l = lib()
    N = keys.numel()
    ka = keys.clone(); va = vals.clone()
    kb = torch.empty_like(ka); vb = torch.empty_like(va)
    ws = torch.zeros(max(l.ghr_sort_workspace_bytes(N), 256), dtype=torch.uint8, device=keys.device)
    sel = torch.zeros(1, dtype=torch.int32, device=keys.device)
    check(l.ghr_sort_pairs_u64(N, bits, ka.data_ptr(), kb.data_ptr(), va.data_ptr(), vb.data_ptr(), ws.data_ptr(),
                               ws.numel(), sel.data_ptr(), torch.cuda.current_stream().cuda_stream), "sort")
    s = int(sel.item())
    return (kb, vb) if s else (ka, va)


def make_keys(N, tiles, seed, depth_lo=0.2, depth_hi=10.0):
    g = torch.Generator().manual_seed(seed)
    tile = torch.randint(0, tiles, (N,), generator=g, dtype=torch.int64)
    depth = torch.rand(N, generator=g) * (depth_hi - depth_lo) + depth_lo
    keys = (tile << 32) | depth.view(torch.int32).to(torch.int64)
    return keys.cuda(), torch.arange(N, dtype=torch.int32).cuda()


# every keys-per-thread regime of the pass kernel (2 / 4 / 8 / 12: up to 2^17 / 3*2^17 / 5*2^17 keys / above), full and
# ragged last tiles
@pytest.mark.parametrize("N", [0, 1, 31, 1000, 1025, 4097, 100003, 300_001, 3 << 17, 600_011, 1 << 20, (1 << 20) + 7,
                               5_000_011])
def test_onesweep_matches_stable_sort(N):
    tiles = 1024
    bits = 32 + 10
    keys, vals = make_keys(N, tiles, seed=N)
    if N > 10:
        keys[N // 2:N // 2 + 5] = keys[N // 2]  # exact ties: stability decides
    k, v = ghr_sort(keys, vals, bits)
    ks, order = torch.sort(keys, stable=True)
    assert torch.equal(k[:N], ks)
    assert torch.equal(v[:N].long(), order)


def test_onesweep_skips_constant_digit_passes_and_matches_cub():
    from oracle import ref_cuda
    N = 300000
    keys, vals = make_keys(N, 4096, seed=1, depth_lo=2.0, depth_hi=3.9)  # top depth byte constant
    k, v = ghr_sort(keys, vals, 32 + 12)
    ks, order = torch.sort(keys, stable=True)
    assert torch.equal(k, ks) and torch.equal(v.long(), order)
    if ref_cuda.available():
        kc, vc, _ = ref_cuda.cub_sort_pairs(keys, vals, 32 + 12)
        torch.cuda.synchronize()
        assert torch.equal(k, kc) and torch.equal(v, vc)


@pytest.mark.parametrize("N", [50_000, 1_500_000])
def test_onesweep_keys_in_runs_and_runs_mixed_with_random_keys(N):
    """The pass kernel ranks a warp's keys with MATCH.ANY when their digits come in runs (a frame: the tiles of one
    Gaussian share their depth) and with one ballot per digit bit otherwise; the choice is per warp, so a sort whose
    first half is runs and second half is random takes both paths in the same launch.  Stability decides inside a run."""
    base, _ = make_keys(N, 1024, seed=7)
    runs = base[: N // 2].repeat_interleave(5)[: N // 2]          # runs of five identical keys
    keys = torch.cat([runs, base[N // 2:]]).contiguous()
    vals = torch.arange(N, dtype=torch.int32).cuda()
    k, v = ghr_sort(keys, vals, 32 + 10)
    ks, order = torch.sort(keys, stable=True)
    assert torch.equal(k[:N], ks) and torch.equal(v[:N].long(), order)


def test_onesweep_full_64bit_and_odd_bits():
    g = torch.Generator().manual_seed(3)
    N = 70001
    keys = torch.randint(-(1 << 62), 1 << 62, (N,), generator=g, dtype=torch.int64).cuda()
    vals = torch.arange(N, dtype=torch.int32).cuda()
    for bits in (64, 37, 9):
        k, v = ghr_sort(keys, vals, bits)
        masked = keys.cpu().numpy().view(np.uint64) & np.uint64((1 << bits) - 1 if bits < 64 else 0xFFFFFFFFFFFFFFFF)
        order = np.argsort(masked, kind="stable")
        assert np.array_equal(v.cpu().numpy(), order.astype(np.int32))


@pytest.mark.parametrize("P", [1, 2, 3, 4, 50, 3000, 4096, 4097, 6890, 70000, 200000])
def test_dist2_knn3(P):
    import oracle as O
    from oracle import ref_cuda
    from mygauhuman_b200.simple_knn._C import distCUDA2
    from tests import common as Cm
    rng = np.random.default_rng(P)
    if P == 6890:
        pts = Cm.smpl()["v_template"].numpy()
    else:
        pts = (rng.normal(size=(P, 3)) * np.array([1.0, 2.0, 0.3])).astype(np.float32)
        if P > 10:
            pts[5] = pts[6]  # duplicate point -> zero distance
    ours = distCUDA2(torch.from_numpy(pts).cuda()).cpu().numpy()
    if P <= 70000:
        ref = O.dist2_knn3(pts)
    else:
        from scipy.spatial import cKDTree
        d, _ = cKDTree(pts.astype(np.float64)).query(pts.astype(np.float64), k=4)
        ref = (d[:, 1:] ** 2).mean(1).astype(np.float32)
    if P < 4:
        assert np.array_equal(ours, ref) and np.all(ours > 1e37)  # FLT_MAX slots, simple_knn.cu:156
    elif P <= 70000:
        np.testing.assert_allclose(ours, ref, rtol=1e-6, atol=0)
    else:
        np.testing.assert_allclose(ours, ref, rtol=1e-4, atol=1e-12)
    if ref_cuda.available() and P >= 4:
        r = ref_cuda.dist_cuda2(torch.from_numpy(pts).cuda()).cpu().numpy()
        np.testing.assert_allclose(ours, r, rtol=1e-6, atol=0)
