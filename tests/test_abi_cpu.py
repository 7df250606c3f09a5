"""CPU checks of the drop-in boundary: libghr.so loads and exports every
symbol include/ghr.h declares; argument validation and layout (no compute)."""
import ctypes as C
import os
import re

import pytest

from mygauhuman_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "ghr.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ghr_[a-z0-9_]+)\s*\(", text)))


def test_library_loads_and_exports_every_declared_symbol():
    l = _lib.lib()
    syms = declared_symbols()
    assert len(syms) >= 11
    for s in syms:
        assert hasattr(l, s), f"{s} declared in include/ghr.h but not exported"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes signature"
    assert l.ghr_version() == 1


def test_error_strings():
    assert _lib.error_string(0) == "ok"
    for code in (-1, -2, -3, -4, -5):
        assert _lib.error_string(code) not in ("ok", "unknown status")
    assert "prefiltered" in _lib.error_string(-5)


def test_layout_is_aligned_and_monotone():
    L = _lib.state_layout(50000, 180000, 512, 512)
    assert L.sort_bits == 43 and L.sort_passes == 5          # ZJU 512^2 -> 1024 tiles
    assert _lib.state_layout(10, 10, 1024, 1024).sort_bits == 45
    assert _lib.state_layout(10, 10, 1920, 1080).sort_bits == 45
    offs = [L.g_header, L.g_scan_status, L.g_rec, L.g_tiles_touched, L.g_point_offsets, L.g_cov3D, L.g_clamped]
    assert offs == sorted(offs) and all(o % 256 == 0 for o in offs) and L.geom_bytes > offs[-1]
    b = [L.b_header, L.b_status, L.b_keys_a, L.b_keys_b, L.b_vals_a, L.b_vals_b]
    assert b == sorted(b) and all(o % 256 == 0 for o in b)
    assert L.b_keys_b - L.b_keys_a >= 8 * 180000 and L.binning_bytes - L.b_vals_b >= 4 * 180000
    assert L.img_bytes >= 512 * 512 * 8 + 1024 * 8
    assert L.capacity == 180000


def test_binning_capacity_is_recoverable_from_the_buffer_size():
    """backward must not depend on a side table for the capacity a binning buffer was laid out for.  The
    buffer size is strictly increasing in the capacity (multiples of 64), across the sort's keys-per-tile regime changes
    at 2^17, 3*2^17 and 5*2^17 pairs too, so ghr_binning_capacity inverts it."""
    l = _lib.lib()
    for (P, W, H) in ((50000, 512, 512), (3_000_000, 1920, 1080), (7, 16, 16)):
        prev = -1
        caps = list(range(0, 64 * 40, 64)) + [(1 << 17) - 64, 1 << 17, (1 << 17) + 64, (3 << 17) - 64, 3 << 17,
                                              (3 << 17) + 64, (5 << 17) - 64, 5 << 17, (5 << 17) + 64, 1 << 20, 12_345_664,
                                              (1 << 30) - 64]
        for cap in sorted(set(caps)):
            nbytes = _lib.state_layout(P, cap, W, H).binning_bytes
            assert nbytes > prev, (P, W, H, cap)
            prev = nbytes
            assert l.ghr_binning_capacity(P, W, H, nbytes) == cap
            assert l.ghr_binning_capacity(P, W, H, nbytes + 1) == -1
    # dense sweep over the regime boundaries
    for lo in ((1 << 17) - 64 * 300, (3 << 17) - 64 * 300, (5 << 17) - 64 * 300):
        sizes = [_lib.state_layout(1000, lo + 64 * i, 512, 512).binning_bytes for i in range(600)]
        assert all(b > a for a, b in zip(sizes, sizes[1:]))


def test_bad_arguments_are_rejected_without_touching_the_gpu():
    l = _lib.lib()
    L = _lib.Layout()
    assert l.ghr_state_layout(-1, 0, 16, 16, C.byref(L)) == _lib.GHR_ERR_BAD_ARG
    assert l.ghr_state_layout(1, 1 << 30, 16, 16, C.byref(L)) == _lib.GHR_ERR_BAD_ARG
    assert l.ghr_state_layout(1, 0, 0, 16, C.byref(L)) == _lib.GHR_ERR_BAD_ARG
    p = _lib.RasterParams()
    p.P, p.W, p.H = 4, 16, 16
    assert l.ghr_forward_preprocess(C.byref(p), None, None, 0, None, None) == _lib.GHR_ERR_BAD_ARG  # null inputs
    assert l.ghr_forward_preprocess(None, None, None, 0, None, None) == _lib.GHR_ERR_BAD_ARG
    p.D = 7
    assert l.ghr_forward_render(C.byref(p), None, 0, None, 0, 0, None, 0, None, None, None, None) == _lib.GHR_ERR_BAD_ARG
    assert l.ghr_mark_visible(-1, None, None, None, None, None) == _lib.GHR_ERR_BAD_ARG
    assert l.ghr_mark_visible(3, None, None, None, None, None) == _lib.GHR_ERR_BAD_ARG
    assert l.ghr_sort_pairs_u64(-1, 32, None, None, None, None, None, 0, None, None) == _lib.GHR_ERR_BAD_ARG
    assert l.ghr_sort_pairs_u64(10, 0, None, None, None, None, None, 0, None, None) == _lib.GHR_ERR_BAD_ARG
    assert l.ghr_sort_pairs_u64(10, 40, None, None, None, None, None, 0, None, None) == _lib.GHR_ERR_BUFFER
    assert l.ghr_dist2_knn3(-5, None, None, None, 0, None) == _lib.GHR_ERR_BAD_ARG
    assert l.ghr_dist2_knn3(5, None, None, None, 0, None) == _lib.GHR_ERR_BAD_ARG
    assert l.ghr_sort_workspace_bytes(1000) > 0 and l.ghr_knn_workspace_bytes(6890) > 0


def test_operator_api_surface_and_argument_errors():
    import torch
    import mygauhuman_b200
    from mygauhuman_b200.diff_gaussian_rasterization import (GaussianRasterizationSettings, GaussianRasterizer,
                                                             rasterize_gaussians, _C)
    assert GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")
    for f in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert callable(getattr(_C, f))
    rs = GaussianRasterizationSettings(16, 16, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0,
                                       torch.zeros(3), False, False)
    r = GaussianRasterizer(rs)
    x = torch.zeros(4, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=x, means2D=x, opacities=x[:, :1])
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=x, means2D=x, opacities=x[:, :1], shs=torch.zeros(4, 1, 3), colors_precomp=x)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=x, means2D=x, opacities=x[:, :1], colors_precomp=x)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=x, means2D=x, opacities=x[:, :1], colors_precomp=x, scales=x, rotations=torch.zeros(4, 4),
          cov3D_precomp=torch.zeros(4, 6))
    with pytest.raises(RuntimeError, match="num_points, 3"):
        _C.rasterize_gaussians(torch.zeros(3), torch.zeros(4, 2), x, x[:, :1], x, x, 1.0, x, torch.eye(4), torch.eye(4),
                               1.0, 1.0, 16, 16, x, 0, torch.zeros(3), False, False)
    with pytest.raises(RuntimeError, match="CUDA"):  # no CPU fallback
        _C.rasterize_gaussians(torch.zeros(3), x, x, x[:, :1], x, torch.zeros(4, 4), 1.0, torch.Tensor([]),
                               torch.eye(4), torch.eye(4), 1.0, 1.0, 16, 16, torch.Tensor([]), 0, torch.zeros(3),
                               False, False)
    dgr, sknn = mygauhuman_b200.install()
    import diff_gaussian_rasterization, simple_knn._C
    assert diff_gaussian_rasterization is dgr and callable(simple_knn._C.distCUDA2)


def test_new_entry_points_reject_bad_arguments_without_touching_the_gpu():
    """View-batched forward, optimiser, checkpoint-row and stand-alone operator entry points: argument validation happens
    before any CUDA call, so it can be checked here."""
    l = _lib.lib()
    BAD, BUF, OK = _lib.GHR_ERR_BAD_ARG, _lib.GHR_ERR_BUFFER, _lib.GHR_OK
    # ghr_forward_views / workspace query
    assert l.ghr_views_workspace_bytes(1000, 6, 32, 32, 4000) > 0
    assert l.ghr_views_workspace_bytes(1000, 0, 32, 32, 4000) == 0          # no views
    assert l.ghr_views_workspace_bytes(1 << 30, 8, 32, 32, 4000) == 0       # n_views * P beyond 2^31
    assert l.ghr_views_workspace_bytes(1000, 6, 32, 32, 1 << 30) == 0       # capacity beyond 2^30
    p = _lib.RasterParams()
    p.P, p.W, p.H = 4, 32, 32
    assert l.ghr_forward_views(None, 1, None, None, None, None, 0, 0, None, None, None, None, None) == BAD
    assert l.ghr_forward_views(C.byref(p), -1, None, None, None, None, 0, 0, None, None, None, None, None) == BAD
    assert l.ghr_forward_views(C.byref(p), 0, None, None, None, None, 0, 0, None, None, None, None, None) == OK
    assert l.ghr_forward_views(C.byref(p), 2, None, None, None, None, 0, 16, None, None, None, None, None) == BAD  # no cameras
    # ghr_adam_step
    t = (_lib.AdamGroup * 1)()
    t[0].count, t[0].step = 4, 1
    assert l.ghr_adam_step(1, t, 0.9, 0.999, 1e-15, 0, None) == BAD          # null pointers
    assert l.ghr_adam_step(17, t, 0.9, 0.999, 1e-15, 0, None) == BAD         # > GHR_ADAM_MAX_GROUPS
    assert l.ghr_adam_step(0, None, 0.9, 0.999, 1e-15, 0, None) == OK
    assert l.ghr_adam_step(1, None, 0.9, 0.999, 1e-15, 0, None) == BAD
    t[0].count = 0
    assert l.ghr_adam_step(1, t, 0.9, 0.999, 1e-15, 0, None) == OK           # empty group: nothing to do
    assert l.ghr_adam_step(1, t, 1.0, 0.999, 1e-15, 0, None) == BAD          # beta1 < 1
    assert l.ghr_adam_step(1, t, 0.9, 0.999, -1.0, 0, None) == BAD           # eps >= 0
    # ghr_adam_step_sharded
    g = (_lib.ShardGroup * 1)()
    g[0].offset, g[0].count, g[0].lr, g[0].step = 0, 8, 1e-3, 1
    assert l.ghr_adam_step_sharded(0, 0, None, None, None, None, 0, 8, 1, g, 0.9, 0.999, 1e-15, None) == BAD   # world >= 1
    assert l.ghr_adam_step_sharded(9, 0, None, None, None, None, 0, 8, 1, g, 0.9, 0.999, 1e-15, None) == BAD   # <= 8 peers
    assert l.ghr_adam_step_sharded(2, 2, None, None, None, None, 0, 8, 1, g, 0.9, 0.999, 1e-15, None) == BAD   # rank < world
    assert l.ghr_adam_step_sharded(2, 0, None, None, None, None, 2, 8, 1, g, 0.9, 0.999, 1e-15, None) == BAD   # unaligned shard
    assert l.ghr_adam_step_sharded(2, 0, None, None, None, None, 8, 8, 1, g, 0.9, 0.999, 1e-15, None) == OK    # empty shard
    assert l.ghr_adam_step_sharded(2, 0, None, None, None, None, 0, 8, 1, g, 0.9, 0.999, 1e-15, None) == BAD   # null tables
    assert l.ghr_adam_step_sharded_synced(2, 0, None, None, None, None, 0, 8, 1, g, 0.9, 0.999, 1e-15, None, 1, None) == BAD  # flags
    f = (C.c_void_p * 2)()
    assert l.ghr_adam_step_sharded_synced(2, 0, None, None, None, None, 0, 8, 1, g, 0.9, 0.999, 1e-15, f, 0, None) == BAD     # epoch >= 1
    assert l.ghr_sharded_flag_words() >= 17
    assert l.ghr_enable_peer_access(-1) == BAD
    assert l.ghr_ipc_open(None, None) == BAD and l.ghr_ipc_close(None) == OK
    # ghr_densify_stats
    assert l.ghr_densify_stats(-1, None, None, 3, None, None, None, None) == BAD
    assert l.ghr_densify_stats(5, None, None, 3, None, None, None, None) == BAD
    assert l.ghr_densify_stats(5, None, None, 1, None, None, None, None) == BAD    # rows of >= 2 floats
    assert l.ghr_densify_stats(0, None, None, 3, None, None, None, None) == OK
    # checkpoint rows
    assert l.ghr_ply_row_floats(16) == 66 and l.ghr_ply_row_floats(1) == 21
    assert l.ghr_ply_row_floats(0) == 0 and l.ghr_ply_row_floats(17) == 0
    assert l.ghr_ply_pack_rows(3, 16, *([None] * 11)) == BAD
    assert l.ghr_ply_pack_rows(0, 16, *([None] * 11)) == OK
    assert l.ghr_ply_pack_rows(3, 0, *([None] * 11)) == BAD
    assert l.ghr_ply_unpack_rows(3, 16, None, 65, None, *([None] * 10)) == BAD     # fewer columns than the format holds
    cols = (C.c_int32 * 66)(*range(66))
    cols[5] = 70
    assert l.ghr_ply_unpack_rows(3, 16, None, 66, cols, *([None] * 10)) == BAD     # column index out of the row
    # stand-alone covariance / SH colours
    assert l.ghr_covariance_forward(4, None, None, 1.0, None, None, None) == BAD
    assert l.ghr_covariance_forward(0, None, None, 1.0, None, None, None) == OK
    assert l.ghr_covariance_backward(4, None, None, 1.0, None, None, None, None, None, None) == BAD
    assert l.ghr_sh_colors_forward(4, 4, 16, None, None, None, None, None, None) == BAD   # degree <= 3
    assert l.ghr_sh_colors_forward(4, 3, 9, None, None, None, None, None, None) == BAD    # M >= (deg+1)^2
    assert l.ghr_sh_colors_forward(4, 3, 16, None, None, None, None, None, None) == BAD   # null inputs
    assert l.ghr_sh_colors_backward(0, 3, 16, None, None, None, None, None, None, None, None) == OK
    assert l.ghr_set_shared_gpu(1) == OK and l.ghr_set_shared_gpu(0) == OK
    assert BUF != BAD
