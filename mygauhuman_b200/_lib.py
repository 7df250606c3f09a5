"""ctypes binding of libghr.so (the C ABI declared in include/ghr.h).

There is no CPU fallback: if the CUDA library is missing, import of the
operators fails loudly (build it with ``python -c "import __graft_entry__ as g; g.build()"``
or ``mygauhuman_b200/csrc/build.sh``).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GHR_LIB_PATH") or os.path.join(_HERE, "csrc", "libghr.so")  # override: tuning variants only

GHR_OK = 0
GHR_ERR_BAD_ARG, GHR_ERR_BUFFER, GHR_ERR_CUDA, GHR_ERR_CAPACITY, GHR_ERR_PREFILTER = -1, -2, -3, -4, -5

c_f32p = C.c_void_p  # device pointers travel as integers


class RasterParams(C.Structure):
    """ghr_raster_params (include/ghr.h)."""
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
        ("scale_modifier", C.c_float), ("tan_fovx", C.c_float), ("tan_fovy", C.c_float),
        ("prefiltered", C.c_int32), ("debug", C.c_int32),
        ("background", c_f32p), ("means3D", c_f32p), ("shs", c_f32p), ("colors_precomp", c_f32p),
        ("opacities", c_f32p), ("scales", c_f32p), ("rotations", c_f32p), ("cov3D_precomp", c_f32p),
        ("viewmatrix", c_f32p), ("projmatrix", c_f32p), ("campos", c_f32p),
    ]


class Layout(C.Structure):
    """ghr_layout (include/ghr.h)."""
    _fields_ = [
        ("geom_bytes", C.c_size_t), ("g_header", C.c_size_t), ("g_rec", C.c_size_t),
        ("g_tiles_touched", C.c_size_t), ("g_point_offsets", C.c_size_t), ("g_cov3D", C.c_size_t),
        ("g_clamped", C.c_size_t), ("g_scan_status", C.c_size_t),
        ("img_bytes", C.c_size_t), ("i_final_T", C.c_size_t), ("i_n_contrib", C.c_size_t), ("i_ranges", C.c_size_t),
        ("binning_bytes", C.c_size_t), ("b_header", C.c_size_t), ("b_keys_a", C.c_size_t), ("b_keys_b", C.c_size_t),
        ("b_vals_a", C.c_size_t), ("b_vals_b", C.c_size_t), ("b_status", C.c_size_t),
        ("capacity", C.c_int64), ("sort_bits", C.c_int32), ("sort_passes", C.c_int32), ("i_order", C.c_size_t),
    ]


class LbsModel(C.Structure):
    """ghr_lbs_model (include/ghr.h)."""
    _fields_ = [("V", C.c_int32), ("NB", C.c_int32), ("weights", c_f32p), ("posedirs", c_f32p), ("shapedirs", c_f32p),
                ("joint_template", c_f32p), ("joint_shapedirs", c_f32p), ("parents", C.c_void_p)]


class PoseView(C.Structure):
    """ghr_pose_view (include/ghr.h)."""
    _fields_ = [("poses", c_f32p), ("shapes", c_f32p), ("correct_Rs", c_f32p), ("R", c_f32p), ("Th", c_f32p),
                ("pose_out", c_f32p), ("offsets_out", c_f32p)]


POSE_VIEWS_MAX = 8


class DeformParams(C.Structure):
    """ghr_deform_params (include/ghr.h)."""
    _fields_ = [("P", C.c_int32), ("V", C.c_int32), ("sh_degree", C.c_int32), ("M", C.c_int32),
                ("scale_modifier", C.c_float), ("xyz", c_f32p), ("normals", c_f32p), ("vert_ids", C.c_void_p),
                ("weights", c_f32p), ("lbs_offsets", c_f32p), ("lbs_stride_p", C.c_int64), ("lbs_stride_j", C.c_int64),
                ("vertex_offsets", c_f32p), ("pose_big", c_f32p), ("pose_target", c_f32p), ("scales", c_f32p),
                ("rotations", c_f32p), ("shs", c_f32p), ("campos", c_f32p), ("scales_are_log", C.c_int32),
                ("opacity_logits", c_f32p), ("opacity_out", c_f32p), ("dL_dopacity", c_f32p),
                ("dL_dopacity_logits", c_f32p), ("accumulate_grads", C.c_int32)]


class AdamGroup(C.Structure):
    """ghr_adam_group (include/ghr.h)."""
    _fields_ = [("param", c_f32p), ("grad", c_f32p), ("exp_avg", c_f32p), ("exp_avg_sq", c_f32p), ("count", C.c_int64),
                ("lr", C.c_double), ("step", C.c_int32), ("pad", C.c_int32)]


class ShardGroup(C.Structure):
    """ghr_shard_group (include/ghr.h)."""
    _fields_ = [("offset", C.c_int64), ("count", C.c_int64), ("lr", C.c_double), ("step", C.c_int32), ("pad", C.c_int32)]


ADAM_MAX_GROUPS = 16
MAX_PEERS = 8
POSE_FLOATS = 1536
POSE_A, POSE_ROT, POSE_PF, POSE_JOINTS, POSE_RGINV, POSE_RG, POSE_TH, POSE_G, POSE_ROD = 0, 288, 504, 720, 800, 816, 832, 848, 1136

# name -> (restype, argtypes); kept in one table so tests can check every
# symbol include/ghr.h declares is exported.
SIGNATURES = {
    "ghr_version": (C.c_int, []),
    "ghr_error_string": (C.c_char_p, [C.c_int]),
    "ghr_state_layout": (C.c_int, [C.c_int32, C.c_int64, C.c_int32, C.c_int32, C.POINTER(Layout)]),
    "ghr_binning_capacity": (C.c_int64, [C.c_int32, C.c_int32, C.c_int32, C.c_size_t]),
    "ghr_forward_preprocess": (C.c_int, [C.POINTER(RasterParams), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
                                         C.c_void_p]),
    "ghr_forward_render": (C.c_int, [C.POINTER(RasterParams), C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                     C.c_int64, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p]),
    "ghr_backward": (C.c_int, [C.POINTER(RasterParams), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                               C.c_int64, C.c_void_p, C.c_size_t] + [C.c_void_p] * 13),
    "ghr_forward_reblend": (C.c_int, [C.POINTER(RasterParams), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                      C.c_int64, C.c_void_p, C.c_size_t] + [C.c_void_p] * 4),
    "ghr_backward_blend": (C.c_int, [C.POINTER(RasterParams), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                     C.c_int64, C.c_void_p, C.c_size_t] + [C.c_void_p] * 6),
    "ghr_backward_finish": (C.c_int, [C.POINTER(RasterParams), C.c_void_p, C.c_void_p, C.c_size_t, C.c_int64]
                            + [C.c_void_p] * 10),
    "ghr_attribute_stride": (C.c_int, [C.c_int]),
    "ghr_attributes_fused": (C.c_int, [C.c_int64]),
    "ghr_pack_attributes": (C.c_int, [C.c_int32, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ghr_unpack_attribute_grads": (C.c_int, [C.c_int32, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ghr_forward_render_multi": (C.c_int, [C.POINTER(RasterParams), C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                           C.c_int64, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ghr_backward_multi": (C.c_int, [C.POINTER(RasterParams), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                     C.c_int64, C.c_void_p, C.c_size_t] + [C.c_void_p] * 3 + [C.c_int]
                           + [C.c_void_p] * 13),
    "ghr_point_list_id_mask": (C.c_uint32, [C.c_int32, C.c_int64]),
    "ghr_mark_visible": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ghr_sort_workspace_bytes": (C.c_size_t, [C.c_int64]),
    "ghr_sort_pairs_u64": (C.c_int, [C.c_int64, C.c_int32] + [C.c_void_p] * 5 + [C.c_size_t, C.c_void_p, C.c_void_p]),
    "ghr_sort_static_max_keys": (C.c_int64, []),
    "ghr_sort_pairs_u64_static": (C.c_int, [C.c_int64, C.c_int32] + [C.c_void_p] * 6),
    "ghr_knn_workspace_bytes": (C.c_size_t, [C.c_int32]),
    "ghr_dist2_knn3": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "ghr_lbs_prepare": (C.c_int, [C.c_int32, C.c_int32] + [C.c_void_p] * 6),
    "ghr_lbs_pose": (C.c_int, [C.POINTER(LbsModel)] + [C.c_void_p] * 7),
    "ghr_lbs_vertex_offsets": (C.c_int, [C.POINTER(LbsModel)] + [C.c_void_p] * 5),
    "ghr_lbs_pose_views": (C.c_int, [C.POINTER(LbsModel), C.c_int32, C.POINTER(PoseView), C.c_void_p, C.c_void_p]),
    "ghr_vertex_grid_workspace_bytes": (C.c_size_t, [C.c_int32]),
    "ghr_vertex_grid_build": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "ghr_nearest_vertex": (C.c_int, [C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ghr_lbs_deform_forward": (C.c_int, [C.POINTER(DeformParams)] + [C.c_void_p] * 10),
    "ghr_lbs_deform_backward": (C.c_int, [C.POINTER(DeformParams)] + [C.c_void_p] * 15),
    "ghr_lbs_pose_backward": (C.c_int, [C.POINTER(LbsModel)] + [C.c_void_p] * 6),
    "ghr_pose_refiner_forward": (C.c_int, [C.c_int32, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_void_p,
                                           C.c_void_p, C.c_void_p]),
    "ghr_lbs_mlp_blob_bytes": (C.c_size_t, []),
    "ghr_lbs_mlp_pack": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p]),
    "ghr_lbs_mlp_forward": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ghr_forward_deform_preprocess": (C.c_int, [C.POINTER(DeformParams), C.POINTER(RasterParams), C.c_void_p, C.c_void_p,
                                                C.c_size_t] + [C.c_void_p] * 7),
    "ghr_forward_render_deformed": (C.c_int, [C.POINTER(RasterParams), C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                              C.c_int64, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_void_p]),
    "ghr_backward_deformed": (C.c_int, [C.POINTER(DeformParams), C.POINTER(RasterParams), C.c_void_p, C.c_void_p,
                                        C.c_size_t, C.c_void_p, C.c_size_t, C.c_int64, C.c_void_p, C.c_size_t]
                              + [C.c_void_p] * 16),
    "ghr_smpl_lbs_vertices": (C.c_int, [C.POINTER(LbsModel)] + [C.c_void_p] * 6),
    "ghr_covariance_forward": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ghr_covariance_backward": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_float] + [C.c_void_p] * 6),
    "ghr_sh_colors_forward": (C.c_int, [C.c_int32, C.c_int32, C.c_int32] + [C.c_void_p] * 6),
    "ghr_sh_colors_backward": (C.c_int, [C.c_int32, C.c_int32, C.c_int32] + [C.c_void_p] * 8),
    "ghr_views_workspace_bytes": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int64]),
    "ghr_forward_views": (C.c_int, [C.POINTER(RasterParams), C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_size_t, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ghr_cubemap_sample_nearest": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                            C.c_void_p]),
    "ghr_bake_occlusion": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_void_p]),
    "ghr_adam_step": (C.c_int, [C.c_int, C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_int, C.c_void_p]),
    "ghr_adam_step_sharded": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                        C.c_int, C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_void_p]),
    "ghr_adam_step_sharded_synced": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                               C.c_int64, C.c_int, C.c_void_p, C.c_double, C.c_double, C.c_double,
                                               C.c_void_p, C.c_uint32, C.c_void_p]),
    "ghr_sharded_flag_words": (C.c_int, []),
    "ghr_sharded_error_word": (C.c_int, []),
    "ghr_sharded_abort": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p]),
    "ghr_enable_peer_access": (C.c_int, [C.c_int]),
    "ghr_ipc_open": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "ghr_ipc_close": (C.c_int, [C.c_void_p]),
    "ghr_densify_stats": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "ghr_ply_row_floats": (C.c_int, [C.c_int32]),
    "ghr_ply_pack_rows": (C.c_int, [C.c_int32, C.c_int32] + [C.c_void_p] * 11),
    "ghr_ply_unpack_rows": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p] + [C.c_void_p] * 10),
    "ghr_launch_count": (C.c_uint64, []),
    "ghr_set_shared_gpu": (C.c_int, [C.c_int]),
    "ghr_shared_gpu": (C.c_int, []),
    "ghr_profile_enable": (C.c_int, [C.c_int]),
    "ghr_profile_read": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
}

STAGES = ("preprocess", "emit_keys", "sort", "tile_ranges", "blend_fwd", "blend_bwd", "preprocess_bwd", "knn",
          "lbs_joints", "lbs_fwd", "lbs_bwd", "misc", "optimizer", "deform_project", "project_deform_bwd")

_lib = None


class GhrError(RuntimeError):
    def __init__(self, status, where):
        self.status = status
        super().__init__(f"libghr {where}: {error_string(status)} (status {status})")


def lib():
    """Load libghr.so once; raise (never fall back) if it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: the CUDA library has not been built "
                "(run mygauhuman_b200/csrc/build.sh). There is no CPU fallback.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def error_string(status):
    return lib().ghr_error_string(int(status)).decode()


def check(status, where):
    if status != GHR_OK:
        raise GhrError(status, where)


def raw_stream(device_index):
    """cudaStream_t of torch's current stream on `device_index` as an int (fast path: no Stream object)."""
    import torch
    return torch._C._cuda_getCurrentRawStream(device_index)


class on_device:
    """`with on_device(dev):` = torch.cuda.device(dev) but free when dev is already current (the common case)."""
    __slots__ = ("idx", "ctx")

    def __init__(self, dev):
        # a device without an index ("cuda") means the CURRENT device, as everywhere in torch -- not device 0
        if dev.index is not None:
            self.idx = dev.index
        else:
            import torch
            self.idx = torch._C._cuda_getDevice()
        self.ctx = None

    def __enter__(self):
        import torch
        if torch._C._cuda_getDevice() != self.idx:
            self.ctx = torch.cuda.device(self.idx)
            self.ctx.__enter__()
        return self

    def __exit__(self, *a):
        if self.ctx is not None:
            self.ctx.__exit__(*a)
            self.ctx = None
        return False


def launch_count():
    """Kernels launched by libghr since it was loaded."""
    return int(lib().ghr_launch_count())


def profile_enable(on=True):
    check(lib().ghr_profile_enable(int(bool(on))), "ghr_profile_enable")


def profile_read():
    """{stage: (device ms summed over spans, span count)} since the last read; synchronises."""
    ms = (C.c_double * len(STAGES))()
    cnt = (C.c_int64 * len(STAGES))()
    check(lib().ghr_profile_read(ms, cnt), "ghr_profile_read")
    return {name: (ms[i], cnt[i]) for i, name in enumerate(STAGES)}


def state_layout(P, capacity, W, H):
    L = Layout()
    check(lib().ghr_state_layout(int(P), int(capacity), int(W), int(H), C.byref(L)), "ghr_state_layout")
    return L
