// Host-side internals of libghr: buffer layouts and per-stage launchers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <atomic>
#include "../../include/ghr.h"

namespace ghr {

// ---- measurement hooks (ghr_profile.cu) ------------------------------------
extern std::atomic<uint64_t> g_launches;
inline void count_launch(uint64_t n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }
// RAII stage span: records CUDA events on `s` when profiling is enabled.
class StageTimer {
public:
    StageTimer(int stage, cudaStream_t s);
    ~StageTimer();
    StageTimer(const StageTimer&) = delete;
    StageTimer& operator=(const StageTimer&) = delete;
private:
    int stage_;
    cudaStream_t stream_;
    cudaEvent_t beg_;
};

constexpr int kMaxPasses = 8;
constexpr int kRadixBits = 9;  // 512 bins: ceil(bits/9) passes (5 for the 43..45-bit keys of 512^2..1080p frames)
constexpr int kRadix = 1 << kRadixBits;
constexpr int kSortThreads = 512;
inline int sort_passes_for(int bits) { return (bits + kRadixBits - 1) / kRadixBits; }

// Control block at the start of the binning buffer (device memory).
struct SortHeader {
    uint32_t sorted_sel;              // buffer holding the sorted result: 0 = A, 1 = B
    uint32_t num_passes;
    uint32_t hist_done;               // histogram CTAs finished (the last one writes the plan)
    uint32_t pad;
    uint32_t pass_trivial[kMaxPasses];  // 1: all keys share this digit -> pass skipped
    uint32_t pass_src[kMaxPasses];      // buffer a pass reads from
    uint32_t ticket[kMaxPasses];        // dynamic tile counters (one per pass)
    uint32_t hist[kMaxPasses][kRadix];  // global digit histograms
    uint32_t bin_base[kMaxPasses][kRadix];  // exclusive scan of hist
};

// keys per onesweep tile for a given capacity (must match between layout and launch)
inline int sort_items_per_thread(int64_t capacity)
{
    static const int forced = [] { const char* e = getenv("GHR_SORT_ITEMS"); return e ? atoi(e) : 0; }();  // tuning only
    if (forced == 2 || forced == 4 || forced == 8 || forced == 12 || forced == 16) return forced;
    // Measured on B200 (45-bit keys; us at 8 / 12 / 16 keys per thread): 5.2e5 keys 81 / 85 / 97, 7.9e5 105 / 95 / 105,
    // 1.6e6 155 / 136 / 146, 3.1e6 222 / 214 / 214, 8.4e6 484 / 433 / 449, 1.7e7 900 / 781 / 823.  A tile's fixed work
    // (512 digit scans and look-backs) is spread over more keys; 16 spills at 64 registers.
    return capacity <= (int64_t(1) << 17) ? 2 : capacity <= (int64_t(3) << 17) ? 4 : capacity <= (int64_t(5) << 17) ? 8 : 12;
}
inline int64_t sort_tiles(int64_t capacity)
{
    const int64_t tile = int64_t(kSortThreads) * sort_items_per_thread(capacity);
    return (capacity + tile - 1) / tile;
}

// Bytes of decoupled-look-back words the onesweep passes use at this capacity, and the never-shrinking figure the
// layout reserves (see ghr_state_layout).
inline size_t sort_status_bytes(int passes, int64_t capacity)
{
    return sizeof(uint32_t) * (size_t)passes * (size_t)sort_tiles(capacity) * kRadix;
}
inline size_t sort_status_bytes_monotone(int passes, int64_t capacity)
{
    size_t b = sort_status_bytes(passes, capacity);
    const int64_t thresholds[3] = {int64_t(1) << 17, int64_t(3) << 17, int64_t(5) << 17};  // regime ends of sort_items_per_thread
    for (int64_t t : thresholds)
        if (capacity > t) { const size_t at = sort_status_bytes(passes, t); b = b > at ? b : at; }
    return b;
}

// Blend kernels: staging configuration by problem size (see ghr_blend_fwd.cu).  `capacity` bounds the number of
// (Gaussian, tile) pairs; GHR_BLEND_REGIME=latency|throughput overrides (tuning).
inline bool blend_throughput_regime(int64_t capacity)
{
    static const int forced = [] {
        const char* e = getenv("GHR_BLEND_REGIME");
        return !e ? 0 : (e[0] == 't' ? 2 : 1);
    }();
    if (forced) return forced == 2;
    // Measured on B200 after the compacted-staging rewrite of the 4x4-block kernels (tools/blend_probe.py, forward +
    // backward incl. the per-entry block masks of tile_ranges_kernel): 2.2 M pairs (300 k Gaussians, 1024^2) 522 us vs
    // 560 us for the wide family; 12.3 M pairs (3 M Gaussians, 1080p, forward only) 1 305 us vs 593 us -- the masks cost
    // 16 ellipse tests per list entry, which the wide family does not pay.
    return capacity > (int64_t(1) << 22);
}

// Throughput regime: the sorted VALUE of a pair carries, above the 24-bit Gaussian id, the 8-bit mask of the tile's 8x4-pixel
// blocks whose conservative bounding-box cull the Gaussian passes (computed once per pair by emit_keys instead of once per
// pair AND block warp inside the blend kernel, and only survivors' records are staged).  Needs ids below 2^24.
inline bool wide_masks_packed(int P, int64_t capacity) { return blend_throughput_regime(capacity) && P < (1 << 24); }

// Small frames: emit_keys also accumulates the sort's digit histograms and writes its plan (one launch and one pass over the
// keys fewer: 58 -> 47 us of sort for 9 -> 16 us of emission at the ZJU shape).  At 12 M pairs its five shared-memory
// atomics per pair cost more than the separate histogram kernel (emission 82 -> 174 us for 62 us of sort): not there.
inline bool sort_histogram_in_emit(int64_t capacity)
{
    static const int forced = [] { const char* e = getenv("GHR_HIST_IN_EMIT"); return e ? atoi(e) + 1 : 0; }();  // tuning only
    if (forced) return forced == 2;
    return !blend_throughput_regime(capacity);
}

// Several views share the GPU (views of a batch on several streams): the blend kernels then take half the staging
// shared memory per warp so that two of them are co-resident on an SM.  Process-wide, set by ghr_set_shared_gpu()
// (MultiLaneRenderStep), GHR_SHARED_GPU=0|1 overrides (tuning).
extern std::atomic<int> g_shared_gpu;
inline int blend_shared_gpu()  // 0: one view owns the GPU (256-entry staging rounds); 1: 128-entry rounds (64: measured, no gain)
{
    static const int forced = [] { const char* e = getenv("GHR_SHARED_GPU"); return e ? atoi(e) + 1 : 0; }();
    if (forced) return forced - 1;
    return g_shared_gpu.load(std::memory_order_relaxed);
}

inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Streaming multiprocessors of the CURRENT device (148 on B200), queried once per device and cached (ghr_profile.cu):
// every persistent / grid-capped launch sizes itself as a multiple of this.
int sm_count();
// true the first time it is called for `mask` with the current device current: per-DEVICE one-time set-up (function
// attributes such as the dynamic shared-memory limit belong to the device's context, not to the process)
inline bool first_use_on_current_device(unsigned long long& mask)
{
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess) return true;
    const unsigned long long bit = 1ull << (d & 63);
    if (mask & bit) return false;
    mask |= bit;
    return true;
}

// rasterizer_impl.cu:35-50 getHigherMsb
inline uint32_t higher_msb(uint32_t n)
{
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

// ---- stage launchers (each enqueues on `s`, returns the launch error) -----
cudaError_t launch_preprocess(const ghr_raster_params& p, int32_t* radii, char* geom, const ghr_layout& L,
                              int32_t* n_host, cudaStream_t s);
cudaError_t launch_emit_keys(const ghr_raster_params& p, char* geom, char* binning,
                             const ghr_layout& L, int64_t capacity, cudaStream_t s);
// Sort N (device word) pairs held in A; header/status must be zeroed beforehand.
// histogram_done: the digit histograms and the plan are already in `hdr` (written by emit_keys): no histogram kernel
cudaError_t launch_sort(SortHeader* hdr, uint32_t* status, const uint32_t* n_dev, int64_t capacity, int bits,
                        uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, cudaStream_t s,
                        const uint32_t* squeeze = nullptr, bool histogram_done = false /* device {depth_hi, ~depth_lo} of a depth|tile frame, or NULL */);
cudaError_t launch_tile_ranges(const ghr_raster_params& p, const char* geom, char* binning, char* img,
                               const ghr_layout& L, int64_t capacity, cudaStream_t s);
cudaError_t launch_blend_forward(const ghr_raster_params& p, const char* geom, const char* binning, char* img,
                                 const ghr_layout& L, float* out_color, float* out_depth, float* out_alpha,
                                 cudaStream_t s);
cudaError_t launch_blend_backward(const ghr_raster_params& p, const char* geom, const char* binning, const char* img,
                                  const ghr_layout& L, const float* dL_dpix, const float* dL_ddepth,
                                  const float* dL_dalpha, float* accum, cudaStream_t s);
// throughput-regime variants (ghr_blend_{fwd,bwd}_wide.cu); called by the two launchers above
cudaError_t launch_blend_forward_wide(const ghr_raster_params& p, const char* geom, const char* binning, char* img,
                                      const ghr_layout& L, float* out_color, float* out_depth, float* out_alpha,
                                      cudaStream_t s);
cudaError_t launch_blend_backward_wide(const ghr_raster_params& p, const char* geom, const char* binning, const char* img,
                                       const ghr_layout& L, const float* dL_dpix, const float* dL_ddepth,
                                       const float* dL_dalpha, float* accum, cudaStream_t s);
cudaError_t launch_preprocess_backward(const ghr_raster_params& p, const int32_t* radii, const char* geom,
                                       const ghr_layout& L, const float* accum, float* dL_dmean2D, float* dL_dcolor,
                                       float* dL_dopacity, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                                       float* dL_dscale, float* dL_drot, cudaStream_t s);
// Packed attribute rows (multi-attribute rendering): N attributes = 3N channels, split in two halves of CHS channels (one
// per pixel slot of the blend kernels), each half padded to XH floats; row stride XS = 2*XH floats, a 16-byte multiple.
// Attribute k lives in half k / (N/2) at floats 3 * (k % (N/2)) .. +2 of that half.
template <int N>
struct AttrLayout {
    static_assert(N >= 2 && N % 2 == 0, "attributes are fused in pairs");
    static constexpr int CH = 3 * N, CHS = CH / 2, XH = (CHS + 1) / 2 * 2, XS = 2 * XH;
    static_assert(XS % 4 == 0, "rows are copied 16 bytes at a time");
};
template <int NX>
__host__ __device__ constexpr int attr_row_floats()  // 0 attributes: no row
{
    if constexpr (NX == 0) return 0;
    else return AttrLayout<NX>::XS;
}
inline int attr_stride(int n) { return n == 2 ? AttrLayout<2>::XS : n == 4 ? AttrLayout<4>::XS : n == 6 ? AttrLayout<6>::XS : 0; }
inline int attr_offset(int n, int k) { return (k / (n / 2)) * (attr_stride(n) / 2) + 3 * (k % (n / 2)); }
struct AttrPtrs { const float* src[6]; float* dst[6]; };
cudaError_t launch_pack_attributes(int P, int n, const AttrPtrs& ptrs, float* packed, cudaStream_t s);
cudaError_t launch_unpack_attributes(int P, int n, const float* packed, const AttrPtrs& ptrs, cudaStream_t s);
cudaError_t launch_blend_forward_multi(const ghr_raster_params& p, const char* geom, const char* binning, char* img,
                                       const ghr_layout& L, float* out_color, float* out_depth, float* out_alpha,
                                       int n_attr, const float* xcol, float* out_x, cudaStream_t s);
cudaError_t launch_blend_backward_multi(const ghr_raster_params& p, const char* geom, const char* binning, const char* img,
                                        const ghr_layout& L, const float* dL_dpix, const float* dL_ddepth,
                                        const float* dL_dalpha, float* accum, int n_attr, const float* xcol,
                                        const float* xdpix, float* xaccum, cudaStream_t s);
// rec[idx].rgb = colors[idx] for every Gaussian (multi-attribute rendering)
cudaError_t launch_recolor(int P, const float* colors, char* geom, const ghr_layout& L, cudaStream_t s);
cudaError_t launch_mark_visible(int P, const float* means3D, const float* V, uint8_t* present, cudaStream_t s);
// ---- LBS (ghr_lbs_pose.cu, ghr_lbs_deform.cu, ghr_knn.cu) -------------------
struct DeformOut {
    float *smpl_pts, *world_pts, *bweights, *transforms, *translation, *world_normals, *cov3D, *colors;
    uint8_t* clamped;
};
struct DeformGrads {
    const uint8_t* clamped;
    const float *g_world, *g_transforms, *g_normals, *g_cov, *g_colors;  // incoming (may be NULL)
    float *d_xyz, *d_normals, *d_offsets, *d_scales, *d_rot, *d_shs;     // per-Gaussian outputs
    float *d_A_target, *d_vertex_offsets;                                 // accumulated (may be NULL)
};
cudaError_t launch_lbs_prepare(int V, int NB, const float* vt, const float* sd, const float* Jreg, float* jt, float* jsd,
                               cudaStream_t s);
cudaError_t launch_lbs_pose(const ghr_lbs_model& m, const float* poses, const float* shapes, const float* correct_Rs,
                            const float* R, const float* Th, float* out, cudaStream_t s);
cudaError_t launch_lbs_vertex_offsets(const ghr_lbs_model& m, const float* pose_big, const float* pose_tgt,
                                      const float* shapes, float* off, cudaStream_t s);
cudaError_t launch_lbs_pose_views(const ghr_lbs_model& m, int n, const ghr_pose_view* views, const float* pose_big,
                                  cudaStream_t s);
cudaError_t launch_lbs_pose_backward(const ghr_lbs_model& m, const float* pose_tgt, const float* dA, const float* doff,
                                     float* dC, float* dpf_scratch, cudaStream_t s);
cudaError_t launch_smpl_lbs_vertices(const ghr_lbs_model& m, const float* vt, const float* pose, const float* offsets,
                                     float* verts, float* T_out, cudaStream_t s);
cudaError_t launch_lbs_deform_forward(const ghr_deform_params& p, const DeformOut& o, cudaStream_t s);
cudaError_t launch_lbs_deform_backward(const ghr_deform_params& p, const DeformGrads& g, cudaStream_t s);
cudaError_t launch_pose_refiner(int width, int n_layers, const float* const* w, const float* const* b, const float* pose,
                                float* Rs, cudaStream_t s);
// LBS-weight offset network on the tensor cores (ghr_mlp.cu)
size_t lbs_mlp_blob_bytes();
cudaError_t launch_lbs_mlp_pack(const float* const* weights, const float* const* biases, void* blob, cudaStream_t s);
cudaError_t launch_lbs_mlp_forward(int P, const float* xyz, const void* blob, float* logits, cudaStream_t s);
size_t lbs_mlp_blob_f16_bytes();
cudaError_t launch_lbs_mlp_pack_f16(const float* const* weights, void* blob_f16, cudaStream_t s);
cudaError_t launch_lbs_mlp_forward_f16(int P, const float* xyz, const void* blob_f16, const float* biases, float* logits,
                                       cudaStream_t s);
// fused LBS + projection (ghr_fused.cu)
cudaError_t launch_deform_project(const ghr_deform_params& dp, const ghr_raster_params& p, int32_t* radii, char* geom,
                                  const ghr_layout& L, float* world_pts, float* world_normals, float* cov3D, float* colors,
                                  float* opacity, int32_t* n_host, cudaStream_t s);
cudaError_t launch_project_deform_backward(const ghr_deform_params& dp, const ghr_raster_params& p, const int32_t* radii,
                                           const char* geom, const ghr_layout& L, const float* accum,
                                           const DeformGrads& g, float* dL_dmean2D, float* max_radii2D,
                                           float* xyz_gradient_accum, float* denom, cudaStream_t s);
size_t vertex_grid_workspace_bytes(int V);
cudaError_t launch_vertex_grid_build(int V, const float* verts, char* ws, cudaStream_t s);
cudaError_t launch_nearest_vertex(int P, const float* query, int V, const char* ws, int32_t* vid, cudaStream_t s);
size_t knn_workspace_bytes(int P);
cudaError_t launch_knn(int P, const float* points, float* out, char* ws, size_t ws_bytes, cudaStream_t s);

}  // namespace ghr
