// extern "C" entry points of libghr.so (declared in include/ghr.h).
// Host side only: argument validation, buffer layout, stage sequencing.
#include <stdio.h>
#include <string.h>
#include "ghr_common.cuh"
#include "ghr_internal.h"

using namespace ghr;

namespace {

inline bool misaligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) != 0; }

int check_cuda(cudaError_t e, const char* what)
{
    if (e == cudaSuccess) return GHR_OK;
    fprintf(stderr, "[libghr] CUDA error in %s: %s\n", what, cudaGetErrorString(e));
    return GHR_ERR_CUDA;
}

// debug=1 mirrors the reference's CHECK_CUDA(.., debug): sync + check after each stage
int stage_done(const ghr_raster_params* prm, cudaError_t launch_err, cudaStream_t s, const char* what)
{
    int r = check_cuda(launch_err, what);
    if (r != GHR_OK) return r;
    if (prm && prm->debug) return check_cuda(cudaStreamSynchronize(s), what);
    return GHR_OK;
}

int validate(const ghr_raster_params* p)
{
    if (!p) return GHR_ERR_BAD_ARG;
    if (p->P < 0 || p->W <= 0 || p->H <= 0 || p->M < 0 || p->D < 0 || p->D > 3) return GHR_ERR_BAD_ARG;
    if (p->P == 0) return GHR_OK;
    if (!p->means3D || !p->opacities || !p->viewmatrix || !p->projmatrix || !p->background) return GHR_ERR_BAD_ARG;
    if (!p->shs && !p->colors_precomp) return GHR_ERR_BAD_ARG;
    if (p->shs && (p->M <= 0 || !p->campos || (p->D + 1) * (p->D + 1) > p->M || p->M > 16)) return GHR_ERR_BAD_ARG;
    if (!p->cov3D_precomp && (!p->scales || !p->rotations)) return GHR_ERR_BAD_ARG;
    if (misaligned(p->means3D, 4) || misaligned(p->cov3D_precomp, 8) || misaligned(p->rotations, 16) ||
        misaligned(p->shs, 16))
        return GHR_ERR_BAD_ARG;
    return GHR_OK;
}

// the camera / image half of ghr_raster_params: what the stages after the projection read
int validate_view(const ghr_raster_params* p)
{
    if (!p) return GHR_ERR_BAD_ARG;
    if (p->P < 0 || p->W <= 0 || p->H <= 0) return GHR_ERR_BAD_ARG;
    if (p->P == 0) return GHR_OK;
    if (!p->viewmatrix || !p->projmatrix || !p->background || !p->campos) return GHR_ERR_BAD_ARG;
    return GHR_OK;
}

}  // namespace

extern "C" {

int ghr_version(void) { return GHR_VERSION; }

const char* ghr_error_string(int status)
{
    switch (status) {
        case GHR_OK: return "ok";
        case GHR_ERR_BAD_ARG: return "bad argument (null/misaligned pointer or bad size)";
        case GHR_ERR_BUFFER: return "state/workspace buffer too small";
        case GHR_ERR_CUDA: return "CUDA runtime or kernel error";
        case GHR_ERR_CAPACITY: return "num_rendered exceeded the binning capacity";
        case GHR_ERR_PREFILTER: return "Point is filtered although prefiltered is set";
        default: return "unknown status";
    }
}

int ghr_state_layout(int32_t P, int64_t capacity, int32_t W, int32_t H, ghr_layout* L)
{
    if (!L || P < 0 || capacity < 0 || W <= 0 || H <= 0 || capacity >= (int64_t(1) << 30)) return GHR_ERR_BAD_ARG;
    memset(L, 0, sizeof(*L));
    const size_t Ps = (size_t)P;
    size_t o = 0;
    L->g_header = o; o += align_up(sizeof(ghr_geom_header));
    L->g_scan_status = o; o += align_up(sizeof(uint64_t) * ((Ps + GHR_BLOCK - 1) / GHR_BLOCK + 1));
    L->g_rec = o; o += align_up(sizeof(float4) * 3 * Ps);
    L->g_tiles_touched = o; o += align_up(sizeof(uint32_t) * Ps);
    L->g_point_offsets = o; o += align_up(sizeof(uint32_t) * Ps);
    L->g_cov3D = o; o += align_up(sizeof(float) * 6 * Ps);
    L->g_clamped = o; o += align_up(Ps);
    L->geom_bytes = o;

    const size_t X = (size_t)W * H;
    const size_t tiles = (size_t)((W + GHR_TILE - 1) / GHR_TILE) * ((H + GHR_TILE - 1) / GHR_TILE);
    o = 0;
    L->i_final_T = o; o += align_up(sizeof(float) * X);
    L->i_n_contrib = o; o += align_up(sizeof(uint32_t) * X);
    L->i_ranges = o; o += align_up(sizeof(uint2) * tiles);
    L->i_order = o; o += align_up(sizeof(uint32_t) * tiles);
    L->img_bytes = o;

    L->capacity = capacity;
    L->sort_bits = 32 + (int32_t)higher_msb((uint32_t)tiles);
    L->sort_passes = sort_passes_for(L->sort_bits);
    const size_t cap = (size_t)capacity;
    o = 0;
    L->b_header = o; o += align_up(sizeof(SortHeader));
    // The look-back area is never smaller than at a smaller capacity (the keys-per-tile regime changes at two capacity
    // thresholds, where the tile count halves): binning_bytes is then strictly increasing in the capacity (multiples of
    // 64), which is what lets ghr_binning_capacity recover the capacity from the buffer size alone.
    L->b_status = o; o += align_up(sort_status_bytes_monotone(L->sort_passes, capacity));
    L->b_keys_a = o; o += align_up(sizeof(uint64_t) * cap);
    L->b_keys_b = o; o += align_up(sizeof(uint64_t) * cap);
    L->b_vals_a = o; o += align_up(sizeof(uint32_t) * cap);
    L->b_vals_b = o; o += align_up(sizeof(uint32_t) * cap);
    L->binning_bytes = o;
    return GHR_OK;
}

int64_t ghr_binning_capacity(int32_t P, int32_t W, int32_t H, size_t binning_bytes)
{
    // binning_bytes is strictly increasing over capacities that are multiples of 64 (ghr_state_layout): bisect
    ghr_layout L;
    int64_t lo = 0, hi = (int64_t(1) << 30) / 64 - 1;
    while (lo < hi) {
        const int64_t mid = (lo + hi) / 2;
        if (ghr_state_layout(P, mid * 64, W, H, &L) != GHR_OK) return -1;
        if (L.binning_bytes < binning_bytes) lo = mid + 1; else hi = mid;
    }
    if (ghr_state_layout(P, lo * 64, W, H, &L) != GHR_OK || L.binning_bytes != binning_bytes) return -1;
    return lo * 64;
}

int ghr_forward_preprocess(const ghr_raster_params* prm, int32_t* radii, void* geom, size_t geom_bytes,
                           int32_t* num_rendered_host, void* stream)
{
    int r = validate(prm);
    if (r != GHR_OK) return r;
    cudaStream_t s = (cudaStream_t)stream;
    ghr_layout L;
    if (ghr_state_layout(prm->P, 0, prm->W, prm->H, &L) != GHR_OK) return GHR_ERR_BAD_ARG;
    if (!geom || geom_bytes < L.geom_bytes || misaligned(geom, 256)) return GHR_ERR_BUFFER;
    if (prm->P == 0) {
        r = check_cuda(cudaMemsetAsync(geom, 0, sizeof(ghr_geom_header), s), "preprocess(P=0)");
        if (num_rendered_host) *num_rendered_host = 0;
        return r;
    }
    if (!radii) return GHR_ERR_BAD_ARG;
    return stage_done(prm, launch_preprocess(*prm, radii, (char*)geom, L, num_rendered_host, s), s, "preprocess");
}

static int multi_layout(const ghr_raster_params* prm, int64_t capacity, const void* geom, size_t geom_bytes,
                        const void* binning, size_t binning_bytes, const void* img, size_t img_bytes, ghr_layout* L);

static int forward_render_impl(const ghr_raster_params* prm, void* geom, size_t geom_bytes, void* binning,
                               size_t binning_bytes, int64_t capacity, void* img, size_t img_bytes, float* out_color,
                               float* out_depth, float* out_alpha, int n_attr, const float* packed, float* out_images,
                               void* stream, bool deformed = false)
{
    int r = deformed ? validate_view(prm) : validate(prm);
    if (r != GHR_OK) return r;
    if (!out_color || !out_depth || !out_alpha) return GHR_ERR_BAD_ARG;
    if (n_attr && (attr_stride(n_attr) == 0 || blend_throughput_regime(capacity) || !packed || !out_images ||
                   misaligned(packed, 16)))
        return GHR_ERR_BAD_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    ghr_layout L;
    if (ghr_state_layout(prm->P, capacity, prm->W, prm->H, &L) != GHR_OK) return GHR_ERR_BAD_ARG;
    if (!geom || geom_bytes < L.geom_bytes || !img || img_bytes < L.img_bytes || misaligned(img, 256))
        return GHR_ERR_BUFFER;
    if (!binning || binning_bytes < L.binning_bytes || misaligned(binning, 256)) return GHR_ERR_BUFFER;
    char* g = (char*)geom; char* b = (char*)binning; char* im = (char*)img;
    if (prm->P == 0) {
        // rasterize_points.cu:83: with no Gaussians the reference returns its zero-filled outputs untouched
        const size_t X = (size_t)prm->W * prm->H;
        r = check_cuda(cudaMemsetAsync(out_color, 0, sizeof(float) * 3 * X, s), "P=0 memset");
        if (r == GHR_OK) r = check_cuda(cudaMemsetAsync(out_depth, 0, sizeof(float) * X, s), "P=0 memset");
        if (r == GHR_OK) r = check_cuda(cudaMemsetAsync(out_alpha, 0, sizeof(float) * X, s), "P=0 memset");
        if (r == GHR_OK && n_attr) r = check_cuda(cudaMemsetAsync(out_images, 0, sizeof(float) * 3 * n_attr * X, s), "P=0 memset");
        return r;
    }
    // one memset clears the sort header (tickets, histograms) and every pass's look-back words
    r = check_cuda(cudaMemsetAsync(b + L.b_header, 0, (L.b_status - L.b_header) + sort_status_bytes(L.sort_passes, capacity), s),
                   "binning memset");
    if (r != GHR_OK) return r;
    const uint32_t* n_dev = reinterpret_cast<const uint32_t*>(g + L.g_header);
    if (capacity > 0 && prm->P > 0) {
        r = stage_done(prm, launch_emit_keys(*prm, g, b, L, capacity, s), s, "emit_keys");
        if (r != GHR_OK) return r;
        r = stage_done(prm,
                       launch_sort(reinterpret_cast<SortHeader*>(b + L.b_header), reinterpret_cast<uint32_t*>(b + L.b_status),
                                   n_dev, capacity, L.sort_bits, reinterpret_cast<uint64_t*>(b + L.b_keys_a),
                                   reinterpret_cast<uint64_t*>(b + L.b_keys_b), reinterpret_cast<uint32_t*>(b + L.b_vals_a),
                                   reinterpret_cast<uint32_t*>(b + L.b_vals_b), s,
                                   &reinterpret_cast<const ghr_geom_header*>(g + L.g_header)->depth_hi,
                                   /*histogram_done=*/sort_histogram_in_emit(capacity)),
                       s, "sort");
        if (r != GHR_OK) return r;
    }
    r = stage_done(prm, launch_tile_ranges(*prm, g, b, im, L, capacity, s), s, "tile_ranges");
    if (r != GHR_OK) return r;
    if (n_attr)
        return stage_done(prm,
                          launch_blend_forward_multi(*prm, g, b, im, L, out_color, out_depth, out_alpha, n_attr, packed,
                                                     out_images, s),
                          s, "blend_forward_multi");
    return stage_done(prm, launch_blend_forward(*prm, g, b, im, L, out_color, out_depth, out_alpha, s), s, "blend_forward");
}

int ghr_forward_render(const ghr_raster_params* prm, void* geom, size_t geom_bytes, void* binning,
                       size_t binning_bytes, int64_t capacity, void* img, size_t img_bytes, float* out_color,
                       float* out_depth, float* out_alpha, void* stream)
{
    return forward_render_impl(prm, geom, geom_bytes, binning, binning_bytes, capacity, img, img_bytes, out_color,
                               out_depth, out_alpha, 0, nullptr, nullptr, stream);
}

int ghr_forward_render_multi(const ghr_raster_params* prm, void* geom, size_t geom_bytes, void* binning,
                             size_t binning_bytes, int64_t capacity, void* img, size_t img_bytes, float* out_color,
                             float* out_depth, float* out_alpha, int n_attr, const float* packed, float* out_images,
                             void* stream)
{
    if (n_attr <= 0) return GHR_ERR_BAD_ARG;
    return forward_render_impl(prm, geom, geom_bytes, binning, binning_bytes, capacity, img, img_bytes, out_color,
                               out_depth, out_alpha, n_attr, packed, out_images, stream);
}

uint32_t ghr_point_list_id_mask(int32_t P, int64_t capacity) { return wide_masks_packed(P, capacity) ? 0x00FFFFFFu : 0xFFFFFFFFu; }

int ghr_attribute_stride(int n_attr) { return attr_stride(n_attr); }
int ghr_attributes_fused(int64_t capacity) { return blend_throughput_regime(capacity) ? 0 : 1; }

static int attr_ptrs(int n_attr, const float* const* src, float* const* dst, AttrPtrs* out)
{
    if (attr_stride(n_attr) == 0) return GHR_ERR_BAD_ARG;
    for (int k = 0; k < 6; k++) {
        out->src[k] = (src && k < n_attr) ? src[k] : nullptr;
        out->dst[k] = (dst && k < n_attr) ? dst[k] : nullptr;
        if (src && k < n_attr && !src[k]) return GHR_ERR_BAD_ARG;
    }
    return GHR_OK;
}

int ghr_pack_attributes(int32_t P, int n_attr, const float* const* colors, float* packed, void* stream)
{
    AttrPtrs ptrs;
    if (P < 0 || !colors || !packed || misaligned(packed, 16) || attr_ptrs(n_attr, colors, nullptr, &ptrs) != GHR_OK)
        return GHR_ERR_BAD_ARG;
    return check_cuda(launch_pack_attributes(P, n_attr, ptrs, packed, (cudaStream_t)stream), "pack_attributes");
}

int ghr_unpack_attribute_grads(int32_t P, int n_attr, const float* packed_grad, float* const* dL_dcolors, void* stream)
{
    AttrPtrs ptrs;
    if (P < 0 || !dL_dcolors || !packed_grad || attr_ptrs(n_attr, nullptr, dL_dcolors, &ptrs) != GHR_OK)
        return GHR_ERR_BAD_ARG;
    return check_cuda(launch_unpack_attributes(P, n_attr, packed_grad, ptrs, (cudaStream_t)stream), "unpack_attributes");
}

int ghr_backward_multi(const ghr_raster_params* prm, const int32_t* radii, const void* geom, size_t geom_bytes,
                       const void* binning, size_t binning_bytes, int64_t capacity, const void* img, size_t img_bytes,
                       const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha, int n_attr,
                       const float* packed, const float* dL_dout_images, float* packed_grad, float* accum,
                       float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                       float* dL_dsh, float* dL_dscales, float* dL_drotations, void* stream)
{
    int r = validate(prm);
    if (r != GHR_OK) return r;
    if (prm->P == 0) return GHR_OK;
    if (!radii || !dL_dout_color || !dL_dout_depth || !dL_dout_alpha || !accum || !dL_dmeans2D || !dL_dcolors ||
        !dL_dopacity || !dL_dmeans3D || !dL_dcov3D || misaligned(accum, 16) || misaligned(dL_drotations, 16))
        return GHR_ERR_BAD_ARG;
    if (prm->shs && !dL_dsh) return GHR_ERR_BAD_ARG;
    const int stride = attr_stride(n_attr);
    if (stride == 0 || blend_throughput_regime(capacity) || !packed || !dL_dout_images || !packed_grad ||
        misaligned(packed, 16))
        return GHR_ERR_BAD_ARG;
    ghr_layout L;
    if ((r = multi_layout(prm, capacity, geom, geom_bytes, binning, binning_bytes, img, img_bytes, &L)) != GHR_OK) return r;
    cudaStream_t s = (cudaStream_t)stream;
    r = check_cuda(cudaMemsetAsync(packed_grad, 0, sizeof(float) * stride * (size_t)prm->P, s), "packed_grad memset");
    if (r != GHR_OK) return r;
    r = stage_done(prm,
                   launch_blend_backward_multi(*prm, (const char*)geom, (const char*)binning, (const char*)img, L,
                                               dL_dout_color, dL_dout_depth, dL_dout_alpha, accum, n_attr, packed,
                                               dL_dout_images, packed_grad, s),
                   s, "blend_backward_multi");
    if (r != GHR_OK) return r;
    return stage_done(prm,
                      launch_preprocess_backward(*prm, radii, (const char*)geom, L, accum, dL_dmeans2D, dL_dcolors,
                                                 dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, s),
                      s, "preprocess_backward");
}

int ghr_backward(const ghr_raster_params* prm, const int32_t* radii, const void* geom, size_t geom_bytes,
                 const void* binning, size_t binning_bytes, int64_t capacity, const void* img, size_t img_bytes,
                 const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha, float* accum,
                 float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                 float* dL_dsh, float* dL_dscales, float* dL_drotations, void* stream)
{
    int r = validate(prm);
    if (r != GHR_OK) return r;
    if (prm->P == 0) return GHR_OK;
    if (!radii || !dL_dout_color || !dL_dout_depth || !dL_dout_alpha || !accum || !dL_dmeans2D || !dL_dcolors ||
        !dL_dopacity || !dL_dmeans3D || !dL_dcov3D || misaligned(accum, 16) || misaligned(dL_drotations, 16))
        return GHR_ERR_BAD_ARG;
    if (prm->shs && !dL_dsh) return GHR_ERR_BAD_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    ghr_layout L;
    if (ghr_state_layout(prm->P, capacity, prm->W, prm->H, &L) != GHR_OK) return GHR_ERR_BAD_ARG;
    if (!geom || geom_bytes < L.geom_bytes || !img || img_bytes < L.img_bytes || !binning ||
        binning_bytes != L.binning_bytes)  // a different size means a different capacity, i.e. different array offsets
        return GHR_ERR_BUFFER;
    r = stage_done(prm,
                   launch_blend_backward(*prm, (const char*)geom, (const char*)binning, (const char*)img, L,
                                         dL_dout_color, dL_dout_depth, dL_dout_alpha, accum, s),
                   s, "blend_backward");
    if (r != GHR_OK) return r;
    return stage_done(prm,
                      launch_preprocess_backward(*prm, radii, (const char*)geom, L, accum, dL_dmeans2D, dL_dcolors,
                                                 dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, s),
                      s, "preprocess_backward");
}

// ---- multi-attribute rendering (include/ghr.h) ------------------------------------------------------------------
static int multi_layout(const ghr_raster_params* prm, int64_t capacity, const void* geom, size_t geom_bytes,
                        const void* binning, size_t binning_bytes, const void* img, size_t img_bytes, ghr_layout* L)
{
    if (ghr_state_layout(prm->P, capacity, prm->W, prm->H, L) != GHR_OK) return GHR_ERR_BAD_ARG;
    if (!geom || geom_bytes < L->geom_bytes || !img || img_bytes < L->img_bytes || !binning ||
        binning_bytes != L->binning_bytes)  // a different size means a different capacity, i.e. different array offsets
        return GHR_ERR_BUFFER;
    return GHR_OK;
}

int ghr_forward_reblend(const ghr_raster_params* prm, const float* colors, void* geom, size_t geom_bytes,
                        const void* binning, size_t binning_bytes, int64_t capacity, void* img, size_t img_bytes,
                        float* out_color, float* out_depth, float* out_alpha, void* stream)
{
    int r = validate(prm);
    if (r != GHR_OK) return r;
    if (!colors || !out_color || !out_depth || !out_alpha) return GHR_ERR_BAD_ARG;
    if (prm->P == 0) return GHR_OK;
    ghr_layout L;
    if ((r = multi_layout(prm, capacity, geom, geom_bytes, binning, binning_bytes, img, img_bytes, &L)) != GHR_OK) return r;
    cudaStream_t s = (cudaStream_t)stream;
    r = check_cuda(launch_recolor(prm->P, colors, (char*)geom, L, s), "recolor");
    if (r != GHR_OK) return r;
    return stage_done(prm,
                      launch_blend_forward(*prm, (const char*)geom, (const char*)binning, (char*)img, L, out_color,
                                           out_depth, out_alpha, s),
                      s, "blend_forward");
}

int ghr_backward_blend(const ghr_raster_params* prm, const float* colors, void* geom, size_t geom_bytes,
                       const void* binning, size_t binning_bytes, int64_t capacity, const void* img, size_t img_bytes,
                       const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha,
                       const float* zeros_plane, float* accum, void* stream)
{
    int r = validate(prm);
    if (r != GHR_OK) return r;
    if (prm->P == 0) return GHR_OK;
    if (!colors || !dL_dout_color || !accum || misaligned(accum, 16)) return GHR_ERR_BAD_ARG;
    if ((!dL_dout_depth || !dL_dout_alpha) && !zeros_plane) return GHR_ERR_BAD_ARG;
    ghr_layout L;
    if ((r = multi_layout(prm, capacity, geom, geom_bytes, binning, binning_bytes, img, img_bytes, &L)) != GHR_OK) return r;
    cudaStream_t s = (cudaStream_t)stream;
    r = check_cuda(launch_recolor(prm->P, colors, (char*)geom, L, s), "recolor");
    if (r != GHR_OK) return r;
    return stage_done(prm,
                      launch_blend_backward(*prm, (const char*)geom, (const char*)binning, (const char*)img, L,
                                            dL_dout_color, dL_dout_depth ? dL_dout_depth : zeros_plane,
                                            dL_dout_alpha ? dL_dout_alpha : zeros_plane, accum, s),
                      s, "blend_backward");
}

int ghr_backward_finish(const ghr_raster_params* prm, const int32_t* radii, const void* geom, size_t geom_bytes,
                        int64_t capacity, const float* accum, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                        float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales, float* dL_drotations,
                        void* stream)
{
    int r = validate(prm);
    if (r != GHR_OK) return r;
    if (prm->P == 0) return GHR_OK;
    if (!radii || !accum || !dL_dmeans2D || !dL_dcolors || !dL_dopacity || !dL_dmeans3D || !dL_dcov3D ||
        misaligned(accum, 16) || misaligned(dL_drotations, 16))
        return GHR_ERR_BAD_ARG;
    if (prm->shs && !dL_dsh) return GHR_ERR_BAD_ARG;
    ghr_layout L;
    if (ghr_state_layout(prm->P, capacity, prm->W, prm->H, &L) != GHR_OK) return GHR_ERR_BAD_ARG;
    if (!geom || geom_bytes < L.geom_bytes) return GHR_ERR_BUFFER;
    cudaStream_t s = (cudaStream_t)stream;
    return stage_done(prm,
                      launch_preprocess_backward(*prm, radii, (const char*)geom, L, accum, dL_dmeans2D, dL_dcolors,
                                                 dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, s),
                      s, "preprocess_backward");
}

int ghr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream)
{
    (void)projmatrix;
    if (P < 0) return GHR_ERR_BAD_ARG;
    if (P == 0) return GHR_OK;
    if (!means3D || !viewmatrix || !present) return GHR_ERR_BAD_ARG;
    return check_cuda(launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream), "mark_visible");
}

size_t ghr_sort_workspace_bytes(int64_t N)
{
    if (N < 0 || N >= (int64_t(1) << 30)) return 0;
    return align_up(sizeof(SortHeader)) + align_up(sizeof(uint32_t) * (size_t)kMaxPasses * (size_t)sort_tiles(N) * kRadix) + 256;
}

int ghr_sort_pairs_u64(int64_t N, int32_t bits, uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a,
                       uint32_t* vals_b, void* workspace, size_t workspace_bytes, uint32_t* sorted_sel_dev,
                       void* stream)
{
    if (N < 0 || N >= (int64_t(1) << 30) || bits <= 0 || bits > 64) return GHR_ERR_BAD_ARG;
    if (!workspace || workspace_bytes < ghr_sort_workspace_bytes(N) || misaligned(workspace, 256)) return GHR_ERR_BUFFER;
    if (N > 0 && (!keys_a || !keys_b || !vals_a || !vals_b)) return GHR_ERR_BAD_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    char* w = (char*)workspace;
    SortHeader* hdr = reinterpret_cast<SortHeader*>(w);
    uint32_t* status = reinterpret_cast<uint32_t*>(w + align_up(sizeof(SortHeader)));
    // clear the header and the look-back words of the passes this key width needs
    const size_t used = align_up(sizeof(SortHeader)) +
                        sizeof(uint32_t) * (size_t)sort_passes_for(bits) * (size_t)sort_tiles(N) * kRadix;
    int r = check_cuda(cudaMemsetAsync(w, 0, used, s), "sort memset");
    if (r != GHR_OK) return r;
    r = check_cuda(launch_sort(hdr, status, nullptr, N, bits, keys_a, keys_b, vals_a, vals_b, s), "sort");
    if (r != GHR_OK) return r;
    if (sorted_sel_dev)
        r = check_cuda(cudaMemcpyAsync(sorted_sel_dev, &hdr->sorted_sel, sizeof(uint32_t), cudaMemcpyDeviceToDevice, s),
                       "sort sel copy");
    return r;
}

size_t ghr_knn_workspace_bytes(int32_t P) { return knn_workspace_bytes(P); }

int ghr_dist2_knn3(int32_t P, const float* points, float* out, void* workspace, size_t workspace_bytes, void* stream)
{
    if (P < 0) return GHR_ERR_BAD_ARG;
    if (P == 0) return GHR_OK;
    if (!points || !out) return GHR_ERR_BAD_ARG;
    if (!workspace || workspace_bytes < knn_workspace_bytes(P) || misaligned(workspace, 256)) return GHR_ERR_BUFFER;
    return check_cuda(launch_knn(P, points, out, (char*)workspace, workspace_bytes, (cudaStream_t)stream), "knn");
}

}  // extern "C"

// ---------------------------------------------------------------------------
// LBS deformation entry points
// ---------------------------------------------------------------------------
namespace {
int validate_model(const ghr_lbs_model* m)
{
    if (!m || m->V <= 0 || m->NB < 0 || m->NB > 32 || !m->weights || !m->posedirs || !m->joint_template || !m->parents)
        return GHR_ERR_BAD_ARG;
    if (m->NB > 0 && (!m->shapedirs || !m->joint_shapedirs)) return GHR_ERR_BAD_ARG;
    if (misaligned(m->weights, 16)) return GHR_ERR_BAD_ARG;
    return GHR_OK;
}
int validate_deform(const ghr_deform_params* p)
{
    if (!p || p->P < 0 || p->V <= 0 || p->M < 0 || p->M > 16 || p->sh_degree < 0 || p->sh_degree > 3) return GHR_ERR_BAD_ARG;
    if (p->P == 0) return GHR_OK;
    if (!p->xyz || !p->vert_ids || !p->weights || !p->vertex_offsets || !p->pose_big || !p->pose_target)
        return GHR_ERR_BAD_ARG;
    if (p->shs && (p->M <= 0 || !p->campos || (p->sh_degree + 1) * (p->sh_degree + 1) > p->M)) return GHR_ERR_BAD_ARG;
    if (p->scales && !p->rotations) return GHR_ERR_BAD_ARG;
    if (misaligned(p->weights, 16) || misaligned(p->rotations, 16)) return GHR_ERR_BAD_ARG;
    return GHR_OK;
}
}  // namespace

extern "C" {

int ghr_lbs_prepare(int32_t V, int32_t NB, const float* v_template, const float* shapedirs, const float* J_regressor,
                    float* joint_template, float* joint_shapedirs, void* stream)
{
    if (V <= 0 || NB < 0 || NB > 32 || !v_template || !J_regressor || !joint_template) return GHR_ERR_BAD_ARG;
    if (NB > 0 && (!shapedirs || !joint_shapedirs)) return GHR_ERR_BAD_ARG;
    return check_cuda(launch_lbs_prepare(V, NB, v_template, shapedirs, J_regressor, joint_template, joint_shapedirs,
                                         (cudaStream_t)stream), "lbs_prepare");
}

int ghr_lbs_pose(const ghr_lbs_model* model, const float* poses, const float* shapes, const float* correct_Rs,
                 const float* R, const float* Th, float* pose_out, void* stream)
{
    int r = validate_model(model);
    if (r != GHR_OK) return r;
    if (!poses || !R || !Th || !pose_out || (model->NB > 0 && !shapes)) return GHR_ERR_BAD_ARG;
    return check_cuda(launch_lbs_pose(*model, poses, shapes, correct_Rs, R, Th, pose_out, (cudaStream_t)stream), "lbs_pose");
}

int ghr_lbs_vertex_offsets(const ghr_lbs_model* model, const float* pose_big, const float* pose_target,
                           const float* shapes, float* offsets, void* stream)
{
    int r = validate_model(model);
    if (r != GHR_OK) return r;
    if (!pose_target || !offsets) return GHR_ERR_BAD_ARG;
    return check_cuda(launch_lbs_vertex_offsets(*model, pose_big, pose_target, shapes, offsets, (cudaStream_t)stream),
                      "lbs_vertex_offsets");
}

int ghr_lbs_pose_views(const ghr_lbs_model* model, int32_t n_views, const ghr_pose_view* views, const float* pose_big,
                       void* stream)
{
    int r = validate_model(model);
    if (r != GHR_OK) return r;
    if (n_views < 0 || n_views > GHR_POSE_VIEWS_MAX || (n_views && !views)) return GHR_ERR_BAD_ARG;
    if (n_views == 0) return GHR_OK;
    for (int i = 0; i < n_views; i++) {
        const ghr_pose_view& v = views[i];
        if (!v.poses || !v.R || !v.Th || !v.pose_out || !v.offsets_out || (model->NB > 0 && !v.shapes)) return GHR_ERR_BAD_ARG;
    }
    return check_cuda(launch_lbs_pose_views(*model, n_views, views, pose_big, (cudaStream_t)stream), "lbs_pose_views");
}

size_t ghr_vertex_grid_workspace_bytes(int32_t V) { return vertex_grid_workspace_bytes(V); }

int ghr_vertex_grid_build(int32_t V, const float* t_vertices, void* workspace, size_t workspace_bytes, void* stream)
{
    if (V <= 0 || !t_vertices) return GHR_ERR_BAD_ARG;
    if (!workspace || workspace_bytes < vertex_grid_workspace_bytes(V) || misaligned(workspace, 256)) return GHR_ERR_BUFFER;
    return check_cuda(launch_vertex_grid_build(V, t_vertices, (char*)workspace, (cudaStream_t)stream), "vertex_grid_build");
}

int ghr_nearest_vertex(int32_t P, const float* query, int32_t V, const void* workspace, int32_t* vert_ids, void* stream)
{
    if (P < 0 || V <= 0) return GHR_ERR_BAD_ARG;
    if (P == 0) return GHR_OK;
    if (!query || !workspace || !vert_ids) return GHR_ERR_BAD_ARG;
    return check_cuda(launch_nearest_vertex(P, query, V, (const char*)workspace, vert_ids, (cudaStream_t)stream), "nearest_vertex");
}

int ghr_lbs_deform_forward(const ghr_deform_params* prm, float* smpl_pts, float* world_pts, float* bweights,
                           float* transforms, float* translation, float* world_normals, float* cov3D, float* colors,
                           uint8_t* clamped, void* stream)
{
    int r = validate_deform(prm);
    if (r != GHR_OK) return r;
    if (prm->P == 0) return GHR_OK;
    if (!world_pts || !transforms) return GHR_ERR_BAD_ARG;
    if (world_normals && !prm->normals) return GHR_ERR_BAD_ARG;
    if (cov3D && !prm->scales) return GHR_ERR_BAD_ARG;
    if (prm->opacity_logits && !prm->opacity_out) return GHR_ERR_BAD_ARG;
    if (colors && !prm->shs) return GHR_ERR_BAD_ARG;
    DeformOut o{smpl_pts, world_pts, bweights, transforms, translation, world_normals, cov3D, colors, clamped};
    return check_cuda(launch_lbs_deform_forward(*prm, o, (cudaStream_t)stream), "lbs_deform_forward");
}

int ghr_lbs_deform_backward(const ghr_deform_params* prm, const uint8_t* clamped, const float* dL_dworld_pts,
                            const float* dL_dtransforms, const float* dL_dworld_normals, const float* dL_dcov3D,
                            const float* dL_dcolors, float* dL_dxyz, float* dL_dnormals, float* dL_dlbs_offsets,
                            float* dL_dscales, float* dL_drotations, float* dL_dshs, float* dL_dA_target,
                            float* dL_dvertex_offsets, void* stream)
{
    int r = validate_deform(prm);
    if (r != GHR_OK) return r;
    if (prm->P == 0) return GHR_OK;
    if (!dL_dxyz) return GHR_ERR_BAD_ARG;
    if (dL_dscales && (!prm->scales || !dL_drotations)) return GHR_ERR_BAD_ARG;
    if (dL_dshs && !prm->shs) return GHR_ERR_BAD_ARG;
    if (misaligned(dL_drotations, 16)) return GHR_ERR_BAD_ARG;
    DeformGrads g{clamped, dL_dworld_pts, dL_dtransforms, dL_dworld_normals, dL_dcov3D, dL_dcolors, dL_dxyz, dL_dnormals,
                  dL_dlbs_offsets, dL_dscales, dL_drotations, dL_dshs, dL_dA_target, dL_dvertex_offsets};
    return check_cuda(launch_lbs_deform_backward(*prm, g, (cudaStream_t)stream), "lbs_deform_backward");
}

int ghr_lbs_pose_backward(const ghr_lbs_model* model, const float* pose_target, const float* correct_Rs,
                          const float* dL_dA_target, const float* dL_dvertex_offsets, float* dL_dcorrect_Rs, void* stream)
{
    (void)correct_Rs;
    int r = validate_model(model);
    if (r != GHR_OK) return r;
    if (!pose_target || !dL_dA_target || !dL_dcorrect_Rs) return GHR_ERR_BAD_ARG;
    // the 207-float pose-feature gradient is staged in the tail of dL_dcorrect_Rs' own buffer: callers pass
    // a buffer of 23*9 + 207 floats (documented in include/ghr.h)
    float* scratch = dL_dcorrect_Rs + 23 * 9;
    return check_cuda(launch_lbs_pose_backward(*model, pose_target, dL_dA_target, dL_dvertex_offsets, dL_dcorrect_Rs,
                                               scratch, (cudaStream_t)stream), "lbs_pose_backward");
}

int ghr_smpl_lbs_vertices(const ghr_lbs_model* model, const float* v_template, const float* pose_buf,
                          const float* offsets, float* verts, float* T_out, void* stream)
{
    int r = validate_model(model);
    if (r != GHR_OK) return r;
    if (!v_template || !pose_buf || !offsets || !verts) return GHR_ERR_BAD_ARG;
    return check_cuda(launch_smpl_lbs_vertices(*model, v_template, pose_buf, offsets, verts, T_out, (cudaStream_t)stream),
                      "smpl_lbs_vertices");
}

// ---- fused LBS + projection chain (include/ghr.h; kernels in ghr_fused.cu) -----------------------------------------
static int validate_fused(const ghr_deform_params* dp, const ghr_raster_params* view)
{
    int r = validate_deform(dp);
    if (r != GHR_OK) return r;
    if ((r = validate_view(view)) != GHR_OK) return r;
    if (view->P != dp->P) return GHR_ERR_BAD_ARG;
    if (dp->P == 0) return GHR_OK;
    if (!dp->scales || !dp->rotations || !dp->shs) return GHR_ERR_BAD_ARG;  // the chain needs covariance and colour
    if (!dp->opacity_logits && !view->opacities) return GHR_ERR_BAD_ARG;
    if (misaligned(dp->shs, 4)) return GHR_ERR_BAD_ARG;
    return GHR_OK;
}

int ghr_forward_deform_preprocess(const ghr_deform_params* dp, const ghr_raster_params* view, int32_t* radii, void* geom,
                                  size_t geom_bytes, float* world_pts, float* world_normals, float* cov3D, float* colors,
                                  float* opacity, int32_t* num_rendered_host, void* stream)
{
    int r = validate_fused(dp, view);
    if (r != GHR_OK) return r;
    cudaStream_t s = (cudaStream_t)stream;
    ghr_layout L;
    if (ghr_state_layout(dp->P, 0, view->W, view->H, &L) != GHR_OK) return GHR_ERR_BAD_ARG;
    if (!geom || geom_bytes < L.geom_bytes || misaligned(geom, 256)) return GHR_ERR_BUFFER;
    if (dp->P == 0) {
        r = check_cuda(cudaMemsetAsync(geom, 0, sizeof(ghr_geom_header), s), "deform_preprocess(P=0)");
        if (num_rendered_host) *num_rendered_host = 0;
        return r;
    }
    if (!radii) return GHR_ERR_BAD_ARG;
    if (world_normals && !dp->normals) return GHR_ERR_BAD_ARG;
    return stage_done(view,
                      launch_deform_project(*dp, *view, radii, (char*)geom, L, world_pts, world_normals, cov3D, colors,
                                            opacity, num_rendered_host, s),
                      s, "deform_project");
}

int ghr_forward_render_deformed(const ghr_raster_params* view, void* geom, size_t geom_bytes, void* binning,
                                size_t binning_bytes, int64_t capacity, void* img, size_t img_bytes, float* out_color,
                                float* out_depth, float* out_alpha, void* stream)
{
    return forward_render_impl(view, geom, geom_bytes, binning, binning_bytes, capacity, img, img_bytes, out_color,
                               out_depth, out_alpha, 0, nullptr, nullptr, stream, true);
}

int ghr_backward_deformed(const ghr_deform_params* dp, const ghr_raster_params* view, const int32_t* radii,
                          const void* geom, size_t geom_bytes, const void* binning, size_t binning_bytes, int64_t capacity,
                          const void* img, size_t img_bytes, const float* dL_dout_color, const float* dL_dout_depth,
                          const float* dL_dout_alpha, float* accum, float* dL_dmeans2D, float* dL_dxyz,
                          float* dL_dlbs_offsets, float* dL_dscales, float* dL_drotations, float* dL_dshs,
                          float* dL_dA_target, float* dL_dvertex_offsets, float* max_radii2D, float* xyz_gradient_accum,
                          float* denom, void* stream)
{
    int r = validate_fused(dp, view);
    if (r != GHR_OK) return r;
    if (dp->P == 0) return GHR_OK;
    if (!radii || !dL_dout_color || !dL_dout_depth || !dL_dout_alpha || !accum || misaligned(accum, 16) || !dL_dxyz ||
        !dL_dscales || !dL_drotations || !dL_dshs || misaligned(dL_drotations, 16))
        return GHR_ERR_BAD_ARG;
    if (dp->opacity_logits && !dp->dL_dopacity_logits) return GHR_ERR_BAD_ARG;
    ghr_layout L;
    if ((r = multi_layout(view, capacity, geom, geom_bytes, binning, binning_bytes, img, img_bytes, &L)) != GHR_OK) return r;
    cudaStream_t s = (cudaStream_t)stream;
    r = stage_done(view,
                   launch_blend_backward(*view, (const char*)geom, (const char*)binning, (const char*)img, L,
                                         dL_dout_color, dL_dout_depth, dL_dout_alpha, accum, s),
                   s, "blend_backward");
    if (r != GHR_OK) return r;
    DeformGrads g{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, dL_dxyz, nullptr, dL_dlbs_offsets, dL_dscales,
                  dL_drotations, dL_dshs, dL_dA_target, dL_dvertex_offsets};
    return stage_done(view, launch_project_deform_backward(*dp, *view, radii, (const char*)geom, L, accum, g, dL_dmeans2D,
                                                                   max_radii2D, xyz_gradient_accum, denom, s),
                      s, "project_deform_backward");
}

// ---- LBS-weight offset network (ghr_mlp.cu) ---------------------------------------------------------------------------
size_t ghr_lbs_mlp_blob_bytes(void) { return lbs_mlp_blob_bytes(); }

int ghr_lbs_mlp_pack(const float* const* weights, const float* const* biases, void* blob, void* stream)
{
    if (!weights || !biases || !blob || misaligned(blob, 16)) return GHR_ERR_BAD_ARG;
    for (int i = 0; i < 5; i++)
        if (!weights[i] || !biases[i]) return GHR_ERR_BAD_ARG;
    return check_cuda(launch_lbs_mlp_pack(weights, biases, blob, (cudaStream_t)stream), "lbs_mlp_pack");
}

int ghr_lbs_mlp_forward(int32_t P, const float* xyz, const void* blob, float* logits, void* stream)
{
    if (P < 0) return GHR_ERR_BAD_ARG;
    if (P == 0) return GHR_OK;
    if (!xyz || !blob || !logits || misaligned(blob, 16) || misaligned(logits, 16)) return GHR_ERR_BAD_ARG;
    return check_cuda(launch_lbs_mlp_forward(P, xyz, blob, logits, (cudaStream_t)stream), "lbs_mlp_forward");
}

int ghr_pose_refiner_forward(int32_t width, int32_t n_layers, const float* const* weights, const float* const* biases,
                             const float* pose69, float* correct_Rs, void* stream)
{
    if (width < 1 || width > GHR_REFINER_MAX_WIDTH || n_layers < 2 || n_layers > GHR_REFINER_MAX_LAYERS || !weights ||
        !biases || !pose69 || !correct_Rs)
        return GHR_ERR_BAD_ARG;
    for (int i = 0; i < n_layers; i++)
        if (!weights[i] || !biases[i]) return GHR_ERR_BAD_ARG;
    return check_cuda(launch_pose_refiner(width, n_layers, weights, biases, pose69, correct_Rs, (cudaStream_t)stream),
                      "pose_refiner");
}

}  // extern "C"
