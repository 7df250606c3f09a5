// View-batched forward rasterisation: many small views of the SAME Gaussians in one set of launches.
//
// The reference's baking.py::bake_set (:136-309) renders, for every occupied grid cell, the six 32x32 faces of a
// cube map with one _C.rasterize_gaussians call each (thousands of calls of ~8 kernels, plus a boolean-mask gather
// of every Gaussian tensor per cell, :237-246), and only keeps the alpha map (:283).  Here a batch of n views is
// one flattened problem:
//   views_preprocess_kernel  (view, Gaussian) pairs -> records + tiles_touched + inclusive scan (look-back);
//                            a Gaussian whose label equals the view's skip label is left out (the render_mask);
//   views_emit_keys_kernel   key = (view * tiles + tile) << 32 | depth bits, value = view * P + Gaussian;
//   onesweep sort            ONE sort over all views (ghr_sort.cu), 32 + msb(n_views * tiles) key bits;
//   views_ranges_kernel      identifyTileRanges over n_views * tiles lists;
//   views_blend_kernel       one CTA per (view, tile), the reference's renderCUDA loop (forward.cu:261-383) with
//                            the pinned arithmetic of the single-view kernels -> images identical to per-view calls.
// All stages are HBM / latency streams; the win is launch count (6 launches per batch instead of ~8 per view) and
// no per-cell gathers.
#include <stdio.h>
#include "ghr_project.cuh"

namespace ghr {
namespace {

constexpr int kViewsSlots = 4;  // (view, Gaussian) slots per thread of views_preprocess_kernel

struct ViewsArgs {
    PreArgs pre;             // shared Gaussian inputs; view / proj / campos unused
    int n_views;
    const float* cameras;    // [n_views][36]: view 16 | proj 16 | campos 3 | pad
    const int32_t* labels;   // [P] or NULL
    const int32_t* skip;     // [n_views] or NULL
    int64_t total;           // n_views * P
};

__global__ void __launch_bounds__(GHR_BLOCK) views_preprocess_kernel(const ViewsArgs va)
{
    __shared__ uint32_t s_ticket;
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_excl;
    const PreArgs& a = va.pre;
    const int tid = threadIdx.x;
    if (tid == 0) s_ticket = atomicAdd(&a.hdr->scan_ticket, 1u);
    __syncthreads();
    const int tile = (int)s_ticket;
    // kViewsSlots consecutive (view, Gaussian) slots per thread: the look-back chain has one link per CTA, and with one
    // slot per thread the 2*10^5 links of a bake (5.5*10^7 slots) were the kernel time (ncu: issue active 14 %)
    const int64_t idx0 = ((int64_t)tile * GHR_BLOCK + tid) * kViewsSlots;

    uint32_t t[kViewsSlots];
    uint32_t tiles = 0;
#pragma unroll 1
    for (int k = 0; k < kViewsSlots; k++) {
        const int64_t idx = idx0 + k;
        uint32_t n = 0;
        if (idx < va.total) {
            const int v = (int)(idx / a.P);
            const int g = (int)(idx - (int64_t)v * a.P);
            int radius = 0;
            bool near_culled;
            const bool skipped = va.labels && va.skip && va.labels[g] == va.skip[v];
            if (!skipped) {
                const float* cam = va.cameras + 36 * (size_t)v;
                n = project_gaussian(a, g, (size_t)idx, cam, cam + 16, cam + 32, radius, near_culled);
            }
            if (a.radii) a.radii[idx] = radius;
            a.tiles_touched[idx] = n;
        }
        t[k] = n;
        tiles += n;
    }

    // block-wide inclusive scan + decoupled look-back (same as preprocess_kernel)
    const uint32_t lane = lane_id(), warp = tid >> 5;
    uint32_t v = tiles;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= (uint32_t)o) v += u;
    }
    if (lane == 31) s_warp[warp] = v;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < 8 ? s_warp[lane] : 0;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= (uint32_t)o) w += u;
        }
        const uint32_t block_total = __shfl_sync(0xffffffffu, w, 7);
        const uint32_t prev = __shfl_up_sync(0xffffffffu, w, 1);
        if (lane < 8) s_warp[lane] = (lane == 0) ? 0u : prev;
        const uint32_t excl = lookback_exclusive(a.scan_status, tile, block_total);
        if (lane == 0) {
            s_excl = excl;
            if (tile == a.nblocks - 1) a.hdr->num_rendered = excl + block_total;
        }
    }
    __syncthreads();
    uint32_t run = s_excl + s_warp[warp] + v - tiles;  // exclusive prefix of this thread's first slot
#pragma unroll
    for (int k = 0; k < kViewsSlots; k++) {
        run += t[k];
        if (idx0 + k < va.total) a.point_offsets[idx0 + k] = run;
    }
}

// rasterizer_impl.cu:70-111 duplicateWithKeys over the flattened (view, Gaussian) space
__global__ void __launch_bounds__(GHR_BLOCK) views_emit_keys_kernel(int64_t total, int P, int W, int H,
                                                                    const uint32_t* __restrict__ tiles_touched,
                                                                    const float4* __restrict__ rec,
                                                                    const uint32_t* __restrict__ point_offsets,
                                                                    const ghr_geom_header* __restrict__ hdr,
                                                                    int64_t capacity, uint64_t* __restrict__ keys,
                                                                    uint32_t* __restrict__ vals)
{
    if ((int64_t)hdr->num_rendered > capacity) return;
    const int64_t idx = (int64_t)blockIdx.x * GHR_BLOCK + threadIdx.x;
    if (idx >= total || tiles_touched[idx] == 0) return;
    const uint32_t gx = (W + GHR_TILE - 1) / GHR_TILE, gy = (H + GHR_TILE - 1) / GHR_TILE;
    const float4 r0 = rec[3 * (size_t)idx];
    uint32_t minx, miny, maxx, maxy;
    get_rect(r0.x, r0.y, __float_as_int(r0.w), gx, gy, minx, miny, maxx, maxy);
    uint32_t off = idx == 0 ? 0u : point_offsets[idx - 1];
    const uint32_t dbits = __float_as_uint(r0.z);
    const uint32_t tile0 = (uint32_t)(idx / P) * gx * gy;
    for (uint32_t ty = miny; ty < maxy; ty++)
        for (uint32_t tx = minx; tx < maxx; tx++) {
            keys[off] = ((uint64_t)(tile0 + ty * gx + tx) << 32) | dbits;
            vals[off] = (uint32_t)idx;
            off++;
        }
}

// rasterizer_impl.cu:116-138 identifyTileRanges; also latches a capacity overflow
__global__ void __launch_bounds__(GHR_BLOCK) views_ranges_kernel(ghr_geom_header* hdr, int64_t capacity,
                                                                 const SortHeader* __restrict__ sh,
                                                                 const uint64_t* __restrict__ keys_a,
                                                                 const uint64_t* __restrict__ keys_b, uint2* ranges)
{
    const int64_t L = hdr->num_rendered;
    if (L > capacity) {
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicMax(&hdr->error, (uint32_t)(-GHR_ERR_CAPACITY));
        return;
    }
    const uint64_t* __restrict__ keys = sh->sorted_sel ? keys_b : keys_a;
    for (int64_t idx = (int64_t)blockIdx.x * GHR_BLOCK + threadIdx.x; idx < L; idx += (int64_t)gridDim.x * GHR_BLOCK) {
        const uint32_t cur = (uint32_t)(keys[idx] >> 32);
        if (idx == 0) ranges[cur].x = 0;
        else {
            const uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
            if (cur != prev) { ranges[prev].y = (uint32_t)idx; ranges[cur].x = (uint32_t)idx; }
        }
        if (idx == L - 1) ranges[cur].y = (uint32_t)L;
    }
}

// forward.cu:261-383 renderCUDA: one CTA per (view, tile), one thread per pixel, rounds of 256 staged entries.
template <bool kColor>
__global__ void __launch_bounds__(GHR_BLOCK) views_blend_kernel(int W, int H, int tiles_per_view,
                                                                const ghr_geom_header* __restrict__ hdr, int64_t capacity,
                                                                const SortHeader* __restrict__ sh,
                                                                const uint32_t* __restrict__ vals_a,
                                                                const uint32_t* __restrict__ vals_b,
                                                                const uint2* __restrict__ ranges,
                                                                const float4* __restrict__ rec,
                                                                const float* __restrict__ bg, float* __restrict__ out_color,
                                                                float* __restrict__ out_depth, float* __restrict__ out_alpha)
{
    if ((int64_t)hdr->num_rendered > capacity) return;
    __shared__ float4 s_r0[GHR_BLOCK], s_r1[GHR_BLOCK];
    __shared__ float4 s_r2[kColor ? GHR_BLOCK : 1];
    const uint32_t* __restrict__ vals = sh->sorted_sel ? vals_b : vals_a;
    const int gx = (W + GHR_TILE - 1) / GHR_TILE;
    const int view = blockIdx.x / tiles_per_view, lt = blockIdx.x - view * tiles_per_view;
    const int tid = threadIdx.x;
    const int pixx = (lt % gx) * GHR_TILE + (tid & 15), pixy = (lt / gx) * GHR_TILE + (tid >> 4);
    const bool inside = pixx < W && pixy < H;
    const float pxf = (float)pixx, pyf = (float)pixy;
    const uint2 range = ranges[blockIdx.x];
    const int rounds = (int)((range.y - range.x + GHR_BLOCK - 1) / GHR_BLOCK);
    int todo = (int)(range.y - range.x);
    bool done = !inside;
    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, wsum = 0.f, Dacc = 0.f;
    for (int i = 0; i < rounds; i++, todo -= GHR_BLOCK) {
        if (__syncthreads_count(done) == GHR_BLOCK) break;
        const uint32_t e = range.x + (uint32_t)i * GHR_BLOCK + tid;
        if (e < range.y) {
            const size_t id = vals[e];
            s_r0[tid] = rec[3 * id];
            s_r1[tid] = rec[3 * id + 1];
            if (kColor) s_r2[tid] = rec[3 * id + 2];
        }
        __syncthreads();
        const int n = min(GHR_BLOCK, todo);
        // Four entries at a time: power / exp / alpha of the four are independent (ILP), the transmittance recurrence
        // then runs over them in list order with the reference's tests, so every pixel sees the reference's sequence.
        // A list is one dependent chain per pixel and the longest list of the batch sets the kernel time, so the
        // latency per entry is what matters here.
        for (int j = 0; !done && j < n; j += 4) {
            float alpha[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float4 r0 = s_r0[j + u], r1 = s_r1[j + u];  // j + u <= 255; entries past n are stale and masked
                const float dx = __fsub_rn(r0.x, pxf), dy = __fsub_rn(r0.y, pyf);
                const float power = pair_power(dx, dy, r1.x, r1.y, r1.z);
                alpha[u] = fminf(0.99f, __fmul_rn(r1.w, expf(power)));
                ok[u] = (j + u < n) && !(power > 0.0f) && !(alpha[u] < 1.0f / 255.0f);
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                if (!ok[u] || done) continue;
                const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha[u]));
                if (test_T < 0.0001f) { done = true; continue; }
                if (kColor) {
                    const float4 r2 = s_r2[j + u];
                    C0 = __fmaf_rn(T, __fmul_rn(alpha[u], r2.x), C0);
                    C1 = __fmaf_rn(T, __fmul_rn(alpha[u], r2.y), C1);
                    C2 = __fmaf_rn(T, __fmul_rn(alpha[u], r2.z), C2);
                    Dacc = __fmaf_rn(T, __fmul_rn(alpha[u], s_r0[j + u].z), Dacc);
                }
                wsum = __fmaf_rn(T, alpha[u], wsum);
                T = test_T;
            }
        }
    }
    if (inside) {
        const size_t HW = (size_t)H * W, pix = (size_t)pixy * W + pixx;
        out_alpha[(size_t)view * HW + pix] = wsum;
        if (kColor) {
            float* c = out_color + (size_t)view * 3 * HW + pix;
            c[0] = __fmaf_rn(T, bg[0], C0);
            c[HW] = __fmaf_rn(T, bg[1], C1);
            c[2 * HW] = __fmaf_rn(T, bg[2], C2);
            if (out_depth) out_depth[(size_t)view * HW + pix] = Dacc;
        }
    }
}

struct ViewsLayout {
    size_t header, scan_status, rec, tiles_touched, point_offsets, ranges, s_header, s_status, keys_a, keys_b, vals_a,
        vals_b, bytes;
    int bits;
};

bool views_layout(int64_t P, int64_t n_views, int W, int H, int64_t capacity, ViewsLayout* L)
{
    if (P < 0 || n_views <= 0 || W <= 0 || H <= 0 || capacity < 0 || capacity >= (int64_t(1) << 30)) return false;
    const int64_t total = P * n_views;
    const int64_t tiles = (int64_t)((W + GHR_TILE - 1) / GHR_TILE) * ((H + GHR_TILE - 1) / GHR_TILE) * n_views;
    if (total >= (int64_t(1) << 31) || tiles >= (int64_t(1) << 31)) return false;
    L->bits = 32 + (int)higher_msb((uint32_t)tiles);
    const size_t T = (size_t)total;
    size_t o = 0;
    L->header = o; o += align_up(sizeof(ghr_geom_header));
    L->scan_status = o; o += align_up(sizeof(uint64_t) * ((T + GHR_BLOCK * kViewsSlots - 1) / (GHR_BLOCK * kViewsSlots) + 1));
    L->rec = o; o += align_up(sizeof(float4) * 3 * T);
    L->tiles_touched = o; o += align_up(sizeof(uint32_t) * T);
    L->point_offsets = o; o += align_up(sizeof(uint32_t) * T);
    L->ranges = o; o += align_up(sizeof(uint2) * (size_t)tiles);
    L->s_header = o; o += align_up(sizeof(SortHeader));
    L->s_status = o; o += align_up(sizeof(uint32_t) * (size_t)sort_passes_for(L->bits) * (size_t)sort_tiles(capacity) * kRadix);
    L->keys_a = o; o += align_up(sizeof(uint64_t) * (size_t)capacity);
    L->keys_b = o; o += align_up(sizeof(uint64_t) * (size_t)capacity);
    L->vals_a = o; o += align_up(sizeof(uint32_t) * (size_t)capacity);
    L->vals_b = o; o += align_up(sizeof(uint32_t) * (size_t)capacity);
    L->bytes = o;
    return true;
}

int cuda_status(cudaError_t e, const char* what)
{
    if (e == cudaSuccess) return GHR_OK;
    fprintf(stderr, "[libghr] CUDA error in %s: %s\n", what, cudaGetErrorString(e));
    return GHR_ERR_CUDA;
}

}  // namespace
}  // namespace ghr

using namespace ghr;

extern "C" {

size_t ghr_views_workspace_bytes(int32_t P, int32_t n_views, int32_t W, int32_t H, int64_t capacity)
{
    ViewsLayout L;
    return views_layout(P, n_views, W, H, capacity, &L) ? L.bytes : 0;
}

int ghr_forward_views(const ghr_raster_params* prm, int32_t n_views, const float* cameras, const int32_t* labels,
                      const int32_t* skip_labels, void* workspace, size_t workspace_bytes, int64_t capacity,
                      float* out_color, float* out_depth, float* out_alpha, int32_t* radii, void* stream)
{
    if (!prm || n_views < 0) return GHR_ERR_BAD_ARG;
    if (n_views == 0) return GHR_OK;
    if (prm->P < 0 || prm->W <= 0 || prm->H <= 0 || prm->M < 0 || prm->D < 0 || prm->D > 3) return GHR_ERR_BAD_ARG;
    if (!cameras || !out_alpha || (out_depth && !out_color)) return GHR_ERR_BAD_ARG;
    if ((labels == nullptr) != (skip_labels == nullptr)) return GHR_ERR_BAD_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t X = (size_t)prm->W * prm->H;
    ViewsLayout L;
    if (!views_layout(prm->P, n_views, prm->W, prm->H, capacity, &L)) return GHR_ERR_BAD_ARG;
    if (!workspace || workspace_bytes < L.bytes || (reinterpret_cast<uintptr_t>(workspace) & 255)) return GHR_ERR_BUFFER;
    char* w = (char*)workspace;
    if (prm->P == 0) {  // rasterize_points.cu:83: no Gaussians -> the zero-filled outputs
        int r = cuda_status(cudaMemsetAsync(out_alpha, 0, sizeof(float) * X * n_views, s), "views P=0");
        if (r == GHR_OK && out_color) r = cuda_status(cudaMemsetAsync(out_color, 0, sizeof(float) * 3 * X * n_views, s), "views P=0");
        if (r == GHR_OK && out_depth) r = cuda_status(cudaMemsetAsync(out_depth, 0, sizeof(float) * X * n_views, s), "views P=0");
        if (r == GHR_OK) r = cuda_status(cudaMemsetAsync(w + L.header, 0, sizeof(ghr_geom_header), s), "views P=0");
        return r;
    }
    if (!prm->means3D || !prm->opacities) return GHR_ERR_BAD_ARG;
    const bool want_color = out_color != nullptr;
    if (want_color && (!prm->background || (!prm->shs && !prm->colors_precomp))) return GHR_ERR_BAD_ARG;
    if (want_color && prm->shs && !prm->colors_precomp && (prm->M <= 0 || (prm->D + 1) * (prm->D + 1) > prm->M || prm->M > 16))
        return GHR_ERR_BAD_ARG;
    if (!prm->cov3D_precomp && (!prm->scales || !prm->rotations)) return GHR_ERR_BAD_ARG;
    if ((reinterpret_cast<uintptr_t>(prm->cov3D_precomp) & 7) || (reinterpret_cast<uintptr_t>(prm->rotations) & 15) ||
        (reinterpret_cast<uintptr_t>(prm->shs) & 15))
        return GHR_ERR_BAD_ARG;

    ViewsArgs va;
    PreArgs& a = va.pre;
    a.P = prm->P; a.D = prm->D; a.M = prm->M; a.W = prm->W; a.H = prm->H;
    a.scale_modifier = prm->scale_modifier; a.tan_fovx = prm->tan_fovx; a.tan_fovy = prm->tan_fovy;
    a.focal_y = prm->H / (2.0f * prm->tan_fovy);  // rasterizer_impl.cu:224-225
    a.focal_x = prm->W / (2.0f * prm->tan_fovx);
    a.prefiltered = 0;
    a.means3D = prm->means3D; a.shs = prm->shs; a.colors_precomp = prm->colors_precomp; a.opacities = prm->opacities;
    a.scales = prm->scales; a.rotations = prm->rotations; a.cov3D_precomp = prm->cov3D_precomp;
    a.view = a.proj = a.campos = nullptr;
    a.radii = radii;
    a.rec = reinterpret_cast<float4*>(w + L.rec);
    a.tiles_touched = reinterpret_cast<uint32_t*>(w + L.tiles_touched);
    a.point_offsets = reinterpret_cast<uint32_t*>(w + L.point_offsets);
    a.cov3D_out = nullptr;
    a.clamped = nullptr;
    a.scan_status = reinterpret_cast<uint64_t*>(w + L.scan_status);
    a.hdr = reinterpret_cast<ghr_geom_header*>(w + L.header);
    a.n_host = nullptr;
    a.skip_rgb = want_color ? 0 : 1;
    a.split_scan = 0;
    va.n_views = n_views; va.cameras = cameras; va.labels = labels; va.skip = skip_labels;
    va.total = (int64_t)prm->P * n_views;
    a.nblocks = (int)((va.total + GHR_BLOCK * kViewsSlots - 1) / (GHR_BLOCK * kViewsSlots));
    const int emit_blocks = (int)((va.total + GHR_BLOCK - 1) / GHR_BLOCK);
    const int tiles_per_view = ((prm->W + GHR_TILE - 1) / GHR_TILE) * ((prm->H + GHR_TILE - 1) / GHR_TILE);
    const int64_t tiles = (int64_t)tiles_per_view * n_views;

    int r;
    // header + look-back words are contiguous; ranges, sort header and its look-back words too
    if ((r = cuda_status(cudaMemsetAsync(w + L.header, 0, L.rec - L.header, s), "views memset")) != GHR_OK) return r;
    if ((r = cuda_status(cudaMemsetAsync(w + L.ranges, 0, L.keys_a - L.ranges, s), "views memset")) != GHR_OK) return r;
    {
        StageTimer t(GHR_STAGE_PREPROCESS, s);
        views_preprocess_kernel<<<a.nblocks, GHR_BLOCK, 0, s>>>(va);
        count_launch();
    }
    if ((r = cuda_status(cudaGetLastError(), "views_preprocess")) != GHR_OK) return r;
    if (capacity > 0) {
        {
            StageTimer t(GHR_STAGE_EMIT_KEYS, s);
            views_emit_keys_kernel<<<emit_blocks, GHR_BLOCK, 0, s>>>(va.total, prm->P, prm->W, prm->H, a.tiles_touched, a.rec,
                                                                   a.point_offsets, a.hdr, capacity,
                                                                   reinterpret_cast<uint64_t*>(w + L.keys_a),
                                                                   reinterpret_cast<uint32_t*>(w + L.vals_a));
            count_launch();
        }
        if ((r = cuda_status(cudaGetLastError(), "views_emit_keys")) != GHR_OK) return r;
        r = cuda_status(launch_sort(reinterpret_cast<SortHeader*>(w + L.s_header), reinterpret_cast<uint32_t*>(w + L.s_status),
                                    &a.hdr->num_rendered, capacity, L.bits, reinterpret_cast<uint64_t*>(w + L.keys_a),
                                    reinterpret_cast<uint64_t*>(w + L.keys_b), reinterpret_cast<uint32_t*>(w + L.vals_a),
                                    reinterpret_cast<uint32_t*>(w + L.vals_b), s),
                        "views sort");
        if (r != GHR_OK) return r;
    }
    {
        StageTimer t(GHR_STAGE_TILE_RANGES, s);
        int64_t grid = (capacity + GHR_BLOCK - 1) / GHR_BLOCK;
        grid = grid > sm_count() * 16 ? sm_count() * 16 : grid < 1 ? 1 : grid;
        views_ranges_kernel<<<(unsigned)grid, GHR_BLOCK, 0, s>>>(a.hdr, capacity, reinterpret_cast<const SortHeader*>(w + L.s_header),
                                                                 reinterpret_cast<const uint64_t*>(w + L.keys_a),
                                                                 reinterpret_cast<const uint64_t*>(w + L.keys_b),
                                                                 reinterpret_cast<uint2*>(w + L.ranges));
        count_launch();
    }
    if ((r = cuda_status(cudaGetLastError(), "views_ranges")) != GHR_OK) return r;
    {
        StageTimer t(GHR_STAGE_BLEND_FWD, s);
        auto kern = want_color ? views_blend_kernel<true> : views_blend_kernel<false>;
        kern<<<(unsigned)tiles, GHR_BLOCK, 0, s>>>(prm->W, prm->H, tiles_per_view, a.hdr, capacity,
                                                   reinterpret_cast<const SortHeader*>(w + L.s_header),
                                                   reinterpret_cast<const uint32_t*>(w + L.vals_a),
                                                   reinterpret_cast<const uint32_t*>(w + L.vals_b),
                                                   reinterpret_cast<const uint2*>(w + L.ranges), a.rec, prm->background, out_color,
                                                   out_depth, out_alpha);
        count_launch();
    }
    return cuda_status(cudaGetLastError(), "views_blend");
}

}  // extern "C"
