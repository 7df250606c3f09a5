// THROUGHPUT-REGIME VARIANT (frames with more than ~1M (Gaussian, tile) pairs: thousands of busy tiles, every SM
// saturated).  There the kernel time is total work, not the heaviest item, so the coarser decomposition below --
// 8x4-pixel blocks, one lane per pixel, eight items per tile, bounding-box cull inside the kernel, no per-entry block
// masks -- does less list scanning per pixel than the 4x4 / two-lanes-per-pixel kernels of ghr_blend_fwd.cu, which
// win when a few heavy tiles set the pace.  The dispatch is in launch_blend_forward (blend_throughput_regime).
//
// Per-tile front-to-back alpha blending with depth and alpha-mask outputs
// (reference forward.cu:261-383 renderCUDA).
//
// Work decomposition (ours): a 16x16 tile is covered by eight 8x4 pixel blocks;
// one (tile, block) pair is a work item processed by ONE warp, independently of
// every other warp (no block-level barrier anywhere).  The kernel is persistent:
// (SM count) x 5 CTAs of 4 warps pull items from a global queue ordered by decreasing
// tile-list length (tile_order_kernel), so the few heavy tiles of a human-shaped
// scene start first and spread over all SMs.  Per warp, Gaussians are staged in
// batches of 64 packed 48-byte records
//   {px,py,depth,radius | conic.xyz,opacity | r,g,b,cull-extent}
// with cp.async (LDGSTS) into a 3-deep private shared-memory ring (list ids are
// fetched one batch ahead of the record copies), so the inner loop never
// touches global memory (the reference gathers colour and depth from global per
// contributing pair, forward.cu:360,362).
//
// Per warp and batch: (1) each lane tests one Gaussian's conservative bounding
// box (pack_cull_extent) against the warp's 8x4 block; a ballot yields the
// survivors.  (2) Survivors are evaluated four at a time: the power/exp/alpha
// of the four are independent (ILP), only the transmittance recurrence is
// sequential.  Skipped Gaussians are exactly those the reference's per-pixel
// tests (power > 0, alpha < 1/255) would skip for every pixel of the block, and
// the per-pair arithmetic is the reference's compiled sequence (pair_power,
// expf, fma accumulation), so final_T / n_contrib are bit-exact.
//
// Bound: fp32 issue + MUFU.EX2 along the longest per-pixel list, not HBM
// (algorithmic bytes 44*N + 28*X).
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

__device__ __forceinline__ void cp_async16w(void* smem, const void* gmem)
{
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit_w() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_w()
{
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

constexpr int kWideWarps = 4;    // warps per CTA (independent workers)
constexpr int kWideBatch = 64;   // Gaussians per staging round and warp
constexpr int kWideStages = 3;
constexpr int kWideCtasPerSm = 5;

__global__ void __launch_bounds__(kWideWarps * 32) blend_forward_wide_kernel(
    int W, int H, int tiles, const uint2* __restrict__ ranges, const uint32_t* __restrict__ order,
    const SortHeader* __restrict__ sh, const uint32_t* __restrict__ vals_a, const uint32_t* __restrict__ vals_b,
    const float4* __restrict__ rec, ghr_geom_header* __restrict__ hdr, int64_t capacity, const float* __restrict__ bg,
    float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_alpha,
    float* __restrict__ final_T, uint32_t* __restrict__ n_contrib)
{
    constexpr int BATCH = kWideBatch, NSTAGE = kWideStages, PER_LANE = BATCH / 32;
    __shared__ __align__(16) float4 s_all[kWideWarps][NSTAGE][BATCH * 3];

    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float4(*s_rec)[BATCH * 3] = s_all[warp];
    const uint32_t gx = (W + GHR_TILE - 1) / GHR_TILE;
    const bool overflow = (int64_t)hdr->num_rendered > capacity;
    const uint32_t* __restrict__ point_list = sh->sorted_sel ? vals_b : vals_a;
    const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];
    const uint32_t n_items = (uint32_t)tiles * 8u;

    for (;;) {
        uint32_t item = 0;
        if (lane == 0) item = atomicAdd(&hdr->blend_fwd_ticket, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= n_items) break;
        const uint32_t tile = order[item >> 3], wb = item & 7u;  // which of the tile's eight 8x4 blocks
        const uint32_t tile_x = tile % gx, tile_y = tile / gx;
        const uint32_t bx0 = tile_x * GHR_TILE + (wb & 1u) * 8u, by0 = tile_y * GHR_TILE + (wb >> 1) * 4u;
        const uint32_t pixx = bx0 + (lane & 7u), pixy = by0 + (lane >> 3);
        const bool inside = pixx < (uint32_t)W && pixy < (uint32_t)H;
        const float pxf = (float)pixx, pyf = (float)pixy;
        const float rx0 = (float)bx0, rx1 = (float)(bx0 + 7u), ry0 = (float)by0, ry1 = (float)(by0 + 3u);
        const uint2 range = overflow ? make_uint2(0, 0) : ranges[tile];
        const int todo_total = (int)(range.y - range.x);
        const int rounds = (todo_total + BATCH - 1) / BATCH;

        // Staging pipeline.  A record gather is two dependent L2 round trips (list entry -> record); the list
        // entries of a round are therefore loaded into registers one round before their records are copied.
        uint32_t nid[PER_LANE];
        auto load_ids = [&](int round) {
#pragma unroll
            for (int u = 0; u < PER_LANE; u++) {
                const uint32_t k = range.x + (uint32_t)round * BATCH + lane + u * 32;
                nid[u] = (round < rounds && k < range.y) ? point_list[k] : 0xFFFFFFFFu;
            }
        };
        auto stage = [&](int buf) {  // copies the records of the ids held in nid[]
#pragma unroll
            for (int u = 0; u < PER_LANE; u++) {
                if (nid[u] != 0xFFFFFFFFu) {
                    const float4* src = rec + 3 * (size_t)nid[u];
                    float4* dst = &s_rec[buf][(lane + u * 32) * 3];
                    cp_async16w(dst, src);
                    cp_async16w(dst + 1, src + 1);
                    cp_async16w(dst + 2, src + 2);
                }
            }
            cp_async_commit_w();
        };

        bool done = !inside;
        float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, wsum = 0.f, Dacc = 0.f;
        uint32_t last_contributor = 0;

        if (rounds > 0) {
            load_ids(0); stage(0);
            load_ids(1); stage(1);
            load_ids(2);
        }
        int toDo = todo_total;
        for (int i = 0; i < rounds; i++, toDo -= BATCH) {
            const int buf = i % NSTAGE;
            stage((i + 2) % NSTAGE);  // records of round i+2 (an empty group past the end keeps the count uniform)
            load_ids(i + 3);
            cp_async_wait_w<2>();       // this lane's copies for rounds <= i have landed
            __syncwarp();             // ... and everybody else's
            if (__all_sync(0xffffffffu, done)) break;  // forward.cu:313-315, per block instead of per tile
            const int n = min(BATCH, toDo);
            const float4* r = s_rec[buf];
            const uint32_t base = (uint32_t)i * BATCH;  // list position of r[0]
            for (int g = 0; g < n; g += 32) {
                // (1) conservative bounding-box cull, one Gaussian per lane
                const int jj = g + (int)lane;
                bool keep = false;
                if (jj < n) {
                    const float4 c0 = r[3 * jj];
                    float ex, ey;
                    unpack_cull_extent(__float_as_uint(r[3 * jj + 2].w), ex, ey);
                    const float ddx = fmaxf(fmaxf(rx0 - c0.x, c0.x - rx1), 0.0f);
                    const float ddy = fmaxf(fmaxf(ry0 - c0.y, c0.y - ry1), 0.0f);
                    keep = !(ddx > ex) && !(ddy > ey);
                }
                uint32_t mask = __ballot_sync(0xffffffffu, keep);
                // (2) survivors, four at a time
                while (mask) {
                    int j[4];
                    bool v[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        v[u] = mask != 0;
                        j[u] = v[u] ? g + __ffs(mask) - 1 : j[0];
                        mask &= mask - 1;
                    }
                    float alpha[4], depth[4], c0[4], c1[4], c2[4];
                    bool ok[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const float4 r0 = r[3 * j[u]], r1 = r[3 * j[u] + 1], r2 = r[3 * j[u] + 2];
                        const float dx = __fsub_rn(r0.x, pxf), dy = __fsub_rn(r0.y, pyf);
                        const float power = pair_power(dx, dy, r1.x, r1.y, r1.z);
                        alpha[u] = fminf(0.99f, __fmul_rn(r1.w, expf(power)));
                        ok[u] = v[u] && !(power > 0.0f) && !(alpha[u] < 1.0f / 255.0f);
                        depth[u] = r0.z; c0[u] = r2.x; c1[u] = r2.y; c2[u] = r2.z;
                    }
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        if (ok[u] && !done) {
                            const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha[u]));
                            if (test_T < 0.0001f) {
                                done = true;
                            } else {
                                C0 = __fmaf_rn(T, __fmul_rn(alpha[u], c0[u]), C0);
                                C1 = __fmaf_rn(T, __fmul_rn(alpha[u], c1[u]), C1);
                                C2 = __fmaf_rn(T, __fmul_rn(alpha[u], c2[u]), C2);
                                wsum = __fmaf_rn(T, alpha[u], wsum);
                                Dacc = __fmaf_rn(T, __fmul_rn(alpha[u], depth[u]), Dacc);
                                T = test_T;
                                last_contributor = base + (uint32_t)j[u] + 1u;
                            }
                        }
                    }
                    if (__all_sync(0xffffffffu, done)) { mask = 0; g = n; }
                }
            }
            __syncwarp();  // buffer `buf` is re-filled by the stage() of the next iteration (round i+3)
        }
        cp_async_wait_w<0>();
        __syncwarp();

        if (inside) {
            const size_t pix = (size_t)pixy * W + pixx, HW = (size_t)H * W;
            final_T[pix] = T;
            n_contrib[pix] = last_contributor;
            out_color[pix] = __fmaf_rn(T, bg0, C0);
            out_color[HW + pix] = __fmaf_rn(T, bg1, C1);
            out_color[2 * HW + pix] = __fmaf_rn(T, bg2, C2);
            out_alpha[pix] = wsum;
            out_depth[pix] = Dacc;
        }
    }
}

// Second form of the kernel above, used when the list values carry the tile's 8-bit block masks (wide_masks_packed,
// emit_keys): a warp reads the (id, mask) words of a round, keeps the entries whose mask has ITS block's bit -- the same
// bounding-box predicate, evaluated once per pair at emission instead of once per pair and block warp here -- and stages
// only those records, compacted in list order.  At 12 M pairs over 1080p four of five (entry, block) visits are culled:
// the first form copied every record of a tile eight times from L2 (4.7 GB) and ran the cull on all of them; this one
// copies the survivors.  Same survivors, same order, same arithmetic: final_T / n_contrib / images unchanged bit for bit.
__global__ void __launch_bounds__(kWideWarps * 32) blend_forward_wide_masked_kernel(
    int W, int H, int tiles, const uint2* __restrict__ ranges, const uint32_t* __restrict__ order,
    const SortHeader* __restrict__ sh, const uint32_t* __restrict__ vals_a, const uint32_t* __restrict__ vals_b,
    const float4* __restrict__ rec, ghr_geom_header* __restrict__ hdr, int64_t capacity, const float* __restrict__ bg,
    float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_alpha,
    float* __restrict__ final_T, uint32_t* __restrict__ n_contrib)
{
    constexpr int BATCH = kWideBatch, NSTAGE = kWideStages, PER_LANE = BATCH / 32;
    __shared__ __align__(16) float4 s_all[kWideWarps][NSTAGE][BATCH * 3];
    __shared__ uint8_t s_pos_all[kWideWarps][NSTAGE][BATCH];  // list position (within the round) of each staged survivor
    __shared__ uint32_t s_cnt_all[kWideWarps][NSTAGE];

    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    float4(*s_rec)[BATCH * 3] = s_all[warp];
    uint8_t(*s_pos)[BATCH] = s_pos_all[warp];
    uint32_t* s_cnt = s_cnt_all[warp];
    const uint32_t gx = (W + GHR_TILE - 1) / GHR_TILE;
    const bool overflow = (int64_t)hdr->num_rendered > capacity;
    const uint32_t* __restrict__ point_list = sh->sorted_sel ? vals_b : vals_a;
    const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];
    const uint32_t n_items = (uint32_t)tiles * 8u;

    for (;;) {
        uint32_t item = 0;
        if (lane == 0) item = atomicAdd(&hdr->blend_fwd_ticket, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= n_items) break;
        const uint32_t tile = order[item >> 3], wb = item & 7u;
        const uint32_t tile_x = tile % gx, tile_y = tile / gx;
        const uint32_t bx0 = tile_x * GHR_TILE + (wb & 1u) * 8u, by0 = tile_y * GHR_TILE + (wb >> 1) * 4u;
        const uint32_t pixx = bx0 + (lane & 7u), pixy = by0 + (lane >> 3);
        const bool inside = pixx < (uint32_t)W && pixy < (uint32_t)H;
        const float pxf = (float)pixx, pyf = (float)pixy;
        const uint2 range = overflow ? make_uint2(0, 0) : ranges[tile];
        const int todo_total = (int)(range.y - range.x);
        const int rounds = (todo_total + BATCH - 1) / BATCH;

        uint32_t nid[PER_LANE];  // id | mask << 24 (0 = nothing: mask 0)
        auto load_ids = [&](int round) {
#pragma unroll
            for (int u = 0; u < PER_LANE; u++) {
                const uint32_t k = range.x + (uint32_t)round * BATCH + lane + u * 32;
                nid[u] = (round < rounds && k < range.y) ? point_list[k] : 0u;
            }
        };
        auto stage = [&](int buf) {  // copies the surviving records of the entries held in nid[], compacted
            uint32_t total = 0;
#pragma unroll
            for (int u = 0; u < PER_LANE; u++) {
                const bool keep = (nid[u] >> (24u + wb)) & 1u;
                const uint32_t bits = __ballot_sync(0xffffffffu, keep);
                if (keep) {
                    const uint32_t pos = total + __popc(bits & lt_mask);
                    const float4* src = rec + 3 * (size_t)(nid[u] & 0x00FFFFFFu);
                    float4* dst = &s_rec[buf][pos * 3];
                    cp_async16w(dst, src);
                    cp_async16w(dst + 1, src + 1);
                    cp_async16w(dst + 2, src + 2);
                    s_pos[buf][pos] = (uint8_t)(lane + u * 32);
                }
                total += __popc(bits);
            }
            if (lane == 0) s_cnt[buf] = total;
            cp_async_commit_w();
        };

        bool done = !inside;
        float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, wsum = 0.f, Dacc = 0.f;
        uint32_t last_contributor = 0;

        if (rounds > 0) {
            load_ids(0); stage(0);
            load_ids(1); stage(1);
            load_ids(2);
        }
        for (int i = 0; i < rounds; i++) {
            const int buf = i % NSTAGE;
            stage((i + 2) % NSTAGE);  // records of round i+2 (an empty group past the end keeps the count uniform)
            load_ids(i + 3);
            cp_async_wait_w<2>();       // this lane's copies for rounds <= i have landed
            __syncwarp();             // ... and everybody else's (and s_pos / s_cnt of round i)
            if (__all_sync(0xffffffffu, done)) break;  // forward.cu:313-315, per block instead of per tile
            const int n = (int)s_cnt[buf];
            const float4* r = s_rec[buf];
            const uint32_t base = (uint32_t)i * BATCH;  // list position of the round's first entry
            for (int s0 = 0; s0 < n; s0 += 4) {
                float alpha[4], depth[4], c0[4], c1[4], c2[4];
                bool ok[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const bool v = s0 + u < n;
                    const int j = v ? s0 + u : s0;
                    const float4 r0 = r[3 * j], r1 = r[3 * j + 1], r2 = r[3 * j + 2];
                    const float dx = __fsub_rn(r0.x, pxf), dy = __fsub_rn(r0.y, pyf);
                    const float power = pair_power(dx, dy, r1.x, r1.y, r1.z);
                    alpha[u] = fminf(0.99f, __fmul_rn(r1.w, expf(power)));
                    ok[u] = v && !(power > 0.0f) && !(alpha[u] < 1.0f / 255.0f);
                    depth[u] = r0.z; c0[u] = r2.x; c1[u] = r2.y; c2[u] = r2.z;
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    if (ok[u] && !done) {
                        const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha[u]));
                        if (test_T < 0.0001f) {
                            done = true;
                        } else {
                            C0 = __fmaf_rn(T, __fmul_rn(alpha[u], c0[u]), C0);
                            C1 = __fmaf_rn(T, __fmul_rn(alpha[u], c1[u]), C1);
                            C2 = __fmaf_rn(T, __fmul_rn(alpha[u], c2[u]), C2);
                            wsum = __fmaf_rn(T, alpha[u], wsum);
                            Dacc = __fmaf_rn(T, __fmul_rn(alpha[u], depth[u]), Dacc);
                            T = test_T;
                            last_contributor = base + (uint32_t)s_pos[buf][s0 + u] + 1u;
                        }
                    }
                }
                if (__all_sync(0xffffffffu, done)) break;
            }
            __syncwarp();  // buffer `buf` is re-filled by the stage() of the next iteration (round i+3)
        }
        cp_async_wait_w<0>();
        __syncwarp();

        if (inside) {
            const size_t pix = (size_t)pixy * W + pixx, HW = (size_t)H * W;
            final_T[pix] = T;
            n_contrib[pix] = last_contributor;
            out_color[pix] = __fmaf_rn(T, bg0, C0);
            out_color[HW + pix] = __fmaf_rn(T, bg1, C1);
            out_color[2 * HW + pix] = __fmaf_rn(T, bg2, C2);
            out_alpha[pix] = wsum;
            out_depth[pix] = Dacc;
        }
    }
}

cudaError_t launch_blend_forward_wide(const ghr_raster_params& p, const char* geom, const char* binning, char* img,
                                 const ghr_layout& L, float* out_color, float* out_depth, float* out_alpha,
                                 cudaStream_t s)
{

    const int tiles = ((p.W + GHR_TILE - 1) / GHR_TILE) * ((p.H + GHR_TILE - 1) / GHR_TILE);
    const int64_t items = (int64_t)tiles * 8;
    int64_t ctas = (items + kWideWarps - 1) / kWideWarps;
    // several views in flight (ghr_set_shared_gpu): leave half of every SM's shared memory to another view's blend
    const int per_sm = blend_shared_gpu() ? (kWideCtasPerSm + 1) / 2 : kWideCtasPerSm;
    if (ctas > (int64_t)sm_count() * per_sm) ctas = (int64_t)sm_count() * per_sm;
    const int64_t capacity = L.capacity;
    ghr_geom_header* hdr = reinterpret_cast<ghr_geom_header*>(const_cast<char*>(geom) + L.g_header);
    cudaError_t e = cudaMemsetAsync(&hdr->blend_fwd_ticket, 0, sizeof(uint32_t), s);  // render may run twice per preprocess
    if (e != cudaSuccess) return e;
    auto kernel = wide_masks_packed(p.P, capacity) ? blend_forward_wide_masked_kernel : blend_forward_wide_kernel;
    kernel<<<(unsigned)ctas, kWideWarps * 32, 0, s>>>(
        p.W, p.H, tiles, reinterpret_cast<const uint2*>(img + L.i_ranges),
        reinterpret_cast<const uint32_t*>(img + L.i_order), reinterpret_cast<const SortHeader*>(binning + L.b_header),
        reinterpret_cast<const uint32_t*>(binning + L.b_vals_a), reinterpret_cast<const uint32_t*>(binning + L.b_vals_b),
        reinterpret_cast<const float4*>(geom + L.g_rec), hdr, capacity, p.background, out_color,
        out_depth, out_alpha, reinterpret_cast<float*>(img + L.i_final_T), reinterpret_cast<uint32_t*>(img + L.i_n_contrib));
    return cudaGetLastError();
}

}  // namespace ghr
