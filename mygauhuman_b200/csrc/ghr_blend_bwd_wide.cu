// THROUGHPUT-REGIME VARIANT (frames with more than ~1M (Gaussian, tile) pairs: thousands of busy tiles, every SM
// saturated).  There the kernel time is total work, not the heaviest item, so the coarser decomposition below --
// 8x4-pixel blocks, one lane per pixel, eight items per tile, bounding-box cull inside the kernel, no per-entry block
// masks -- does less list scanning per pixel than the 4x4 / two-lanes-per-pixel kernels of ghr_blend_bwd.cu, which
// win when a few heavy tiles set the pace.  The dispatch is in launch_blend_backward (blend_throughput_regime).
//
// Backward of the per-tile blend (reference backward.cu:399-587 renderCUDA).
//
// The reference issues 9 fp32 global atomicAdds per contributing
// (pixel, Gaussian) pair (backward.cu:538,574-584) -- its hottest spot.  Here
// every warp (a 16x2 pixel strip) first votes whether any of its pixels is
// touched by the Gaussian (ballot -> skip), then reduces the 9 partial
// gradients across its 32 pixels with a split butterfly (12 shuffles instead
// of 45: at each xor step half of the lanes keeps half of the values), and
// only then 9 lanes issue one atomic each into a packed per-Gaussian
// accumulator [P][12] (one 48-byte record -> the 9 atomics of a Gaussian hit
// at most two 32-byte sectors).  The per-Gaussian preprocess backward reads
// that record, so no separate dL_dconic/dL_dmean2D arrays are zeroed.
//
// accum record: 0-2 dL_dcolor, 3-4 dL_dmean2D.xy, 5-7 dL_dconic (x,y,w),
//               8 dL_dopacity, 9-11 unused.
// Bound: fp32 issue / MUFU / SHFL, not HBM (44*N + 28*X + 36*V bytes).
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

__device__ __forceinline__ void cp_async16w(void* smem, const void* gmem)
{
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}

// Sum v[0..8] over the warp.  Returns the total of value `slot` (see
// reduce_slot_w()) in every lane; lanes whose slot is >= 9 return garbage-free 0.
__device__ __forceinline__ float warp_reduce9_w(const float (&v)[9], uint32_t lane)
{
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
    float a[6];
#pragma unroll
    for (int k = 0; k < 5; k++) {
        const float lo = v[k], hi = (k < 4) ? v[5 + k] : 0.0f;
        const float keep = b4 ? hi : lo, send = b4 ? lo : hi;
        a[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    a[5] = 0.0f;
    float c[4];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float keep = b3 ? a[3 + k] : a[k], send = b3 ? a[k] : a[3 + k];
        c[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    c[3] = 0.0f;
    float d[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const float keep = b2 ? c[2 + k] : c[k], send = b2 ? c[k] : c[2 + k];
        d[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    const float keep = b1 ? d[1] : d[0], send = b1 ? d[0] : d[1];
    float e = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    e += __shfl_xor_sync(0xffffffffu, e, 1);
    return e;
}

// Which of the 9 values a lane holds after warp_reduce9_w (>= 9: none).
__device__ __forceinline__ int reduce_slot_w(uint32_t lane)
{
    const int w4 = ((lane & 4) ? 2 : 0) + ((lane & 2) ? 1 : 0);  // position inside the padded group of 4
    if (w4 >= 3) return 99;
    const int local = ((lane & 8) ? 3 : 0) + w4;                 // position inside the padded group of 6
    if (lane & 16) return local < 4 ? 5 + local : 99;
    return local < 5 ? local : 99;
}

constexpr int kWideBwdWarps = 4;     // independent warps per CTA
constexpr int kWideBwdBatch = 64;    // Gaussians per staging round and warp
constexpr int kWideBwdStages = 3;
constexpr int kWideBwdCtasPerSm = 4;

// Persistent kernel: warps pull (tile, 8x4 block) work items from a queue ordered by decreasing tile-list
// length (tile_order_kernel) and work independently (no block-level barrier); see ghr_blend_fwd.cu.
__global__ void __launch_bounds__(kWideBwdWarps * 32) blend_backward_wide_kernel(
    int W, int H, int tiles, const uint2* __restrict__ ranges, const uint32_t* __restrict__ order,
    const SortHeader* __restrict__ sh, const uint32_t* __restrict__ vals_a, const uint32_t* __restrict__ vals_b,
    const float4* __restrict__ rec, ghr_geom_header* __restrict__ hdr, int64_t capacity, const float* __restrict__ bg,
    const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
    const float* __restrict__ dL_ddepths, const float* __restrict__ dL_dalphas, float* __restrict__ accum,
    uint32_t id_mask /* 0x00FFFFFF when the list values carry block masks (wide_masks_packed) */)
{
    constexpr int BATCH = kWideBwdBatch, NSTAGE = kWideBwdStages, PER_LANE = BATCH / 32;
    __shared__ __align__(16) float4 s_all[kWideBwdWarps][NSTAGE][BATCH * 3];
    __shared__ uint32_t s_id_all[kWideBwdWarps][NSTAGE][BATCH];

    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float4(*s_rec)[BATCH * 3] = s_all[warp];
    uint32_t(*s_id)[BATCH] = s_id_all[warp];
    const uint32_t gx = (W + GHR_TILE - 1) / GHR_TILE;
    const bool overflow = (int64_t)hdr->num_rendered > capacity;
    const uint32_t* __restrict__ point_list = sh->sorted_sel ? vals_b : vals_a;
    const float bgc0 = bg[0], bgc1 = bg[1], bgc2 = bg[2];
    const float ddelx_dx = (float)(0.5 * W), ddely_dy = (float)(0.5 * H);
    const int slot = reduce_slot_w(lane);
    const bool writer = slot < 9 && !(lane & 1);
    const size_t HW = (size_t)H * W;
    const uint32_t n_items = (uint32_t)tiles * 8u;

    for (;;) {
        uint32_t item = 0;
        if (lane == 0) item = atomicAdd(&hdr->blend_bwd_ticket, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= n_items) break;
        const uint32_t tile = order[item >> 3], wb = item & 7u;  // same 8x4 pixel blocks as the forward
        const uint2 range = overflow ? make_uint2(0, 0) : ranges[tile];
        const int todo_total = (int)(range.y - range.x);
        if (todo_total == 0) break;  // the queue is sorted by length: everything after this is empty too
        const uint32_t tile_x = tile % gx, tile_y = tile / gx;
        const uint32_t bx0 = tile_x * GHR_TILE + (wb & 1u) * 8u, by0 = tile_y * GHR_TILE + (wb >> 1) * 4u;
        const uint32_t pixx = bx0 + (lane & 7u), pixy = by0 + (lane >> 3);
        const bool inside = pixx < (uint32_t)W && pixy < (uint32_t)H;
        const float pxf = (float)pixx, pyf = (float)pixy;
        const float rx0 = (float)bx0, rx1 = (float)(bx0 + 7u), ry0 = (float)by0, ry1 = (float)(by0 + 3u);

        const size_t pix = (size_t)pixy * W + pixx;
        const float T_final = inside ? final_Ts[pix] : 0.0f;
        float T = T_final;
        const int last_contributor = inside ? (int)n_contrib[pix] : 0;
        // the warp only walks the part of the list some pixel of it still needs: positions [0, warp_last)
        const int warp_last = __reduce_max_sync(0xffffffffu, last_contributor);
        const int walk = min(todo_total, warp_last);
        const int rounds = (walk + BATCH - 1) / BATCH;
        if (rounds == 0) continue;

        float dpix0 = 0.f, dpix1 = 0.f, dpix2 = 0.f, dL_ddepth = 0.f, dL_dalpha = 0.f;
        if (inside) {
            dpix0 = dL_dpixels[pix]; dpix1 = dL_dpixels[HW + pix]; dpix2 = dL_dpixels[2 * HW + pix];
            dL_ddepth = dL_ddepths[pix]; dL_dalpha = dL_dalphas[pix];
        }
        const float bg_dot = bgc0 * dpix0 + bgc1 * dpix1 + bgc2 * dpix2;
        float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;
        float accum_depth_rec = 0.f, accum_alpha_rec = 0.f, last_alpha = 0.f, last_depth = 0.f;

        // round r stages list positions walk-1-r*BATCH-slot (back to front).  Ids are fetched one round ahead of
        // the record copies (two dependent L2 round trips otherwise), records two rounds ahead of their use.
        uint32_t nid[PER_LANE];
        auto load_ids = [&](int round) {
#pragma unroll
            for (int u = 0; u < PER_LANE; u++) {
                const int progress = round * BATCH + (int)lane + u * 32;
                nid[u] = (round < rounds && progress < walk) ? (point_list[range.x + (uint32_t)(walk - 1 - progress)] & id_mask) : 0xFFFFFFFFu;
            }
        };
        auto stage = [&](int buf) {
#pragma unroll
            for (int u = 0; u < PER_LANE; u++) {
                const int sl = (int)lane + u * 32;
                if (nid[u] != 0xFFFFFFFFu) {
                    s_id[buf][sl] = nid[u];
                    const float4* src = rec + 3 * (size_t)nid[u];
                    float4* dst = &s_rec[buf][sl * 3];
                    cp_async16w(dst, src);
                    cp_async16w(dst + 1, src + 1);
                    cp_async16w(dst + 2, src + 2);
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };

        load_ids(0); stage(0);
        load_ids(1); stage(1);
        load_ids(2);
        int toDo = walk;
        for (int i = 0; i < rounds; i++, toDo -= BATCH) {
            const int buf = i % NSTAGE;
            stage((i + 2) % NSTAGE);
            load_ids(i + 3);
            asm volatile("cp.async.wait_group 2;" ::: "memory");
            __syncwarp();
            const int n = min(BATCH, toDo);
            const float4* r = s_rec[buf];
            // entry j of this round sits at list position pos(j) = walk-1-(i*BATCH+j); the reference's
            // `contributor` for it equals pos(j) (backward.cu:499-501: skip if contributor >= last_contributor)
            const int pos0 = walk - 1 - i * BATCH;
            for (int g = 0; g < n; g += 32) {
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
                // Survivors four at a time.  Phase 1 (independent: ILP) evaluates G/alpha of the four; phase 2
                // runs the transmittance / accumulated-colour recurrences in list order; phase 3 reduces the
                // 4 x 9 partial gradients with four interleaved butterflies.
                while (mask) {
                    int j[4];
                    bool v_ok[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        v_ok[u] = mask != 0;
                        j[u] = v_ok[u] ? g + __ffs(mask) - 1 : j[0];
                        mask &= mask - 1;
                    }
                    float G[4], alpha[4], dx[4], dy[4];
                    float4 r0[4], r1[4];
                    bool valid[4];
                    uint32_t vm[4];
    #pragma unroll
                    for (int u = 0; u < 4; u++) {
                        r0[u] = r[3 * j[u]]; r1[u] = r[3 * j[u] + 1];
                        dx[u] = __fsub_rn(r0[u].x, pxf); dy[u] = __fsub_rn(r0[u].y, pyf);
                        const float power = pair_power(dx[u], dy[u], r1[u].x, r1[u].y, r1[u].z);
                        G[u] = expf(power);
                        alpha[u] = fminf(0.99f, __fmul_rn(r1[u].w, G[u]));
                        valid[u] = v_ok[u] && (pos0 - j[u] < last_contributor) && !(power > 0.0f) && !(alpha[u] < 1.0f / 255.0f);
                        vm[u] = __ballot_sync(0xffffffffu, valid[u]);
                    }
                    if ((vm[0] | vm[1] | vm[2] | vm[3]) == 0) continue;
                    float v[4][9];
    #pragma unroll
                    for (int u = 0; u < 4; u++) {
    #pragma unroll
                        for (int k = 0; k < 9; k++) v[u][k] = 0.0f;
                        if (valid[u]) {
                            const float4 r2 = r[3 * j[u] + 2];
                            const float a = alpha[u];
                            T = T / (1.0f - a);
                            const float dchannel_dcolor = a * T;
                            float dL_dopa = 0.0f;
                            acc0 = last_alpha * lc0 + (1.0f - last_alpha) * acc0; lc0 = r2.x;
                            dL_dopa += (r2.x - acc0) * dpix0; v[u][0] = dchannel_dcolor * dpix0;
                            acc1 = last_alpha * lc1 + (1.0f - last_alpha) * acc1; lc1 = r2.y;
                            dL_dopa += (r2.y - acc1) * dpix1; v[u][1] = dchannel_dcolor * dpix1;
                            acc2 = last_alpha * lc2 + (1.0f - last_alpha) * acc2; lc2 = r2.z;
                            dL_dopa += (r2.z - acc2) * dpix2; v[u][2] = dchannel_dcolor * dpix2;
                            accum_depth_rec = last_alpha * last_depth + (1.0f - last_alpha) * accum_depth_rec;
                            last_depth = r0[u].z;
                            dL_dopa += (r0[u].z - accum_depth_rec) * dL_ddepth;
                            accum_alpha_rec = last_alpha + (1.0f - last_alpha) * accum_alpha_rec;
                            dL_dopa += (1.0f - accum_alpha_rec) * dL_dalpha;
                            dL_dopa *= T;
                            last_alpha = a;
                            dL_dopa += (-T_final / (1.0f - a)) * bg_dot;
                            const float dL_dG = r1[u].w * dL_dopa;
                            const float gdx = G[u] * dx[u], gdy = G[u] * dy[u];
                            const float dG_ddelx = -gdx * r1[u].x - gdy * r1[u].y;
                            const float dG_ddely = -gdy * r1[u].z - gdx * r1[u].y;
                            v[u][3] = dL_dG * dG_ddelx * ddelx_dx;
                            v[u][4] = dL_dG * dG_ddely * ddely_dy;
                            v[u][5] = -0.5f * gdx * dx[u] * dL_dG;
                            v[u][6] = -0.5f * gdx * dy[u] * dL_dG;
                            v[u][7] = -0.5f * gdy * dy[u] * dL_dG;
                            v[u][8] = G[u] * dL_dopa;
                        }
                    }
                    float total[4];
    #pragma unroll
                    for (int u = 0; u < 4; u++) total[u] = warp_reduce9_w(v[u], lane);
    #pragma unroll
                    for (int u = 0; u < 4; u++)
                        if (vm[u] != 0 && writer) atomicAdd(accum + 12 * (size_t)s_id[buf][j[u]] + slot, total[u]);
                }
            }
            __syncwarp();  // buffer `buf` is re-filled by the next iteration's stage()
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
    }
}

cudaError_t launch_blend_backward_wide(const ghr_raster_params& p, const char* geom, const char* binning, const char* img,
                                  const ghr_layout& L, const float* dL_dpix, const float* dL_ddepth,
                                  const float* dL_dalpha, float* accum, cudaStream_t s)
{

    const int tiles = ((p.W + GHR_TILE - 1) / GHR_TILE) * ((p.H + GHR_TILE - 1) / GHR_TILE);
    ghr_geom_header* hdr = reinterpret_cast<ghr_geom_header*>(const_cast<char*>(geom) + L.g_header);
    // the queue head is scratch state inside the (otherwise read-only) geometry buffer
    cudaError_t e = cudaMemsetAsync(&hdr->blend_bwd_ticket, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    const int64_t items = (int64_t)tiles * 8;
    int64_t ctas = (items + kWideBwdWarps - 1) / kWideBwdWarps;
    const int per_sm = blend_shared_gpu() ? kWideBwdCtasPerSm / 2 : kWideBwdCtasPerSm;  // see ghr_blend_fwd_wide.cu
    if (ctas > (int64_t)sm_count() * per_sm) ctas = (int64_t)sm_count() * per_sm;
    blend_backward_wide_kernel<<<(unsigned)ctas, kWideBwdWarps * 32, 0, s>>>(
        p.W, p.H, tiles, reinterpret_cast<const uint2*>(img + L.i_ranges),
        reinterpret_cast<const uint32_t*>(img + L.i_order), reinterpret_cast<const SortHeader*>(binning + L.b_header),
        reinterpret_cast<const uint32_t*>(binning + L.b_vals_a), reinterpret_cast<const uint32_t*>(binning + L.b_vals_b),
        reinterpret_cast<const float4*>(geom + L.g_rec), hdr, L.capacity, p.background,
        reinterpret_cast<const float*>(img + L.i_final_T), reinterpret_cast<const uint32_t*>(img + L.i_n_contrib), dL_dpix,
        dL_ddepth, dL_dalpha, accum, wide_masks_packed(p.P, L.capacity) ? 0x00FFFFFFu : 0xFFFFFFFFu);
    return cudaGetLastError();
}

}  // namespace ghr
