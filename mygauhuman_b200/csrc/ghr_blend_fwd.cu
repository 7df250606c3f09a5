// Per-tile front-to-back alpha blending with depth and alpha-mask outputs
// (reference forward.cu:261-383 renderCUDA).
//
// Work decomposition (ours): a 16x16 tile is covered by sixteen 4x4 pixel
// blocks; one (tile, block) pair is a work item processed by ONE warp,
// independently of every other warp (no block-level barrier anywhere).  Two
// lanes serve each pixel ("slots"): lane p and lane p+16 evaluate different
// Gaussians of the pixel's list at the same time.  The kernel is persistent:
// (SM count) x 8 single-warp CTAs pull items from a global queue ordered by
// decreasing tile-list length (tile_order_kernel), so the few heavy tiles of a
// human-shaped scene start first and spread over all SMs.  Per warp, list
// entries are staged in rounds of 256 (128 when several views share the GPU,
// ghr_set_shared_gpu) packed 48-byte records
//   {px,py,depth,radius | conic.xyz,opacity | r,g,b,cull-extent}
// with cp.async (LDGSTS) into a 2-deep private shared-memory ring (list entries
// are fetched one round ahead of the record copies), so the inner loop never
// touches global memory (the reference gathers colour and depth from global per
// contributing pair, forward.cu:360,362).
//
// Per warp and round: (1) only entries whose per-entry 4x4-block mask
// (tile_ranges_kernel: bounding box, then the exact ellipse-vs-block test) has
// this block's bit are staged; their bits form the round's survivor words.
// (2) Survivors are taken KSTEP = 8 at a time in list order, across the 32-entry
// words; slot 0 evaluates the even ones, slot 1 the odd ones (power/exp/alpha,
// independent -> ILP), the alphas are exchanged with one shuffle each, and both
// lanes run the transmittance recurrence over all eight.
// The five accumulators are split between the two lanes of a pixel (slot 0:
// red, green, alpha-mask; slot 1: blue, depth), each accumulated in list order.
// Skipped Gaussians are exactly those the reference's per-pixel tests
// (power > 0, alpha < 1/255) would skip for every pixel of the block, and the
// per-pair arithmetic is the reference's compiled sequence (pair_power, expf,
// fma accumulation in list order), so images, final_T and n_contrib are
// bit-exact.
//
// Bound: fp32 issue + MUFU.EX2 along the longest per-pixel list, not HBM
// (algorithmic bytes 44*N + 28*X).
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem)
{
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

#ifdef GHR_TRACE  // tuning builds only (tools/build_variant.sh): per-item begin/end timestamps and phase cycle counts
__device__ unsigned long long g_fwd_trace[1 << 16][8];
__device__ __forceinline__ unsigned long long gtime()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ uint32_t smid()
{
    uint32_t v;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(v));
    return v;
}
#endif

// Staging: 256-entry rounds in a 2-deep per-warp ring, one warp per CTA, 8 CTAs per SM.  Long rounds amortise the
// per-round cost and few co-resident warps leave the heavy items the issue slots (measured, tools/fwd_trace.py).
// Frames with millions of pairs take the other kernel family (ghr_blend_fwd_wide.cu), see launch_blend_forward.
#ifndef GHR_FWD_KSTEP
#define GHR_FWD_KSTEP 8
#endif
constexpr int KSTEP = GHR_FWD_KSTEP;  // survivors per step (half of them evaluated by each slot)
// NX > 0: NX more [P,3] attributes are blended in the same walk (multi-attribute rendering, include/ghr.h).  Their
// colours come packed as AttrLayout<NX>::XS floats per Gaussian (ghr_pack_attributes): the first half of a row holds the
// channels accumulated by slot 0, the second half those of slot 1, each channel with the arithmetic of the main colour
// channels (fmul alpha*c, fma with T, list order), so every attribute image equals a separate pass bit for bit.
template <int BATCH, int NSTAGE, int NX>
__host__ __device__ constexpr size_t blend_fwd_warp_smem()
{
    return ((sizeof(float4) * 3 + sizeof(float) * attr_row_floats<NX>()) * BATCH * NSTAGE +
            (sizeof(uint8_t) * BATCH + sizeof(uint32_t) * 4) * NSTAGE + 15) / 16 * 16;
}

// One 8-byte list entry {Gaussian id, block mask}, read-only path.  Always a single 64-bit load that nothing consumes
// until the next staging step: the loads of a whole round are in flight together and their latency hides behind the
// evaluation of the current round.  (Written as `const uint2 e = entries[k]; if ((e.y >> wb) & 1) id = e.x;` the
// compiler split every entry into two DEPENDENT 32-bit loads behind branches: sixteen exposed L2 round trips per
// 256-entry round, 41 % of the kernel's warp time in the per-item trace.)
__device__ __forceinline__ uint2 ld_entry(const uint2* p)
{
    uint2 v;
    asm volatile("ld.global.nc.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}

template <int BATCH, int NSTAGE, int WARPS, int NX>
__global__ void __launch_bounds__(WARPS * 32) blend_forward_kernel(
    int W, int H, int tiles, const uint2* __restrict__ ranges, const uint32_t* __restrict__ order,
    const SortHeader* __restrict__ sh, const uint2* __restrict__ ent_a, const uint2* __restrict__ ent_b,
    const float4* __restrict__ rec, ghr_geom_header* __restrict__ hdr, int64_t capacity, const float* __restrict__ bg,
    float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_alpha,
    float* __restrict__ final_T, uint32_t* __restrict__ n_contrib, const float* __restrict__ xcol,
    float* __restrict__ out_x)
{
    constexpr int PER_LANE = BATCH / 32;
    static_assert(BATCH <= 256, "staged positions are bytes");
    using AL = AttrLayout<(NX ? NX : 2)>;
    constexpr int XS = attr_row_floats<NX>(), XH = AL::XH, CHS = AL::CHS;
    // per-warp staging ring in dynamic shared memory.  The survivors of a round (entries whose block mask has this
    // block's bit) are staged COMPACTED, in list order: records (, attribute rows), then per staged record its position
    // inside the round (a byte) and per stage the survivor count.  The evaluation loop walks 0..count-1 and never looks
    // at a mask word.
    extern __shared__ __align__(16) unsigned char fwd_smem[];

    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t slot = lane >> 4, pl = lane & 15u;  // which of the pixel's two lanes; pixel inside the 4x4 block
    const uint32_t lt_mask = (1u << lane) - 1u;
    unsigned char* wbase = fwd_smem + warp * blend_fwd_warp_smem<BATCH, NSTAGE, NX>();
    float4(*s_rec)[BATCH * 3] = reinterpret_cast<float4(*)[BATCH * 3]>(wbase);
    float* s_x = reinterpret_cast<float*>(wbase + sizeof(float4) * 3 * BATCH * NSTAGE);  // [NSTAGE][BATCH][XS]
    uint8_t(*s_pos)[BATCH] = reinterpret_cast<uint8_t(*)[BATCH]>(
        wbase + (sizeof(float4) * 3 + sizeof(float) * XS) * BATCH * NSTAGE);
    uint32_t* s_cnt = reinterpret_cast<uint32_t*>(wbase + (sizeof(float4) * 3 + sizeof(float) * XS + 1) * BATCH * NSTAGE);
    const uint32_t gx = (W + GHR_TILE - 1) / GHR_TILE;
    const bool overflow = (int64_t)hdr->num_rendered > capacity;
    // {Gaussian id, 4x4-block mask} per list entry, written by tile_ranges_kernel into the key buffer the sort left unused
    const uint2* __restrict__ entries = sh->sorted_sel ? ent_a : ent_b;
    const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];
    const uint32_t n_items = (uint32_t)tiles * 16u;

    for (;;) {
        uint32_t item = 0;
        if (lane == 0) item = atomicAdd(&hdr->blend_fwd_ticket, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= n_items) break;
#ifdef GHR_TRACE
        const unsigned long long t_begin = gtime();
        int trace_rounds = 0, trace_surv = 0, trace_steps = 0, trace_seq = 0;
        long long c_wait = 0, c_cull = 0, c_eval = 0, c_mark = clock64();
#define TRACE_LAP(acc) { const long long now_ = clock64(); acc += now_ - c_mark; c_mark = now_; }
#else
#define TRACE_LAP(acc)
#endif
        const uint32_t tile = order[item >> 4], wb = item & 15u;  // which of the tile's sixteen 4x4 blocks
        const uint32_t tile_x = tile % gx, tile_y = tile / gx;
        const uint32_t bx0 = tile_x * GHR_TILE + (wb & 3u) * 4u, by0 = tile_y * GHR_TILE + (wb >> 2) * 4u;
        const uint32_t pixx = bx0 + (pl & 3u), pixy = by0 + (pl >> 2);
        const bool inside = pixx < (uint32_t)W && pixy < (uint32_t)H;
        const float pxf = (float)pixx, pyf = (float)pixy;
        const uint2 range = overflow ? make_uint2(0, 0) : ranges[tile];
        const int todo_total = (int)(range.y - range.x);
        const int rounds = (todo_total + BATCH - 1) / BATCH;

        // Staging pipeline.  A record gather is two dependent L2 round trips (list entry -> record); the entries of a
        // round are therefore loaded into registers one round before their records are copied -- issued only, see
        // ld_entry.  Only the records of Gaussians whose block mask has this block's bit are copied, compacted.
        uint2 ent[PER_LANE];
        auto load_ids = [&](int round) {
#pragma unroll
            for (int u = 0; u < PER_LANE; u++) {
                const uint32_t k = range.x + (uint32_t)round * BATCH + lane + u * 32;
                ent[u] = ld_entry(entries + min(k, range.y - 1u));  // (range.y > range.x whenever this runs)
            }
        };
        auto stage = [&](int buf, int round) {  // copies the surviving records of the entries held in ent[] (= round's)
            uint32_t total = 0;
#pragma unroll
            for (int u = 0; u < PER_LANE; u++) {
                const uint32_t k = range.x + (uint32_t)round * BATCH + lane + u * 32;
                const bool keep = round < rounds && k < range.y && ((ent[u].y >> wb) & 1u);
                const uint32_t bits = __ballot_sync(0xffffffffu, keep);
                if (keep) {
                    const uint32_t pos = total + __popc(bits & lt_mask);
                    const float4* src = rec + 3 * (size_t)ent[u].x;
                    float4* dst = &s_rec[buf][pos * 3];
                    cp_async16(dst, src);
                    cp_async16(dst + 1, src + 1);
                    cp_async16(dst + 2, src + 2);
                    if constexpr (NX > 0) {
                        const float* xsrc = xcol + (size_t)ent[u].x * XS;
                        float* xdst = s_x + ((size_t)buf * BATCH + pos) * XS;
#pragma unroll
                        for (int c = 0; c < XS; c += 4) cp_async16(xdst + c, xsrc + c);
                    }
                    s_pos[buf][pos] = (uint8_t)(lane + u * 32);
                }
                total += __popc(bits);
            }
            if (lane == 0) s_cnt[buf] = total;
            cp_async_commit();
        };

        // Both lanes of a pixel carry T / done / last_contributor (identical); A0, A1, A2 are this lane's share of the
        // accumulators: slot 0 -> red, green, alpha-mask; slot 1 -> blue, depth, (unused).
        bool done = !inside;
        float T = 1.0f, A0 = 0.f, A1 = 0.f, A2 = 0.f;
        uint32_t last_contributor = 0;
        float X[CHS];  // NX > 0: this lane's half of the attribute channels
#pragma unroll
        for (int c = 0; c < CHS; c++) X[c] = 0.f;

        if (rounds > 0) {
            // prologue: all list entries of the first NSTAGE-1 rounds are requested before the first record copy
            uint2 pre[NSTAGE - 1][PER_LANE];
#pragma unroll
            for (int q = 0; q < NSTAGE - 1; q++) {
                load_ids(q);
#pragma unroll
                for (int u = 0; u < PER_LANE; u++) pre[q][u] = ent[u];
            }
#pragma unroll
            for (int q = 0; q < NSTAGE - 1; q++) {
#pragma unroll
                for (int u = 0; u < PER_LANE; u++) ent[u] = pre[q][u];
                stage(q, q);
            }
            load_ids(NSTAGE - 1);
        }
        for (int i = 0; i < rounds; i++) {
            const int buf = i % NSTAGE;
            stage((i + NSTAGE - 1) % NSTAGE, i + NSTAGE - 1);  // records of round i+NSTAGE-1 (an empty group past the end keeps the count uniform)
            if (i + NSTAGE < rounds) load_ids(i + NSTAGE);
            cp_async_wait<NSTAGE - 1>();       // this lane's copies for rounds <= i have landed
            __syncwarp();             // ... and everybody else's
            if (__all_sync(0xffffffffu, done)) break;  // forward.cu:313-315, per block instead of per tile
#ifdef GHR_TRACE
            trace_rounds++;
#endif
            TRACE_LAP(c_wait)
            const float4* r = s_rec[buf];
            const float* rf = reinterpret_cast<const float*>(r);
            const uint32_t base = (uint32_t)i * BATCH;  // list position of the round's first entry
            const int n = (int)s_cnt[buf];               // survivors of the round, staged contiguously in list order
            int last_k = -1;                             // staged index of the last contributor inside this round
            // Survivors are consumed KSTEP at a time in list order: even ones are evaluated by slot 0, odd ones by slot 1.
            for (int s0 = 0; s0 < n; s0 += KSTEP) {
#ifdef GHR_TRACE
                trace_steps++;
                trace_surv += min(KSTEP, n - s0);
#endif
                float mine[KSTEP / 2];
#pragma unroll
                for (int u = 0; u < KSTEP / 2; u++) {
                    const int jq = s0 + 2 * u + (int)slot;
                    const bool vm = jq < n;
                    const int jm = vm ? jq : s0;
                    const float4 r0 = r[3 * jm], r1 = r[3 * jm + 1];
                    const float dx = __fsub_rn(r0.x, pxf), dy = __fsub_rn(r0.y, pyf);
                    const float power = pair_power(dx, dy, r1.x, r1.y, r1.z);
                    const float al = fminf(0.99f, __fmul_rn(r1.w, expf(power)));
                    mine[u] = (vm && !(power > 0.0f) && !(al < 1.0f / 255.0f)) ? al : 0.0f;  // 0 = skipped
                }
                float alpha[KSTEP];
#pragma unroll
                for (int u = 0; u < KSTEP / 2; u++) {
                    const float theirs = __shfl_xor_sync(0xffffffffu, mine[u], 16);
                    alpha[2 * u] = slot ? theirs : mine[u];
                    alpha[2 * u + 1] = slot ? mine[u] : theirs;
                }
                // Sequential part, straight-line and branch-free: only T -> test_T -> T is a dependent chain
                // (one multiply and one compare per Gaussian); everything else is computed up front.
                float x0[KSTEP], x1[KSTEP], om[KSTEP];
#pragma unroll
                for (int u = 0; u < KSTEP; u++) {
                    const int ju = min(s0 + u, n - 1);
                    const float4 r2 = r[3 * ju + 2];
                    const float depth = rf[12 * ju + 2];
                    x0[u] = __fmul_rn(alpha[u], slot ? r2.z : r2.x);
                    x1[u] = __fmul_rn(alpha[u], slot ? depth : r2.y);
                    om[u] = __fsub_rn(1.0f, alpha[u]);
                }
#pragma unroll
                for (int u = 0; u < KSTEP; u++) {
                    const bool act = alpha[u] > 0.0f && !done;
                    const float test_T = __fmul_rn(T, om[u]);
                    const bool stop = act && test_T < 0.0001f;
                    const bool acc = act && !stop;
                    done = done || stop;
                    A0 = acc ? __fmaf_rn(T, x0[u], A0) : A0;
                    A1 = acc ? __fmaf_rn(T, x1[u], A1) : A1;
                    A2 = acc ? __fmaf_rn(T, alpha[u], A2) : A2;
                    if constexpr (NX > 0) {
                        const int ju = min(s0 + u, n - 1);
                        const float2* cx = reinterpret_cast<const float2*>(s_x + ((size_t)buf * BATCH + ju) * XS + slot * XH);
#pragma unroll
                        for (int c = 0; c < CHS; c += 2) {
                            const float2 cc = cx[c >> 1];
                            X[c] = acc ? __fmaf_rn(T, __fmul_rn(alpha[u], cc.x), X[c]) : X[c];
                            if (c + 1 < CHS) X[c + 1] = acc ? __fmaf_rn(T, __fmul_rn(alpha[u], cc.y), X[c + 1]) : X[c + 1];
                        }
                    }
                    last_k = acc ? s0 + u : last_k;
                    T = acc ? test_T : T;
                }
                TRACE_LAP(c_eval)
                if (__all_sync(0xffffffffu, done)) break;
            }
            if (last_k >= 0) last_contributor = base + (uint32_t)s_pos[buf][last_k] + 1u;
            __syncwarp();  // buffer `buf` is re-filled by the stage() of the next iteration (round i+NSTAGE)
        }
        cp_async_wait<0>();
        __syncwarp();

        if (inside) {
            const size_t pix = (size_t)pixy * W + pixx, HW = (size_t)H * W;
            if (slot == 0) {
                final_T[pix] = T;
                n_contrib[pix] = last_contributor;
                out_color[pix] = __fmaf_rn(T, bg0, A0);
                out_color[HW + pix] = __fmaf_rn(T, bg1, A1);
                out_alpha[pix] = A2;
            } else {
                out_color[2 * HW + pix] = __fmaf_rn(T, bg2, A0);
                out_depth[pix] = A1;
            }
            if constexpr (NX > 0) {
#pragma unroll
                for (int c = 0; c < CHS; c++) {
                    const int ch = (int)slot * CHS + c;  // channel ch = colour (ch % 3) of attribute ch / 3
                    const float b = (ch % 3 == 0) ? bg0 : (ch % 3 == 1) ? bg1 : bg2;
                    out_x[(size_t)ch * HW + pix] = __fmaf_rn(T, b, X[c]);
                }
            }
        }
#ifdef GHR_TRACE
        if (lane == 0 && item < (1u << 16)) {
            g_fwd_trace[item][0] = t_begin;
            g_fwd_trace[item][1] = gtime();
            g_fwd_trace[item][2] = ((unsigned long long)smid() << 32) | (uint32_t)todo_total;
            g_fwd_trace[item][3] = (unsigned long long)trace_rounds;
            g_fwd_trace[item][4] = ((unsigned long long)trace_surv << 32) | (uint32_t)trace_steps;
            g_fwd_trace[item][5] = (unsigned long long)c_wait;
            g_fwd_trace[item][6] = (unsigned long long)c_cull;
            g_fwd_trace[item][7] = (unsigned long long)c_eval;
        }
#endif
    }
}
#ifdef GHR_TRACE
}  // namespace ghr
extern "C" int ghr_debug_fwd_trace(void* out, size_t bytes)
{
    return (int)cudaMemcpyFromSymbol(out, ghr::g_fwd_trace, bytes);
}
namespace ghr {
#endif

template <int BATCH, int NSTAGE, int WARPS, int NX>
static cudaError_t launch_fwd_config(int ctas_per_sm, const ghr_raster_params& p, const char* geom, const char* binning,
                                     char* img, const ghr_layout& L, float* out_color, float* out_depth, float* out_alpha,
                                     const float* xcol, float* out_x, cudaStream_t s)
{
    const int tiles = ((p.W + GHR_TILE - 1) / GHR_TILE) * ((p.H + GHR_TILE - 1) / GHR_TILE);
    const int64_t items = (int64_t)tiles * 16;
    int64_t ctas = (items + WARPS - 1) / WARPS;
    if (ctas > (int64_t)sm_count() * ctas_per_sm) ctas = (int64_t)sm_count() * ctas_per_sm;
    ghr_geom_header* hdr = reinterpret_cast<ghr_geom_header*>(const_cast<char*>(geom) + L.g_header);
    auto kernel = blend_forward_kernel<BATCH, NSTAGE, WARPS, NX>;
    const size_t smem = blend_fwd_warp_smem<BATCH, NSTAGE, NX>() * WARPS;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kernel<<<(unsigned)ctas, WARPS * 32, smem, s>>>(
        p.W, p.H, tiles, reinterpret_cast<const uint2*>(img + L.i_ranges),
        reinterpret_cast<const uint32_t*>(img + L.i_order), reinterpret_cast<const SortHeader*>(binning + L.b_header),
        reinterpret_cast<const uint2*>(binning + L.b_keys_a), reinterpret_cast<const uint2*>(binning + L.b_keys_b),
        reinterpret_cast<const float4*>(geom + L.g_rec), hdr, L.capacity, p.background, out_color,
        out_depth, out_alpha, reinterpret_cast<float*>(img + L.i_final_T), reinterpret_cast<uint32_t*>(img + L.i_n_contrib),
        xcol, out_x);
    return cudaGetLastError();
}

cudaError_t launch_blend_forward(const ghr_raster_params& p, const char* geom, const char* binning, char* img,
                                 const ghr_layout& L, float* out_color, float* out_depth, float* out_alpha,
                                 cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_BLEND_FWD, s);
    count_launch();
    if (blend_throughput_regime(L.capacity))
        return launch_blend_forward_wide(p, geom, binning, img, L, out_color, out_depth, out_alpha, s);
    ghr_geom_header* hdr = reinterpret_cast<ghr_geom_header*>(const_cast<char*>(geom) + L.g_header);
    cudaError_t e = cudaMemsetAsync(&hdr->blend_fwd_ticket, 0, sizeof(uint32_t), s);  // render may run twice per preprocess
    if (e != cudaSuccess) return e;
    if (blend_shared_gpu())  // several views in flight (MultiLaneRenderStep): half the staging per warp, so that the
                             // blend kernels of two views fit an SM together
        return launch_fwd_config<128, 2, 1, 0>(8, p, geom, binning, img, L, out_color, out_depth, out_alpha, nullptr, nullptr, s);
#ifdef GHR_FWD_BATCH  // tuning builds only
    return launch_fwd_config<GHR_FWD_BATCH, GHR_FWD_NSTAGE, 1, 0>(GHR_FWD_CTAS, p, geom, binning, img, L, out_color, out_depth, out_alpha, nullptr, nullptr, s);
#else
    return launch_fwd_config<256, 2, 1, 0>(8, p, geom, binning, img, L, out_color, out_depth, out_alpha, nullptr, nullptr, s);
#endif
}

// Main image + n_attr (2, 4 or 6) packed attributes in one walk; latency regime only (ghr_api.cu checks).
cudaError_t launch_blend_forward_multi(const ghr_raster_params& p, const char* geom, const char* binning, char* img,
                                       const ghr_layout& L, float* out_color, float* out_depth, float* out_alpha,
                                       int n_attr, const float* xcol, float* out_x, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_BLEND_FWD, s);
    count_launch();
    ghr_geom_header* hdr = reinterpret_cast<ghr_geom_header*>(const_cast<char*>(geom) + L.g_header);
    cudaError_t e = cudaMemsetAsync(&hdr->blend_fwd_ticket, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    switch (n_attr) {
    case 2: return launch_fwd_config<128, 2, 1, 2>(8, p, geom, binning, img, L, out_color, out_depth, out_alpha, xcol, out_x, s);
    case 4: return launch_fwd_config<128, 2, 1, 4>(8, p, geom, binning, img, L, out_color, out_depth, out_alpha, xcol, out_x, s);
#ifdef GHR_MF_BATCH  // tuning builds only
    case 6: return launch_fwd_config<GHR_MF_BATCH, GHR_MF_NSTAGE, 1, 6>(GHR_MF_CTAS, p, geom, binning, img, L, out_color, out_depth, out_alpha, xcol, out_x, s);
#else
    case 6: return launch_fwd_config<96, 2, 1, 6>(8, p, geom, binning, img, L, out_color, out_depth, out_alpha, xcol, out_x, s);
#endif
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace ghr
