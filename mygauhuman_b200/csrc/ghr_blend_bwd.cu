// Backward of the per-tile blend (reference backward.cu:399-587 renderCUDA).
//
// The reference issues 9 fp32 global atomicAdds per contributing
// (pixel, Gaussian) pair (backward.cu:538,574-584) -- its hottest spot.  Here a
// work item is a (tile, 4x4 pixel block) pair processed by one warp, and two
// lanes serve each pixel ("slots", as in ghr_blend_fwd.cu): the warp walks the
// tile list back to front eight surviving Gaussians at a time; slot 0 owns the
// 1st, 3rd, 5th, 7th of them and slot 1 the others.  A lane evaluates
// G / alpha only for the Gaussians it owns, the alphas are exchanged with one
// shuffle each, both lanes run the (cheap, strictly sequential) transmittance /
// accumulated-colour recurrences over all eight, and each lane computes the 9
// partial gradients of its own four.  These are reduced over the 16 pixels of
// the slot with a split butterfly (11 shuffles for 9 values: at each xor step
// half of the lanes keeps half of the values) -- the two slots reduce two
// different Gaussians in the same instructions -- and 9 lanes per slot issue
// one atomic each into a packed per-Gaussian accumulator [P][12] (one 48-byte
// record -> the 9 atomics of a Gaussian hit at most two 32-byte sectors).  The
// per-Gaussian preprocess backward reads that record, so no separate
// dL_dconic/dL_dmean2D arrays are zeroed.
// Why slots: a warp alone issues ~0.4 instructions per cycle on this dependent
// code, and the kernel time is the instruction count along the heaviest item's
// list (measured: all items start at t=0, the longest one ends the kernel).
// Two slots halve the instructions per Gaussian on that path.
//
// accum record: 0-2 dL_dcolor, 3-4 dL_dmean2D.xy, 5-7 dL_dconic (x,y,w),
//               8 dL_dopacity, 9-11 unused.
// Bound: fp32 issue / MUFU / SHFL, not HBM (44*N + 28*X + 36*V bytes).
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

__device__ __forceinline__ void cp_async16b(void* smem, const void* gmem)
{
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}

// Sum v[0..8] over the 16 lanes of a slot (xor 8,4,2,1 never leaves the half-warp).  Afterwards every lane holds the
// total of exactly one value, half_slot(lane) (>= 9: a padding position, holds 0).
__device__ __forceinline__ float half_reduce9(const float (&v)[9], uint32_t lane)
{
    const bool b3 = lane & 8, b2 = lane & 4, b1 = lane & 2, b0 = lane & 1;
    float a[6];
#pragma unroll
    for (int k = 0; k < 5; k++) {
        const float lo = v[k], hi = (k < 4) ? v[5 + k] : 0.0f;
        const float keep = b3 ? hi : lo, send = b3 ? lo : hi;
        a[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    a[5] = 0.0f;
    float c[4];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float keep = b2 ? a[3 + k] : a[k], send = b2 ? a[k] : a[3 + k];
        c[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    c[3] = 0.0f;
    float d[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const float keep = b1 ? c[2 + k] : c[k], send = b1 ? c[k] : c[2 + k];
        d[k] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    const float keep = b0 ? d[1] : d[0], send = b0 ? d[0] : d[1];
    return keep + __shfl_xor_sync(0xffffffffu, send, 1);
}

// Which of the 9 values a lane holds after half_reduce9 (>= 9: none).
__device__ __forceinline__ int half_slot(uint32_t lane)
{
    const int w4 = ((lane & 2) ? 2 : 0) + ((lane & 1) ? 1 : 0);  // position inside the padded group of 4
    if (w4 >= 3) return 99;
    const int local = ((lane & 4) ? 3 : 0) + w4;                 // position inside the padded group of 6
    if (lane & 8) return local < 4 ? 5 + local : 99;
    return local < 5 ? local : 99;
}

// Split butterfly for N values over the 16 lanes of a slot: at each xor step (8, 4, 2, 1) a lane keeps one half of its
// values and sends the other, so N values cost about N shuffles instead of 4N.  Afterwards a lane holds the totals of
// (at most) two of the values: out[0], out[1] are the totals, reduce_index<N>(lane, t) says of which value (-1: padding).
template <int N>
__device__ __forceinline__ void half_reduce_n(const float (&v)[N], uint32_t lane, float (&out)[2])
{
    constexpr int N1 = (N + 1) / 2, N2 = (N1 + 1) / 2, N3 = (N2 + 1) / 2, N4 = (N3 + 1) / 2;
    static_assert(N4 <= 2, "at most two totals per lane");
    const bool b3 = lane & 8, b2 = lane & 4, b1 = lane & 2, b0 = lane & 1;
    float a[N1 + 1], c[N2 + 1], d[N3 + 1], e[N4 + 1];
#pragma unroll
    for (int k = 0; k < N1; k++) {
        const float lo = v[k], hi = (N1 + k < N) ? v[(N1 + k < N) ? N1 + k : 0] : 0.0f;
        a[k] = (b3 ? hi : lo) + __shfl_xor_sync(0xffffffffu, b3 ? lo : hi, 8);
    }
#pragma unroll
    for (int k = 0; k < N2; k++) {
        const float lo = a[k], hi = (N2 + k < N1) ? a[(N2 + k < N1) ? N2 + k : 0] : 0.0f;
        c[k] = (b2 ? hi : lo) + __shfl_xor_sync(0xffffffffu, b2 ? lo : hi, 4);
    }
#pragma unroll
    for (int k = 0; k < N3; k++) {
        const float lo = c[k], hi = (N3 + k < N2) ? c[(N3 + k < N2) ? N3 + k : 0] : 0.0f;
        d[k] = (b1 ? hi : lo) + __shfl_xor_sync(0xffffffffu, b1 ? lo : hi, 2);
    }
#pragma unroll
    for (int k = 0; k < N4; k++) {
        const float lo = d[k], hi = (N4 + k < N3) ? d[(N4 + k < N3) ? N4 + k : 0] : 0.0f;
        e[k] = (b0 ? hi : lo) + __shfl_xor_sync(0xffffffffu, b0 ? lo : hi, 1);
    }
    out[0] = e[0];
    out[1] = N4 > 1 ? e[N4 > 1 ? 1 : 0] : 0.0f;
}

template <int N>
__device__ __forceinline__ int reduce_index(uint32_t lane, int t)
{
    constexpr int N1 = (N + 1) / 2, N2 = (N1 + 1) / 2, N3 = (N2 + 1) / 2, N4 = (N3 + 1) / 2;
    if (t >= N4) return -1;
    const int i4 = ((lane & 1) ? N4 : 0) + t;
    if (i4 >= N3) return -1;
    const int i3 = ((lane & 2) ? N3 : 0) + i4;
    if (i3 >= N2) return -1;
    const int i2 = ((lane & 4) ? N2 : 0) + i3;
    if (i2 >= N1) return -1;
    const int i1 = ((lane & 8) ? N1 : 0) + i2;
    return i1 < N ? i1 : -1;
}

#ifdef GHR_TRACE  // tuning builds only (tools/build_variant.sh): per-item begin/end timestamps and phase cycle counts
__device__ unsigned long long g_bwd_trace[1 << 16][8];
__device__ __forceinline__ unsigned long long gtime_b()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define BTRACE_LAP(acc) { const long long now_ = clock64(); acc += now_ - c_mark; c_mark = now_; }
#else
#define BTRACE_LAP(acc)
#endif
constexpr int kBwdWarps = 4;  // independent warps per CTA
// Staging as in ghr_blend_fwd.cu: 256-entry rounds, 2-deep ring, 8 warps per SM; frames with millions of pairs take
// ghr_blend_bwd_wide.cu instead (launch_blend_backward).
// NX > 0 (multi-attribute rendering, include/ghr.h): NX more [P,3] attributes were blended by the forward; their packed
// colour rows (AttrLayout<NX>) are staged next to the records.  The two lanes of a pixel split the attribute channels
// (slot 0 the first half of a row, slot 1 the second): each runs the accumulated-colour recurrence of its channels over
// BOTH Gaussians of a pair, the partial sums of (c - accum) * dL_dpixel are exchanged with one shuffle and join the
// main channels' dL_dalpha, and the 2 * CHS colour gradients of the pair go through half_reduce_n into xaccum (same row
// layout as the colours).
template <int BATCH, int NSTAGE, int NX>
__host__ __device__ constexpr size_t blend_bwd_warp_smem()
{
    return ((sizeof(float4) * 3 + sizeof(uint32_t) + sizeof(float) * attr_row_floats<NX>()) * NSTAGE * BATCH +
            (sizeof(uint8_t) * BATCH + sizeof(uint32_t) * 4) * NSTAGE + 15) / 16 * 16;
}

// see ghr_blend_fwd.cu: one 64-bit load per list entry, consumed only at the next staging step
__device__ __forceinline__ uint2 ld_entry_b(const uint2* p)
{
    uint2 v;
    asm volatile("ld.global.nc.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}

// Persistent kernel: warps pull (tile, 4x4 block) work items from a queue ordered by decreasing tile-list
// length (tile_order_kernel) and work independently (no block-level barrier); see ghr_blend_fwd.cu.
template <int BATCH, int NSTAGE, int NX>
__global__ void __launch_bounds__(kBwdWarps * 32) blend_backward_kernel(
    int W, int H, int tiles, const uint2* __restrict__ ranges, const uint32_t* __restrict__ order,
    const SortHeader* __restrict__ sh, const uint2* __restrict__ ent_a, const uint2* __restrict__ ent_b,
    const float4* __restrict__ rec, ghr_geom_header* __restrict__ hdr, int64_t capacity, const float* __restrict__ bg,
    const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
    const float* __restrict__ dL_ddepths, const float* __restrict__ dL_dalphas, float* __restrict__ accum,
    const float* __restrict__ xcol, const float* __restrict__ xdpix, float* __restrict__ xaccum)
{
    constexpr int PER_LANE = BATCH / 32;
    static_assert(BATCH <= 256, "staged positions are bytes");
    using AL = AttrLayout<(NX ? NX : 2)>;
    constexpr int XS = attr_row_floats<NX>(), XH = AL::XH, CHS = AL::CHS;
    // per-warp staging ring in dynamic shared memory; the survivors of a round are staged COMPACTED in walk order
    // (ghr_blend_fwd.cu): records, ids, per staged record its position inside the round (a byte), per stage the
    // survivor count(, attribute rows)
    extern __shared__ __align__(16) unsigned char bwd_smem[];
    constexpr size_t kWarpBytes = blend_bwd_warp_smem<BATCH, NSTAGE, NX>();
    static_assert(kWarpBytes % 16 == 0, "per-warp staging area must keep float4 alignment");
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t slot = lane >> 4, pl = lane & 15u;  // which of the pixel's two lanes; pixel inside the 4x4 block
    const uint32_t lt_mask = (1u << lane) - 1u;
    unsigned char* wbase = bwd_smem + warp * kWarpBytes;
    float4(*s_rec)[BATCH * 3] = reinterpret_cast<float4(*)[BATCH * 3]>(wbase);
    uint32_t(*s_id)[BATCH] = reinterpret_cast<uint32_t(*)[BATCH]>(wbase + sizeof(float4) * NSTAGE * BATCH * 3);
    float* s_x = reinterpret_cast<float*>(wbase + (sizeof(float4) * 3 + sizeof(uint32_t)) * NSTAGE * BATCH);  // [NSTAGE][BATCH][XS]
    uint8_t(*s_pos)[BATCH] = reinterpret_cast<uint8_t(*)[BATCH]>(
        wbase + (sizeof(float4) * 3 + sizeof(uint32_t) + sizeof(float) * XS) * NSTAGE * BATCH);
    uint32_t* s_cnt = reinterpret_cast<uint32_t*>(
        wbase + (sizeof(float4) * 3 + sizeof(uint32_t) + sizeof(float) * XS + 1) * NSTAGE * BATCH);
    int xidx[2];  // which of the pair's 2*CHS attribute colour gradients this lane holds after half_reduce_n
    xidx[0] = reduce_index<2 * CHS>(lane, 0);
    xidx[1] = reduce_index<2 * CHS>(lane, 1);
    const uint32_t gx = (W + GHR_TILE - 1) / GHR_TILE;
    const bool overflow = (int64_t)hdr->num_rendered > capacity;
    // {Gaussian id, 4x4-block mask} per list entry (tile_ranges_kernel; lives in the key buffer the sort left unused)
    const uint2* __restrict__ entries = sh->sorted_sel ? ent_a : ent_b;
    const float bgc0 = bg[0], bgc1 = bg[1], bgc2 = bg[2];
    const float ddelx_dx = (float)(0.5 * W), ddely_dy = (float)(0.5 * H);
    const int vslot = half_slot(lane);  // which of the 9 reduced values this lane ends up holding
    const bool writer = vslot < 9;
    const uint32_t half_mask = slot ? 0xFFFF0000u : 0x0000FFFFu;
    const size_t HW = (size_t)H * W;
    const uint32_t n_items = (uint32_t)tiles * 16u;

    for (;;) {
        uint32_t item = 0;
        if (lane == 0) item = atomicAdd(&hdr->blend_bwd_ticket, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= n_items) break;
#ifdef GHR_TRACE
        const unsigned long long t_begin = gtime_b();
        int trace_rounds = 0, trace_steps = 0, trace_live = 0;
        long long c_wait = 0, c_p1 = 0, c_p23 = 0, c_mark = clock64();
#endif
        const uint32_t tile = order[item >> 4], wb = item & 15u;  // same 4x4 pixel blocks as the forward
        const uint2 range = overflow ? make_uint2(0, 0) : ranges[tile];
        const int todo_total = (int)(range.y - range.x);
        if (todo_total == 0) break;  // the queue is sorted by length: everything after this is empty too
        const uint32_t tile_x = tile % gx, tile_y = tile / gx;
        const uint32_t bx0 = tile_x * GHR_TILE + (wb & 3u) * 4u, by0 = tile_y * GHR_TILE + (wb >> 2) * 4u;
        const uint32_t pixx = bx0 + (pl & 3u), pixy = by0 + (pl >> 2);
        const bool inside = pixx < (uint32_t)W && pixy < (uint32_t)H;
        const float pxf = (float)pixx, pyf = (float)pixy;

        const size_t pix = (size_t)pixy * W + pixx;
        const float T_final = inside ? final_Ts[pix] : 0.0f;
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
        float bg_dot = bgc0 * dpix0 + bgc1 * dpix1 + bgc2 * dpix2;
        float dpx[CHS], accx[CHS], lcx[CHS];  // NX > 0: this lane's half of the attribute channels
        if constexpr (NX > 0) {
            float part = 0.f;
#pragma unroll
            for (int c = 0; c < CHS; c++) {
                const int ch = (int)slot * CHS + c;
                dpx[c] = inside ? xdpix[(size_t)ch * HW + pix] : 0.f;
                accx[c] = 0.f; lcx[c] = 0.f;
                part = fmaf((ch % 3 == 0) ? bgc0 : (ch % 3 == 1) ? bgc1 : bgc2, dpx[c], part);
            }
            bg_dot += part + __shfl_xor_sync(0xffffffffu, part, 16);
        }
        // per-pixel recurrence state, carried identically by both lanes of the pixel
        float T = T_final;
        float Bacc = T_final * bg_dot;  // NX == 0: the scalar buffer B of phase 2, seeded with the background term
        float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;
        float accum_depth_rec = 0.f, accum_alpha_rec = 0.f, last_alpha = 0.f, last_depth = 0.f;

        // round r stages list positions walk-1-r*BATCH-slot (back to front).  Entries are fetched one round ahead of
        // the record copies (two dependent L2 round trips otherwise), records two rounds ahead of their use; only the
        // records of Gaussians whose block mask has this block's bit are copied.
        uint2 ent[PER_LANE];
        auto load_ids = [&](int round) {  // issue only (ld_entry_b)
#pragma unroll
            for (int u = 0; u < PER_LANE; u++) {
                const int progress = min(round * BATCH + (int)lane + u * 32, walk - 1);
                ent[u] = ld_entry_b(entries + range.x + (uint32_t)(walk - 1 - progress));
            }
        };
        auto stage = [&](int buf, int round) {
            uint32_t total = 0;
#pragma unroll
            for (int u = 0; u < PER_LANE; u++) {
                const int sl = (int)lane + u * 32;
                const bool keep = round < rounds && round * BATCH + sl < walk && ((ent[u].y >> wb) & 1u);
                const uint32_t bits = __ballot_sync(0xffffffffu, keep);
                if (keep) {
                    const uint32_t pos = total + __popc(bits & lt_mask);
                    s_id[buf][pos] = ent[u].x;
                    s_pos[buf][pos] = (uint8_t)sl;
                    const float4* src = rec + 3 * (size_t)ent[u].x;
                    float4* dst = &s_rec[buf][pos * 3];
                    cp_async16b(dst, src);
                    cp_async16b(dst + 1, src + 1);
                    cp_async16b(dst + 2, src + 2);
                    if constexpr (NX > 0) {
                        const float* xsrc = xcol + (size_t)ent[u].x * XS;
                        float* xdst = s_x + ((size_t)buf * BATCH + pos) * XS;
#pragma unroll
                        for (int c = 0; c < XS; c += 4) cp_async16b(xdst + c, xsrc + c);
                    }
                }
                total += __popc(bits);
            }
            if (lane == 0) s_cnt[buf] = total;
            asm volatile("cp.async.commit_group;" ::: "memory");
        };

        {
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
            stage((i + NSTAGE - 1) % NSTAGE, i + NSTAGE - 1);
            if (i + NSTAGE < rounds) load_ids(i + NSTAGE);
            asm volatile("cp.async.wait_group %0;" ::"n"(NSTAGE - 1) : "memory");
            __syncwarp();
            const float4* r = s_rec[buf];
            const float* rf = reinterpret_cast<const float*>(r);
#ifdef GHR_TRACE
            trace_rounds++;
#endif
            BTRACE_LAP(c_wait)
            // staged record j of this round came from round entry s_pos[j], i.e. list position
            // pos0 - s_pos[j]; the reference's `contributor` for it equals that position (backward.cu:499-501: skip if
            // contributor >= last_contributor)
            const int pos0 = walk - 1 - i * BATCH;
            const int n = (int)s_cnt[buf];  // survivors of the round, contiguous, in walk order
            for (int s0 = 0; s0 < n; s0 += 8) {
                {
                    // eight survivors in walk order; slot 0 owns j[0] j[2] j[4] j[6], slot 1 the others
                    int j[8];
                    bool v_ok[8];
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        v_ok[u] = s0 + u < n;
                        j[u] = v_ok[u] ? s0 + u : s0;
                    }
                    // phase 1 (independent: ILP): G / alpha of the four Gaussians this lane owns; alpha 0 = not valid here
                    float G[4], al[4], dx[4], dy[4], depth_own[4];
                    float4 con[4];
                    int jm[4];
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        jm[q] = slot ? j[2 * q + 1] : j[2 * q];
                        const bool ok = slot ? v_ok[2 * q + 1] : v_ok[2 * q];
                        const float4 r0 = r[3 * jm[q]];
                        con[q] = r[3 * jm[q] + 1];
                        depth_own[q] = r0.z;
                        dx[q] = __fsub_rn(r0.x, pxf); dy[q] = __fsub_rn(r0.y, pyf);
                        const float power = pair_power(dx[q], dy[q], con[q].x, con[q].y, con[q].z);
                        G[q] = expf(power);
                        const float alpha = fminf(0.99f, __fmul_rn(con[q].w, G[q]));
                        const bool valid = ok && (pos0 - (int)s_pos[buf][jm[q]] < last_contributor) && !(power > 0.0f) &&
                                           !(alpha < 1.0f / 255.0f);
                        al[q] = valid ? alpha : 0.0f;
                    }
                    float a[8];
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        const float theirs = __shfl_xor_sync(0xffffffffu, al[q], 16);
                        a[2 * q] = slot ? theirs : al[q];
                        a[2 * q + 1] = slot ? al[q] : theirs;
                    }
#ifdef GHR_TRACE
                    trace_steps++;
#endif
                    BTRACE_LAP(c_p1)
                    if (!__any_sync(0xffffffffu, (a[0] > 0.f) | (a[1] > 0.f) | (a[2] > 0.f) | (a[3] > 0.f) | (a[4] > 0.f) |
                                                     (a[5] > 0.f) | (a[6] > 0.f) | (a[7] > 0.f)))
                        continue;
                    if constexpr (NX == 0) {
                        // Single-attribute kernel.  Phase 2, the sequential part, carries TWO scalars per pixel instead of
                        // the reference's six recurrences (backward.cu:508-560): with g = (dL_dpixel, dL_ddepth, dL_dalpha)
                        // and s_k = <(colour_k, depth_k, 1), g>, the reference's
                        //     dL_dalpha_k = T_k <(c_k - accum_rec_k), g> - T_final / (1 - alpha_k) <bg, dL_dpixel>
                        // equals  s_k T_k - B_k / (1 - alpha_k)  where  B_k = T_final <bg, dL_dpixel> + sum_{j behind k} s_j alpha_j T_j
                        // (accum_rec_k T_k = (sum_{j behind k} c_j alpha_j T_j) / (1 - alpha_k)).  alpha = 0 (skipped here)
                        // leaves T and B unchanged without any select, so the chain is branch-free: per Gaussian one
                        // reciprocal, the dot product s, and three multiply-adds.  Phase 3 (gradients of this lane's own
                        // four Gaussians, 16-lane reductions, atomics) is straight-line code over the four, so the four
                        // independent butterfly reductions overlap.
                        float Town[8], Bown[8], sown[8], rown[8];
#pragma unroll
                        for (int k = 0; k < 8; k++) {
                            const float4 c = r[3 * j[k] + 2];
                            const float dep = rf[12 * j[k] + 2];
                            sown[k] = fmaf(c.x, dpix0, fmaf(c.y, dpix1, fmaf(c.z, dpix2, fmaf(dep, dL_ddepth, dL_dalpha))));
                            rown[k] = __fdividef(1.0f, 1.0f - a[k]);
                            T *= rown[k];
                            Town[k] = T; Bown[k] = Bacc;
                            Bacc = fmaf(sown[k], a[k] * T, Bacc);
                        }
                        uint32_t live[4];
                        float total[4];
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            live[q] = __ballot_sync(0xffffffffu, al[q] > 0.0f);
#ifdef GHR_TRACE
                            trace_live += ((live[q] & 0xFFFFu) != 0) + ((live[q] >> 16) != 0);
#endif
                            const float Tm = slot ? Town[2 * q + 1] : Town[2 * q], Bm = slot ? Bown[2 * q + 1] : Bown[2 * q];
                            const float sm = slot ? sown[2 * q + 1] : sown[2 * q], rm = slot ? rown[2 * q + 1] : rown[2 * q];
                            const float aq = al[q];
                            const float dchannel_dcolor = aq * Tm;
                            float dL_dopa = fmaf(sm, Tm, -rm * Bm);
                            dL_dopa = aq > 0.0f ? dL_dopa : 0.0f;
                            const float dL_dG = con[q].w * dL_dopa;
                            const float gdx = G[q] * dx[q], gdy = G[q] * dy[q];
                            const float dG_ddelx = -gdx * con[q].x - gdy * con[q].y;
                            const float dG_ddely = -gdy * con[q].z - gdx * con[q].y;
                            float v[9];
                            v[0] = dchannel_dcolor * dpix0;
                            v[1] = dchannel_dcolor * dpix1;
                            v[2] = dchannel_dcolor * dpix2;
                            v[3] = dL_dG * dG_ddelx * ddelx_dx;
                            v[4] = dL_dG * dG_ddely * ddely_dy;
                            v[5] = -0.5f * gdx * dx[q] * dL_dG;
                            v[6] = -0.5f * gdx * dy[q] * dL_dG;
                            v[7] = -0.5f * gdy * dy[q] * dL_dG;
                            v[8] = G[q] * dL_dopa;
                            total[q] = half_reduce9(v, lane);
                        }
#pragma unroll
                        for (int q = 0; q < 4; q++)
                            if ((live[q] & half_mask) != 0 && writer)
                                atomicAdd(accum + 12 * (size_t)s_id[buf][jm[q]] + vslot, total[q]);
                        BTRACE_LAP(c_p23)
                    } else {
                    // (multi-attribute kernels) phases 2 and 3, one pair of Gaussians (one per slot) at a time
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        // state after the pair's first Gaussian (S1: what slot 0 differentiates) and after its second (S2)
                        float Ts[2], a0s[2], a1s[2], a2s[2], ads[2], aas[2];
                        float xpart[2];  // this lane's channels' share of sum_ch (c - accum) * dL_dpixel, per Gaussian
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            const int k = 2 * q + h;
                            const float4 c = r[3 * j[k] + 2];
                            const float dep = rf[12 * j[k] + 2];
                            const bool on = a[k] > 0.0f;
                            const float inv = __fdividef(1.0f, 1.0f - a[k]);
                            const float om = 1.0f - last_alpha;
                            T = on ? T * inv : T;
                            acc0 = on ? fmaf(last_alpha, lc0, om * acc0) : acc0;
                            acc1 = on ? fmaf(last_alpha, lc1, om * acc1) : acc1;
                            acc2 = on ? fmaf(last_alpha, lc2, om * acc2) : acc2;
                            accum_depth_rec = on ? fmaf(last_alpha, last_depth, om * accum_depth_rec) : accum_depth_rec;
                            accum_alpha_rec = on ? fmaf(om, accum_alpha_rec, last_alpha) : accum_alpha_rec;
                            if constexpr (NX > 0) {
                                const float2* cx =
                                    reinterpret_cast<const float2*>(s_x + ((size_t)buf * BATCH + j[k]) * XS + slot * XH);
                                float part = 0.f;
#pragma unroll
                                for (int c = 0; c < CHS; c++) {
                                    const float2 cc = cx[c >> 1];
                                    const float col = (c & 1) ? cc.y : cc.x;
                                    accx[c] = on ? fmaf(last_alpha, lcx[c], om * accx[c]) : accx[c];
                                    lcx[c] = on ? col : lcx[c];
                                    part = fmaf(col - accx[c], dpx[c], part);
                                }
                                xpart[h] = part;
                            }
                            lc0 = on ? c.x : lc0; lc1 = on ? c.y : lc1; lc2 = on ? c.z : lc2;
                            last_depth = on ? dep : last_depth;
                            last_alpha = on ? a[k] : last_alpha;
                            Ts[h] = T; a0s[h] = acc0; a1s[h] = acc1; a2s[h] = acc2; ads[h] = accum_depth_rec; aas[h] = accum_alpha_rec;
                        }
                        const uint32_t live = __ballot_sync(0xffffffffu, al[q] > 0.0f);
#ifdef GHR_TRACE
                        trace_live += ((live & 0xFFFFu) != 0) + ((live >> 16) != 0);
#endif
                        if (live == 0) continue;
                        float xdot = 0.f;
                        if constexpr (NX > 0) {
                            // colour gradients of both Gaussians of the pair, this lane's channels
                            float xv[2 * CHS];
                            const float w0 = a[2 * q] * Ts[0], w1 = a[2 * q + 1] * Ts[1];
#pragma unroll
                            for (int c = 0; c < CHS; c++) { xv[c] = w0 * dpx[c]; xv[CHS + c] = w1 * dpx[c]; }
                            float tot[2];
                            half_reduce_n<2 * CHS>(xv, lane, tot);
#pragma unroll
                            for (int t = 0; t < 2; t++) {
                                if (xidx[t] >= 0) {
                                    const int h = xidx[t] >= CHS, c = xidx[t] - h * CHS;
                                    if ((live & (h ? 0xFFFF0000u : 0x0000FFFFu)) != 0)
                                        atomicAdd(xaccum + (size_t)s_id[buf][h ? j[2 * q + 1] : j[2 * q]] * XS + slot * XH + c, tot[t]);
                                }
                            }
                            // the other lane of the pixel holds the other half of the channels of MY Gaussian
                            const float theirs = __shfl_xor_sync(0xffffffffu, slot ? xpart[0] : xpart[1], 16);
                            xdot = (slot ? xpart[1] : xpart[0]) + theirs;
                        }
                        // this lane's own Gaussian of the pair: partial gradients of its pixel (zero where not valid)
                        const float Tm = slot ? Ts[1] : Ts[0];
                        const float m0 = slot ? a0s[1] : a0s[0], m1 = slot ? a1s[1] : a1s[0], m2 = slot ? a2s[1] : a2s[0];
                        const float md = slot ? ads[1] : ads[0], ma = slot ? aas[1] : aas[0];
                        const float4 c = r[3 * jm[q] + 2];
                        const float aq = al[q];
                        float v[9];
                        {
                            const float dchannel_dcolor = aq * Tm;
                            float dL_dopa = (c.x - m0) * dpix0;
                            dL_dopa = fmaf(c.y - m1, dpix1, dL_dopa);
                            dL_dopa = fmaf(c.z - m2, dpix2, dL_dopa);
                            dL_dopa = fmaf(depth_own[q] - md, dL_ddepth, dL_dopa);
                            dL_dopa = fmaf(1.0f - ma, dL_dalpha, dL_dopa);
                            if constexpr (NX > 0) dL_dopa += xdot;
                            dL_dopa *= Tm;
                            dL_dopa = fmaf(-T_final * __fdividef(1.0f, 1.0f - aq), bg_dot, dL_dopa);
                            dL_dopa = aq > 0.0f ? dL_dopa : 0.0f;
                            const float dL_dG = con[q].w * dL_dopa;
                            const float gdx = G[q] * dx[q], gdy = G[q] * dy[q];
                            const float dG_ddelx = -gdx * con[q].x - gdy * con[q].y;
                            const float dG_ddely = -gdy * con[q].z - gdx * con[q].y;
                            v[0] = dchannel_dcolor * dpix0;
                            v[1] = dchannel_dcolor * dpix1;
                            v[2] = dchannel_dcolor * dpix2;
                            v[3] = dL_dG * dG_ddelx * ddelx_dx;
                            v[4] = dL_dG * dG_ddely * ddely_dy;
                            v[5] = -0.5f * gdx * dx[q] * dL_dG;
                            v[6] = -0.5f * gdx * dy[q] * dL_dG;
                            v[7] = -0.5f * gdy * dy[q] * dL_dG;
                            v[8] = G[q] * dL_dopa;
                        }
                        const float total = half_reduce9(v, lane);
                        if ((live & half_mask) != 0 && writer)
                            atomicAdd(accum + 12 * (size_t)s_id[buf][jm[q]] + vslot, total);
                    }
                    BTRACE_LAP(c_p23)
                    }  // NX > 0
                }
            }
            __syncwarp();  // buffer `buf` is re-filled by the next iteration's stage()
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
#ifdef GHR_TRACE
        if (lane == 0 && item < (1u << 16)) {
            g_bwd_trace[item][0] = t_begin;
            g_bwd_trace[item][1] = gtime_b();
            g_bwd_trace[item][2] = (uint32_t)todo_total;
            g_bwd_trace[item][3] = (unsigned long long)trace_rounds;
            g_bwd_trace[item][4] = ((unsigned long long)trace_live << 32) | (uint32_t)trace_steps;
            g_bwd_trace[item][5] = (unsigned long long)c_wait;
            g_bwd_trace[item][6] = (unsigned long long)c_p1;
            g_bwd_trace[item][7] = (unsigned long long)c_p23;
        }
#endif
    }
}
#ifdef GHR_TRACE
}  // namespace ghr
extern "C" int ghr_debug_bwd_trace(void* out, size_t bytes)
{
    return (int)cudaMemcpyFromSymbol(out, ghr::g_bwd_trace, bytes);
}
namespace ghr {
#endif

template <int BATCH, int NSTAGE, int NX>
static cudaError_t launch_bwd_config(int ctas_per_sm, const ghr_raster_params& p, const char* geom, const char* binning,
                                     const char* img, const ghr_layout& L, const float* dL_dpix, const float* dL_ddepth,
                                     const float* dL_dalpha, float* accum, const float* xcol, const float* xdpix,
                                     float* xaccum, cudaStream_t s)
{
    const int tiles = ((p.W + GHR_TILE - 1) / GHR_TILE) * ((p.H + GHR_TILE - 1) / GHR_TILE);
    ghr_geom_header* hdr = reinterpret_cast<ghr_geom_header*>(const_cast<char*>(geom) + L.g_header);
    const int64_t items = (int64_t)tiles * 16;
    int64_t ctas = (items + kBwdWarps - 1) / kBwdWarps;
    if (ctas > (int64_t)sm_count() * ctas_per_sm) ctas = (int64_t)sm_count() * ctas_per_sm;
    auto kernel = blend_backward_kernel<BATCH, NSTAGE, NX>;
    const size_t smem = blend_bwd_warp_smem<BATCH, NSTAGE, NX>() * kBwdWarps;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kernel<<<(unsigned)ctas, kBwdWarps * 32, smem, s>>>(
        p.W, p.H, tiles, reinterpret_cast<const uint2*>(img + L.i_ranges),
        reinterpret_cast<const uint32_t*>(img + L.i_order), reinterpret_cast<const SortHeader*>(binning + L.b_header),
        reinterpret_cast<const uint2*>(binning + L.b_keys_a), reinterpret_cast<const uint2*>(binning + L.b_keys_b),
        reinterpret_cast<const float4*>(geom + L.g_rec), hdr, L.capacity, p.background,
        reinterpret_cast<const float*>(img + L.i_final_T), reinterpret_cast<const uint32_t*>(img + L.i_n_contrib), dL_dpix,
        dL_ddepth, dL_dalpha, accum, xcol, xdpix, xaccum);
    return cudaGetLastError();
}

cudaError_t launch_blend_backward(const ghr_raster_params& p, const char* geom, const char* binning, const char* img,
                                  const ghr_layout& L, const float* dL_dpix, const float* dL_ddepth,
                                  const float* dL_dalpha, float* accum, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_BLEND_BWD, s);
    count_launch();
    if (blend_throughput_regime(L.capacity))
        return launch_blend_backward_wide(p, geom, binning, img, L, dL_dpix, dL_ddepth, dL_dalpha, accum, s);
    ghr_geom_header* hdr = reinterpret_cast<ghr_geom_header*>(const_cast<char*>(geom) + L.g_header);
    // the queue head is scratch state inside the (otherwise read-only) geometry buffer
    cudaError_t e = cudaMemsetAsync(&hdr->blend_bwd_ticket, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    if (blend_shared_gpu())
        return launch_bwd_config<128, 2, 0>(2, p, geom, binning, img, L, dL_dpix, dL_ddepth, dL_dalpha, accum, nullptr,
                                            nullptr, nullptr, s);
#ifdef GHR_BWD_BATCH  // tuning builds only
    return launch_bwd_config<GHR_BWD_BATCH, GHR_BWD_NSTAGE, 0>(GHR_BWD_CTAS, p, geom, binning, img, L, dL_dpix, dL_ddepth, dL_dalpha, accum, nullptr, nullptr,
                                        nullptr, s);
#else
    return launch_bwd_config<256, 2, 0>(2, p, geom, binning, img, L, dL_dpix, dL_ddepth, dL_dalpha, accum, nullptr, nullptr,
                                        nullptr, s);
#endif
}

// Main image + n_attr (2, 4 or 6) packed attributes in one walk; latency regime only (ghr_api.cu checks).
cudaError_t launch_blend_backward_multi(const ghr_raster_params& p, const char* geom, const char* binning, const char* img,
                                        const ghr_layout& L, const float* dL_dpix, const float* dL_ddepth,
                                        const float* dL_dalpha, float* accum, int n_attr, const float* xcol,
                                        const float* xdpix, float* xaccum, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_BLEND_BWD, s);
    count_launch();
    ghr_geom_header* hdr = reinterpret_cast<ghr_geom_header*>(const_cast<char*>(geom) + L.g_header);
    cudaError_t e = cudaMemsetAsync(&hdr->blend_bwd_ticket, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    switch (n_attr) {
    case 2: return launch_bwd_config<128, 2, 2>(2, p, geom, binning, img, L, dL_dpix, dL_ddepth, dL_dalpha, accum, xcol, xdpix, xaccum, s);
    case 4: return launch_bwd_config<128, 2, 4>(2, p, geom, binning, img, L, dL_dpix, dL_ddepth, dL_dalpha, accum, xcol, xdpix, xaccum, s);
#ifdef GHR_MB_BATCH  // tuning builds only
    case 6: return launch_bwd_config<GHR_MB_BATCH, GHR_MB_NSTAGE, 6>(GHR_MB_CTAS, p, geom, binning, img, L, dL_dpix, dL_ddepth, dL_dalpha, accum, xcol, xdpix, xaccum, s);
#else
    case 6: return launch_bwd_config<96, 2, 6>(2, p, geom, binning, img, L, dL_dpix, dL_ddepth, dL_dalpha, accum, xcol, xdpix, xaccum, s);
#endif
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace ghr
