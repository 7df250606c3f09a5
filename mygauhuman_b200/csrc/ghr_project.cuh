// Per-Gaussian projection shared by the single-view preprocess kernel (ghr_preprocess.cu) and the view-batched one
// (ghr_views.cu): reference forward.cu:156-256 preprocessCUDA, and the decoupled look-back scan that replaces
// cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:279).  The fp32 operation sequence is pinned with intrinsics
// (ghr_common.cuh), so both kernels produce the reference's radii / rects bit for bit.
#pragma once
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

struct PreArgs {
    int P, D, M, W, H;
    float scale_modifier, tan_fovx, tan_fovy, focal_x, focal_y;
    int prefiltered;
    const float *means3D, *shs, *colors_precomp, *opacities, *scales, *rotations, *cov3D_precomp;
    const float *view, *proj, *campos;
    int32_t* radii;
    float4* rec;
    uint32_t* tiles_touched;
    uint32_t* point_offsets;
    float* cov3D_out;
    uint8_t* clamped;
    uint64_t* scan_status;
    ghr_geom_header* hdr;
    int32_t* n_host;
    int nblocks;
    int skip_rgb;  // view-batched alpha-only rendering: the SH -> RGB evaluation is left out (rgb = 0)
    int split_scan;  // 1: large P -- blocks only publish their tile count (no look-back); two small kernels finish the scan
};

constexpr uint64_t kFlagAgg = 1ull << 32;
constexpr uint64_t kFlagIncl = 2ull << 32;

// Decoupled look-back over 64-bit {flag,value} words; called by all lanes of warp 0.
__device__ __forceinline__ uint32_t lookback_exclusive(uint64_t* status, int tile, uint32_t aggregate)
{
    const uint32_t lane = lane_id();
    if (tile == 0) {
        if (lane == 0) st_relaxed_u64(&status[0], kFlagIncl | aggregate);
        return 0;
    }
    if (lane == 0) st_relaxed_u64(&status[tile], kFlagAgg | aggregate);
    uint32_t excl = 0;
    int base = tile - 1;
    while (true) {
        const int j = base - (int)lane;
        const uint64_t s = (j >= 0) ? ld_relaxed_u64(&status[j]) : kFlagIncl;
        const uint32_t flag = (uint32_t)(s >> 32);
        const uint32_t invalid = __ballot_sync(0xffffffffu, flag == 0);
        const uint32_t incl = __ballot_sync(0xffffffffu, flag == 2);
        if (incl) {
            const int fi = __ffs(incl) - 1;
            if (invalid & ((1u << fi) - 1u)) continue;  // a nearer predecessor is not published yet
            uint32_t v = ((int)lane <= fi) ? (uint32_t)s : 0u;
            excl += __reduce_add_sync(0xffffffffu, v);
            break;
        }
        if (invalid) continue;
        excl += __reduce_add_sync(0xffffffffu, (uint32_t)s);
        base -= 32;
    }
    if (lane == 0) st_relaxed_u64(&status[tile], kFlagIncl | (uint64_t)(excl + aggregate));
    return excl;
}

// The part of a preprocess block that follows the per-Gaussian projection: depth range of the visible Gaussians for the
// sort, block-wide inclusive scan of `tiles`, decoupled look-back over the blocks, point_offsets.  `tile` is the block's
// ticket, `idx` this thread's Gaussian; dbits_hi / dbits_lo_inv the depth bit pattern of a visible Gaussian and its
// complement (0 otherwise).  s_warp[8] and s_excl are block-shared scratch.  Called by all GHR_BLOCK threads.
__device__ __forceinline__ void finish_preprocess_block(const PreArgs& a, int tile, int idx, uint32_t tiles,
                                                        uint32_t dbits_hi, uint32_t dbits_lo_inv, uint32_t* s_warp,
                                                        uint32_t& s_excl)
{
    const int tid = threadIdx.x;
    // depth range of the visible Gaussians, for the sort (ghr_sort.cu: squeezed digits): depth > 0.2, so the bit
    // patterns order like the values
    dbits_hi = __reduce_max_sync(0xffffffffu, dbits_hi);
    dbits_lo_inv = __reduce_max_sync(0xffffffffu, dbits_lo_inv);
    if ((tid & 31) == 0 && dbits_hi) {
        atomicMax(&a.hdr->depth_hi, dbits_hi);
        atomicMax(&a.hdr->depth_lo_inv, dbits_lo_inv);
    }

    // block-wide inclusive scan of `tiles`
    const uint32_t lane = lane_id(), warp = tid >> 5;
    uint32_t v = tiles;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= (uint32_t)o) v += t;
    }
    if (lane == 31) s_warp[warp] = v;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < 8 ? s_warp[lane] : 0;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= (uint32_t)o) w += t;
        }
        const uint32_t block_total = __shfl_sync(0xffffffffu, w, 7);
        // exclusive prefix over warps = inclusive of the previous warp
        const uint32_t prev = __shfl_up_sync(0xffffffffu, w, 1);
        if (lane < 8) s_warp[lane] = (lane == 0) ? 0u : prev;
        if (a.split_scan) {  // the block's total goes to its status word; scan_block_totals_kernel / add_block_prefix_kernel follow
            if (lane == 0) { a.scan_status[tile] = block_total; s_excl = 0; }
        } else {
        const uint32_t excl = lookback_exclusive(a.scan_status, tile, block_total);
        if (lane == 0) {
            s_excl = excl;
            if (tile == a.nblocks - 1) {
                a.hdr->num_rendered = excl + block_total;
                if (a.n_host) *a.n_host = (int32_t)(excl + block_total);
            }
        }
        }
    }
    __syncthreads();
    if (idx < a.P) a.point_offsets[idx] = s_excl + s_warp[warp] + v;
}

// Projects a Gaussian with mean (x, y, z) and opacity opac_in with the camera (V = view, Pm = proj; column-major as the
// reference reads them) and, if it survives the culls, writes its 48-byte blend record to a.rec[o].  Returns tiles
// touched (0 = culled), the integer radius and whether the near-plane cull (z <= 0.2) took it.  Where the remaining
// inputs come from is the caller's business (global arrays in the rasteriser's own preprocess kernels, registers in the
// fused LBS + projection kernel):  hint(projx, projy) runs once the NDC centre is known (prefetch),  cov3(c3) delivers
// the packed 3D covariance of a Gaussian that passed the near cull,  rgbf(rgb) the colour of one that passed EVERY cull.
template <class HintFn, class Cov3Fn, class RgbFn>
__device__ __forceinline__ uint32_t project_core(const PreArgs& a, size_t o, float x, float y, float z, float opac_in,
                                                 const float* V, const float* Pm, int& radius, bool& near_culled,
                                                 HintFn hint, Cov3Fn cov3, RgbFn rgbf, float* depth_out = nullptr)
{
    uint32_t tiles = 0;
    radius = 0;
    near_culled = false;
    const float pvz = affine_row(x, y, z, V[2], V[6], V[10], V[14]);
    if (pvz <= 0.2f) {
        near_culled = true;
    } else {
        const float hx = affine_row(x, y, z, Pm[0], Pm[4], Pm[8], Pm[12]);
        const float hy = affine_row(x, y, z, Pm[1], Pm[5], Pm[9], Pm[13]);
        const float hw = affine_row(x, y, z, Pm[3], Pm[7], Pm[11], Pm[15]);
        const float p_w = __frcp_rn(__fadd_rn(hw, 0.0000001f));
        const float projx = __fmul_rn(hx, p_w), projy = __fmul_rn(hy, p_w);
        hint(projx, projy);
        float c3[6];
        cov3(c3);
        Cov2DCtx ctx;
        cov2d_T(x, y, z, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, V, ctx);
        float c00, c01, c11;
        cov2d(ctx, c3, c00, c01, c11);
        const float det = __fmaf_rn(c00, c11, -__fmul_rn(c01, c01));
        if (det != 0.0f) {
            const float det_inv = __frcp_rn(det);
            const float con0 = __fmul_rn(c11, det_inv), con1 = __fmul_rn(det_inv, -c01),
                        con2 = __fmul_rn(c00, det_inv);
            const float mid = __fmul_rn(__fadd_rn(c00, c11), 0.5f);
            const float sq = __fsqrt_rn(fmaxf(__fmaf_rn(mid, mid, -det), 0.1f));
            const float lambda1 = __fadd_rn(mid, sq), lambda2 = __fsub_rn(mid, sq);
            const float my_radius = ceilf(__fmul_rn(__fsqrt_rn(fmaxf(lambda1, lambda2)), 3.0f));
            const float px = ndc2pix(projx, a.W), py = ndc2pix(projy, a.H);
            const uint32_t gx = (a.W + GHR_TILE - 1) / GHR_TILE, gy = (a.H + GHR_TILE - 1) / GHR_TILE;
            uint32_t minx, miny, maxx, maxy;
            const int ri = __float2int_rz(my_radius);
            get_rect(px, py, ri, gx, gy, minx, miny, maxx, maxy);
            const uint32_t cnt = (maxx - minx) * (maxy - miny);
            if (cnt != 0) {
                float rgb[3];
                rgbf(rgb);
                radius = ri;
                tiles = cnt;
                float4* r = a.rec + 3 * o;
                r[0] = make_float4(px, py, pvz, __int_as_float(ri));
                if (depth_out) *depth_out = pvz;  // (reading it back from the record is a store -> load round trip to L2)
                const float opac = opac_in;
                r[1] = make_float4(con0, con1, con2, opac);
                r[2] = make_float4(rgb[0], rgb[1], rgb[2], __uint_as_float(pack_cull_extent(con0, con1, con2, opac)));
            }
        }
    }
    return tiles;
}

// project_core over the rasteriser's own input arrays (reference forward.cu:156-256 preprocessCUDA): Gaussian `idx`,
// record / cov3D / clamp slot `o`.
__device__ __forceinline__ uint32_t project_gaussian(const PreArgs& a, int idx, size_t o, const float* V, const float* Pm,
                                                     const float* cam, int& radius, bool& near_culled,
                                                     float* depth_out = nullptr)
{
    const float x = a.means3D[3 * idx], y = a.means3D[3 * idx + 1], z = a.means3D[3 * idx + 2];
    // Every per-Gaussian input that does not depend on the culls is requested here, next to the mean, so that the
    // kernel has ONE exposed DRAM round trip per Gaussian instead of three dependent ones (mean -> scale/rotation ->
    // SH): at 3 M Gaussians the kernel ran at 2.2 TB/s with 20 % issue activity and 24 warps per SM -- latency, not
    // bandwidth.  The near cull (z <= 0.2) removes almost nothing, so nothing extra is read.
    float4 q_rot = make_float4(0.f, 0.f, 0.f, 0.f);
    float s_x = 0.f, s_y = 0.f, s_z = 0.f;
    if (!a.cov3D_precomp) {
        q_rot = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
        s_x = __ldg(a.scales + 3 * idx); s_y = __ldg(a.scales + 3 * idx + 1); s_z = __ldg(a.scales + 3 * idx + 2);
    }
    const float opac_in = __ldg(a.opacities + idx);
    auto hint = [&](float projx, float projy) {
        // SH coefficients (192 bytes at degree 3) are only read for Gaussians that survive every cull, ~150 instructions
        // from here: if the centre lands in or near the image, pull the row into L2 now (no register is tied up), so
        // the loads below find it there.  A Gaussian that is culled after all costs two wasted sector fetches.
        if (a.shs && !a.colors_precomp && !a.skip_rgb && fabsf(projx) < 1.25f && fabsf(projy) < 1.25f) {
            const char* sp = reinterpret_cast<const char*>(a.shs + (size_t)idx * a.M * 3);
            const int nb = a.M * 12;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(sp));
            if (nb > 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(sp + 128));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(sp + nb - 4));
        }
    };
    auto cov3 = [&](float* c3) {
        if (a.cov3D_precomp) {
            const float2* cp = reinterpret_cast<const float2*>(a.cov3D_precomp + 6 * (size_t)idx);
            const float2 c01 = __ldg(cp), c23 = __ldg(cp + 1), c45 = __ldg(cp + 2);
            c3[0] = c01.x; c3[1] = c01.y; c3[2] = c23.x; c3[3] = c23.y; c3[4] = c45.x; c3[5] = c45.y;
        } else {
            cov3d_from_scale_rot(s_x, s_y, s_z, a.scale_modifier, q_rot.x, q_rot.y, q_rot.z, q_rot.w, c3);
            float2* co = reinterpret_cast<float2*>(a.cov3D_out + 6 * o);
            if (a.cov3D_out) {
                co[0] = make_float2(c3[0], c3[1]); co[1] = make_float2(c3[2], c3[3]); co[2] = make_float2(c3[4], c3[5]);
            }
        }
    };
    auto rgbf = [&](float* rgb) {
        if (a.colors_precomp) {
            rgb[0] = a.colors_precomp[3 * idx]; rgb[1] = a.colors_precomp[3 * idx + 1];
            rgb[2] = a.colors_precomp[3 * idx + 2];
        } else if (a.skip_rgb) {
            rgb[0] = rgb[1] = rgb[2] = 0.0f;
        } else {
            float sh[48];
            const int nf = min(a.M * 3, 48);
            const float* sp = a.shs + (size_t)idx * a.M * 3;
            if ((nf & 3) == 0) {
                const float4* sp4 = reinterpret_cast<const float4*>(sp);
#pragma unroll
                for (int k = 0; k < 12; k++)
                    if (4 * k < nf) {
                        const float4 v = __ldg(sp4 + k);
                        sh[4 * k] = v.x; sh[4 * k + 1] = v.y; sh[4 * k + 2] = v.z; sh[4 * k + 3] = v.w;
                    }
            } else {
#pragma unroll
                for (int k = 0; k < 48; k++)
                    if (k < nf) sh[k] = __ldg(sp + k);
            }
            uint32_t cm;
            sh_to_rgb(a.D, sh, x - cam[0], y - cam[1], z - cam[2], rgb, cm);
            if (a.clamped) a.clamped[o] = (uint8_t)cm;
        }
    };
    return project_core(a, o, x, y, z, opac_in, V, Pm, radius, near_culled, hint, cov3, rgbf, depth_out);
}

}  // namespace ghr
