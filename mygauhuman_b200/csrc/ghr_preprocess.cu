// Stage 1: per-Gaussian projection / EWA covariance / radius / tile rect /
// SH->RGB (reference forward.cu:156-256 preprocessCUDA) fused with the
// prefix sum over tiles_touched (reference rasterizer_impl.cu:279,
// cub::DeviceScan::InclusiveSum) as a single-pass decoupled look-back scan,
// so num_rendered is known when this one kernel retires.  Also
// checkFrustum (rasterizer_impl.cu:54-66).
//
// Roofline: HBM stream.  Algorithmic bytes per Gaussian:
// 48*P + 28*V (precomputed colour + cov3D) or 244*P + 67*V (SH deg3 + scale/rot);
// we additionally write the packed 48-B blend record per visible Gaussian.
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

__global__ void __launch_bounds__(GHR_BLOCK) preprocess_kernel(const PreArgs a)
{
    __shared__ float s_cam[36];  // view[16] proj[16] campos[3]
    __shared__ uint32_t s_ticket;
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_excl;

    const int tid = threadIdx.x;
    if (tid < 16) s_cam[tid] = a.view[tid];
    else if (tid < 32) s_cam[tid] = a.proj[tid - 16];
    else if (tid < 35) s_cam[tid] = a.campos[tid - 32];
    if (tid == 0) s_ticket = atomicAdd(&a.hdr->scan_ticket, 1u);
    __syncthreads();
    const int tile = (int)s_ticket;
    const int idx = tile * GHR_BLOCK + tid;
    const float* V = s_cam;
    const float* Pm = s_cam + 16;

    uint32_t tiles = 0;
    if (idx < a.P) {
        int radius = 0;
        const float x = a.means3D[3 * idx], y = a.means3D[3 * idx + 1], z = a.means3D[3 * idx + 2];
        const float pvz = affine_row(x, y, z, V[2], V[6], V[10], V[14]);
        if (pvz <= 0.2f) {
            if (a.prefiltered) atomicMax(&a.hdr->error, (uint32_t)(-GHR_ERR_PREFILTER));
        } else {
            const float hx = affine_row(x, y, z, Pm[0], Pm[4], Pm[8], Pm[12]);
            const float hy = affine_row(x, y, z, Pm[1], Pm[5], Pm[9], Pm[13]);
            const float hw = affine_row(x, y, z, Pm[3], Pm[7], Pm[11], Pm[15]);
            const float p_w = __frcp_rn(__fadd_rn(hw, 0.0000001f));
            const float projx = __fmul_rn(hx, p_w), projy = __fmul_rn(hy, p_w);
            float c3[6];
            if (a.cov3D_precomp) {
                const float2* cp = reinterpret_cast<const float2*>(a.cov3D_precomp + 6 * (size_t)idx);
                const float2 c01 = __ldg(cp), c23 = __ldg(cp + 1), c45 = __ldg(cp + 2);
                c3[0] = c01.x; c3[1] = c01.y; c3[2] = c23.x; c3[3] = c23.y; c3[4] = c45.x; c3[5] = c45.y;
            } else {
                const float4 q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
                cov3d_from_scale_rot(a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2],
                                     a.scale_modifier, q.x, q.y, q.z, q.w, c3);
                float2* co = reinterpret_cast<float2*>(a.cov3D_out + 6 * (size_t)idx);
                co[0] = make_float2(c3[0], c3[1]); co[1] = make_float2(c3[2], c3[3]); co[2] = make_float2(c3[4], c3[5]);
            }
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
                    if (a.colors_precomp) {
                        rgb[0] = a.colors_precomp[3 * idx]; rgb[1] = a.colors_precomp[3 * idx + 1];
                        rgb[2] = a.colors_precomp[3 * idx + 2];
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
                        sh_to_rgb(a.D, sh, x - s_cam[32], y - s_cam[33], z - s_cam[34], rgb, cm);
                        a.clamped[idx] = (uint8_t)cm;
                    }
                    radius = ri;
                    tiles = cnt;
                    float4* r = a.rec + 3 * (size_t)idx;
                    r[0] = make_float4(px, py, pvz, __int_as_float(ri));
                    const float opac = a.opacities[idx];
                    r[1] = make_float4(con0, con1, con2, opac);
                    r[2] = make_float4(rgb[0], rgb[1], rgb[2], __uint_as_float(pack_cull_extent(con0, con1, con2, opac)));
                }
            }
        }
        a.radii[idx] = radius;
        a.tiles_touched[idx] = tiles;
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
        const uint32_t excl = lookback_exclusive(a.scan_status, tile, block_total);
        if (lane == 0) {
            s_excl = excl;
            if (tile == a.nblocks - 1) {
                a.hdr->num_rendered = excl + block_total;
                if (a.n_host) *a.n_host = (int32_t)(excl + block_total);
            }
        }
    }
    __syncthreads();
    if (idx < a.P) a.point_offsets[idx] = s_excl + s_warp[warp] + v;
}

__global__ void __launch_bounds__(GHR_BLOCK) mark_visible_kernel(int P, const float* __restrict__ means3D,
                                                                 const float* __restrict__ V, uint8_t* present)
{
    const int idx = blockIdx.x * GHR_BLOCK + threadIdx.x;
    if (idx >= P) return;
    const float z = affine_row(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2], V[2], V[6], V[10], V[14]);
    present[idx] = !(z <= 0.2f);
}

cudaError_t launch_preprocess(const ghr_raster_params& p, int32_t* radii, char* geom, const ghr_layout& L,
                              int32_t* n_host, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_PREPROCESS, s);
    PreArgs a;
    a.P = p.P; a.D = p.D; a.M = p.M; a.W = p.W; a.H = p.H;
    a.scale_modifier = p.scale_modifier; a.tan_fovx = p.tan_fovx; a.tan_fovy = p.tan_fovy;
    a.focal_y = p.H / (2.0f * p.tan_fovy);  // rasterizer_impl.cu:224-225
    a.focal_x = p.W / (2.0f * p.tan_fovx);
    a.prefiltered = p.prefiltered;
    a.means3D = p.means3D; a.shs = p.shs; a.colors_precomp = p.colors_precomp; a.opacities = p.opacities;
    a.scales = p.scales; a.rotations = p.rotations; a.cov3D_precomp = p.cov3D_precomp;
    a.view = p.viewmatrix; a.proj = p.projmatrix; a.campos = p.campos;
    a.radii = radii;
    a.rec = reinterpret_cast<float4*>(geom + L.g_rec);
    a.tiles_touched = reinterpret_cast<uint32_t*>(geom + L.g_tiles_touched);
    a.point_offsets = reinterpret_cast<uint32_t*>(geom + L.g_point_offsets);
    a.cov3D_out = reinterpret_cast<float*>(geom + L.g_cov3D);
    a.clamped = reinterpret_cast<uint8_t*>(geom + L.g_clamped);
    a.scan_status = reinterpret_cast<uint64_t*>(geom + L.g_scan_status);
    a.hdr = reinterpret_cast<ghr_geom_header*>(geom + L.g_header);
    a.n_host = n_host;
    a.nblocks = (p.P + GHR_BLOCK - 1) / GHR_BLOCK;
    // header + look-back words are contiguous: one memset
    cudaError_t e = cudaMemsetAsync(geom + L.g_header, 0, L.g_rec - L.g_header, s);
    if (e != cudaSuccess) return e;
    preprocess_kernel<<<a.nblocks, GHR_BLOCK, 0, s>>>(a);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_mark_visible(int P, const float* means3D, const float* V, uint8_t* present, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_MISC, s);
    count_launch();
    mark_visible_kernel<<<(P + GHR_BLOCK - 1) / GHR_BLOCK, GHR_BLOCK, 0, s>>>(P, means3D, V, present);
    return cudaGetLastError();
}

}  // namespace ghr
