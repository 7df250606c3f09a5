// Per-tile front-to-back alpha blending with depth and alpha-mask outputs
// (reference forward.cu:261-383 renderCUDA).
//
// One CTA per 16x16 tile, one thread per pixel (thread_rank = ty*16+tx as in
// the reference).  Gaussians are staged in batches of 256 packed 48-byte
// records {px,py,depth,_ | conic.xyz,opacity | r,g,b,_} with cp.async
// (LDGSTS) into a double-buffered shared-memory ring, so the inner loop never
// touches global memory (the reference gathers colour and depth from global
// per contributing pair, forward.cu:360,362).
//
// Bound: fp32 issue + MUFU.EX2, not HBM (algorithmic bytes 44*N + 28*X).
// The arithmetic sequence per pair is the reference's compiled sequence
// (pair_power, expf, fma accumulation) so final_T/n_contrib are bit-exact.
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

constexpr int kBatch = 256;

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

__global__ void __launch_bounds__(256) blend_forward_kernel(int W, int H, const uint2* __restrict__ ranges,
                                                            const SortHeader* __restrict__ sh,
                                                            const uint32_t* __restrict__ vals_a,
                                                            const uint32_t* __restrict__ vals_b,
                                                            const float4* __restrict__ rec,
                                                            const ghr_geom_header* __restrict__ hdr, int64_t capacity,
                                                            const float* __restrict__ bg, float* __restrict__ out_color,
                                                            float* __restrict__ out_depth, float* __restrict__ out_alpha,
                                                            float* __restrict__ final_T, uint32_t* __restrict__ n_contrib)
{
    __shared__ __align__(16) float4 s_rec[2][kBatch * 3];

    const uint32_t tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const uint32_t gx = (W + GHR_TILE - 1) / GHR_TILE;
    const uint32_t pixx = blockIdx.x * GHR_TILE + tx, pixy = blockIdx.y * GHR_TILE + ty;
    const bool inside = pixx < (uint32_t)W && pixy < (uint32_t)H;
    const float pxf = (float)pixx, pyf = (float)pixy;
    const bool overflow = (int64_t)hdr->num_rendered > capacity;
    uint2 range = overflow ? make_uint2(0, 0) : ranges[blockIdx.y * gx + blockIdx.x];
    const uint32_t* __restrict__ point_list = sh->sorted_sel ? vals_b : vals_a;

    const int todo_total = (int)(range.y - range.x);
    const int rounds = (todo_total + kBatch - 1) / kBatch;

    auto stage = [&](int round, int buf) {
        const uint32_t k = range.x + (uint32_t)round * kBatch + threadIdx.x;
        if (k < range.y) {
            const uint32_t id = point_list[k];
            const float4* src = rec + 3 * (size_t)id;
            float4* dst = &s_rec[buf][threadIdx.x * 3];
            cp_async16(dst, src);
            cp_async16(dst + 1, src + 1);
            cp_async16(dst + 2, src + 2);
        }
        cp_async_commit();
    };

    bool done = !inside;
    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, wsum = 0.f, Dacc = 0.f;
    uint32_t contributor = 0, last_contributor = 0;

    if (rounds > 0) stage(0, 0);
    int toDo = todo_total;
    for (int i = 0; i < rounds; i++, toDo -= kBatch) {
        const int buf = i & 1;
        if (i + 1 < rounds) { stage(i + 1, buf ^ 1); cp_async_wait<1>(); }
        else cp_async_wait<0>();
        // barrier doubles as the "whole tile done" vote (forward.cu:313-315)
        const int num_done = __syncthreads_count(done);
        if (num_done == 256) break;
        const int n = min(kBatch, toDo);
        const float4* r = s_rec[buf];
        for (int j = 0; !done && j < n; j++) {
            contributor++;
            const float4 r0 = r[3 * j], r1 = r[3 * j + 1];
            const float dx = __fsub_rn(r0.x, pxf), dy = __fsub_rn(r0.y, pyf);
            const float power = pair_power(dx, dy, r1.x, r1.y, r1.z);
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, __fmul_rn(r1.w, expf(power)));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
            if (test_T < 0.0001f) { done = true; continue; }
            const float4 r2 = r[3 * j + 2];
            C0 = __fmaf_rn(T, __fmul_rn(alpha, r2.x), C0);
            C1 = __fmaf_rn(T, __fmul_rn(alpha, r2.y), C1);
            C2 = __fmaf_rn(T, __fmul_rn(alpha, r2.z), C2);
            wsum = __fmaf_rn(T, alpha, wsum);
            Dacc = __fmaf_rn(T, __fmul_rn(alpha, r0.z), Dacc);
            T = test_T;
            last_contributor = contributor;
        }
        __syncthreads();  // buffer `buf` is re-filled two rounds later by stage(i+2)
    }
    cp_async_wait<0>();

    if (inside) {
        const size_t pix = (size_t)pixy * W + pixx, HW = (size_t)H * W;
        final_T[pix] = T;
        n_contrib[pix] = last_contributor;
        out_color[pix] = __fmaf_rn(T, bg[0], C0);
        out_color[HW + pix] = __fmaf_rn(T, bg[1], C1);
        out_color[2 * HW + pix] = __fmaf_rn(T, bg[2], C2);
        out_alpha[pix] = wsum;
        out_depth[pix] = Dacc;
    }
}

cudaError_t launch_blend_forward(const ghr_raster_params& p, const char* geom, const char* binning, char* img,
                                 const ghr_layout& L, float* out_color, float* out_depth, float* out_alpha,
                                 cudaStream_t s)
{
    dim3 grid((p.W + GHR_TILE - 1) / GHR_TILE, (p.H + GHR_TILE - 1) / GHR_TILE);
    const int64_t capacity = L.capacity;
    blend_forward_kernel<<<grid, 256, 0, s>>>(
        p.W, p.H, reinterpret_cast<const uint2*>(img + L.i_ranges),
        reinterpret_cast<const SortHeader*>(binning + L.b_header),
        reinterpret_cast<const uint32_t*>(binning + L.b_vals_a), reinterpret_cast<const uint32_t*>(binning + L.b_vals_b),
        reinterpret_cast<const float4*>(geom + L.g_rec), reinterpret_cast<const ghr_geom_header*>(geom + L.g_header),
        capacity, p.background, out_color, out_depth, out_alpha, reinterpret_cast<float*>(img + L.i_final_T),
        reinterpret_cast<uint32_t*>(img + L.i_n_contrib));
    return cudaGetLastError();
}

}  // namespace ghr
