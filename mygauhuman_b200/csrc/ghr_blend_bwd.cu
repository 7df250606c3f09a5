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

constexpr int kBatchB = 256;

__device__ __forceinline__ void cp_async16b(void* smem, const void* gmem)
{
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}

// Sum v[0..8] over the warp.  Returns the total of value `slot` (see
// reduce_slot()) in every lane; lanes whose slot is >= 9 return garbage-free 0.
__device__ __forceinline__ float warp_reduce9(const float (&v)[9], uint32_t lane)
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

// Which of the 9 values a lane holds after warp_reduce9 (>= 9: none).
__device__ __forceinline__ int reduce_slot(uint32_t lane)
{
    const int w4 = ((lane & 4) ? 2 : 0) + ((lane & 2) ? 1 : 0);  // position inside the padded group of 4
    if (w4 >= 3) return 99;
    const int local = ((lane & 8) ? 3 : 0) + w4;                 // position inside the padded group of 6
    if (lane & 16) return local < 4 ? 5 + local : 99;
    return local < 5 ? local : 99;
}

__global__ void __launch_bounds__(256) blend_backward_kernel(
    int W, int H, const uint2* __restrict__ ranges, const SortHeader* __restrict__ sh,
    const uint32_t* __restrict__ vals_a, const uint32_t* __restrict__ vals_b, const float4* __restrict__ rec,
    const ghr_geom_header* __restrict__ hdr, int64_t capacity, const float* __restrict__ bg,
    const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
    const float* __restrict__ dL_ddepths, const float* __restrict__ dL_dalphas, float* __restrict__ accum)
{
    __shared__ __align__(16) float4 s_rec[2][kBatchB * 3];
    __shared__ uint32_t s_id[2][kBatchB];

    const uint32_t tx = threadIdx.x & 15, ty = threadIdx.x >> 4, lane = threadIdx.x & 31;
    const uint32_t gx = (W + GHR_TILE - 1) / GHR_TILE;
    const uint32_t pixx = blockIdx.x * GHR_TILE + tx, pixy = blockIdx.y * GHR_TILE + ty;
    const bool inside = pixx < (uint32_t)W && pixy < (uint32_t)H;
    const float pxf = (float)pixx, pyf = (float)pixy;
    const bool overflow = (int64_t)hdr->num_rendered > capacity;
    const uint2 range = overflow ? make_uint2(0, 0) : ranges[blockIdx.y * gx + blockIdx.x];
    const uint32_t* __restrict__ point_list = sh->sorted_sel ? vals_b : vals_a;
    const int todo_total = (int)(range.y - range.x);
    const int rounds = (todo_total + kBatchB - 1) / kBatchB;
    if (rounds == 0) return;

    const size_t pix = (size_t)pixy * W + pixx, HW = (size_t)H * W;
    const float T_final = inside ? final_Ts[pix] : 0.0f;
    float T = T_final;
    const int last_contributor = inside ? (int)n_contrib[pix] : 0;
    float dpix0 = 0.f, dpix1 = 0.f, dpix2 = 0.f, dL_ddepth = 0.f, dL_dalpha = 0.f;
    if (inside) {
        dpix0 = dL_dpixels[pix]; dpix1 = dL_dpixels[HW + pix]; dpix2 = dL_dpixels[2 * HW + pix];
        dL_ddepth = dL_ddepths[pix]; dL_dalpha = dL_dalphas[pix];
    }
    const float bg_dot = bg[0] * dpix0 + bg[1] * dpix1 + bg[2] * dpix2;
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;
    float accum_depth_rec = 0.f, accum_alpha_rec = 0.f, last_alpha = 0.f, last_depth = 0.f;
    const float ddelx_dx = (float)(0.5 * W), ddely_dy = (float)(0.5 * H);
    // highest list position any pixel of this warp still needs
    const int warp_last = __reduce_max_sync(0xffffffffu, last_contributor);
    const int slot = reduce_slot(lane);
    const bool writer = slot < 9 && !(lane & 1);

    auto stage = [&](int round, int buf) {
        const int progress = round * kBatchB + (int)threadIdx.x;
        if (progress < todo_total) {
            const uint32_t id = point_list[range.y - 1 - (uint32_t)progress];  // back to front
            s_id[buf][threadIdx.x] = id;
            const float4* src = rec + 3 * (size_t)id;
            float4* dst = &s_rec[buf][threadIdx.x * 3];
            cp_async16b(dst, src);
            cp_async16b(dst + 1, src + 1);
            cp_async16b(dst + 2, src + 2);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    stage(0, 0);
    int contributor = todo_total;
    int toDo = todo_total;
    for (int i = 0; i < rounds; i++, toDo -= kBatchB) {
        const int buf = i & 1;
        if (i + 1 < rounds) { stage(i + 1, buf ^ 1); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
        else asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        const int n = min(kBatchB, toDo);
        const float4* r = s_rec[buf];
        // entries with contributor-1 >= warp_last are beyond every pixel's last contributor
        int j0 = contributor - warp_last;
        j0 = j0 < 0 ? 0 : (j0 > n ? n : j0);
        contributor -= j0;
        for (int j = j0; j < n; j++) {
            contributor--;
            const float4 r0 = r[3 * j], r1 = r[3 * j + 1];
            const float dx = __fsub_rn(r0.x, pxf), dy = __fsub_rn(r0.y, pyf);
            const float power = pair_power(dx, dy, r1.x, r1.y, r1.z);
            const float G = expf(power);
            const float alpha = fminf(0.99f, __fmul_rn(r1.w, G));
            const bool valid = (contributor < last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
            if (!__any_sync(0xffffffffu, valid)) continue;
            float v[9];
#pragma unroll
            for (int k = 0; k < 9; k++) v[k] = 0.0f;
            if (valid) {
                const float4 r2 = r[3 * j + 2];
                T = T / (1.0f - alpha);
                const float dchannel_dcolor = alpha * T;
                float dL_dopa = 0.0f;
                acc0 = last_alpha * lc0 + (1.0f - last_alpha) * acc0; lc0 = r2.x;
                dL_dopa += (r2.x - acc0) * dpix0; v[0] = dchannel_dcolor * dpix0;
                acc1 = last_alpha * lc1 + (1.0f - last_alpha) * acc1; lc1 = r2.y;
                dL_dopa += (r2.y - acc1) * dpix1; v[1] = dchannel_dcolor * dpix1;
                acc2 = last_alpha * lc2 + (1.0f - last_alpha) * acc2; lc2 = r2.z;
                dL_dopa += (r2.z - acc2) * dpix2; v[2] = dchannel_dcolor * dpix2;
                accum_depth_rec = last_alpha * last_depth + (1.0f - last_alpha) * accum_depth_rec;
                last_depth = r0.z;
                dL_dopa += (r0.z - accum_depth_rec) * dL_ddepth;
                accum_alpha_rec = last_alpha + (1.0f - last_alpha) * accum_alpha_rec;
                dL_dopa += (1.0f - accum_alpha_rec) * dL_dalpha;
                dL_dopa *= T;
                last_alpha = alpha;
                dL_dopa += (-T_final / (1.0f - alpha)) * bg_dot;
                const float dL_dG = r1.w * dL_dopa;
                const float gdx = G * dx, gdy = G * dy;
                const float dG_ddelx = -gdx * r1.x - gdy * r1.y;
                const float dG_ddely = -gdy * r1.z - gdx * r1.y;
                v[3] = dL_dG * dG_ddelx * ddelx_dx;
                v[4] = dL_dG * dG_ddely * ddely_dy;
                v[5] = -0.5f * gdx * dx * dL_dG;
                v[6] = -0.5f * gdx * dy * dL_dG;
                v[7] = -0.5f * gdy * dy * dL_dG;
                v[8] = G * dL_dopa;
            }
            const float total = warp_reduce9(v, lane);
            if (writer) atomicAdd(accum + 12 * (size_t)s_id[buf][j] + slot, total);
        }
        __syncthreads();
    }
}

cudaError_t launch_blend_backward(const ghr_raster_params& p, const char* geom, const char* binning, const char* img,
                                  const ghr_layout& L, const float* dL_dpix, const float* dL_ddepth,
                                  const float* dL_dalpha, float* accum, cudaStream_t s)
{
    dim3 grid((p.W + GHR_TILE - 1) / GHR_TILE, (p.H + GHR_TILE - 1) / GHR_TILE);
    blend_backward_kernel<<<grid, 256, 0, s>>>(
        p.W, p.H, reinterpret_cast<const uint2*>(img + L.i_ranges),
        reinterpret_cast<const SortHeader*>(binning + L.b_header),
        reinterpret_cast<const uint32_t*>(binning + L.b_vals_a), reinterpret_cast<const uint32_t*>(binning + L.b_vals_b),
        reinterpret_cast<const float4*>(geom + L.g_rec), reinterpret_cast<const ghr_geom_header*>(geom + L.g_header),
        L.capacity, p.background, reinterpret_cast<const float*>(img + L.i_final_T),
        reinterpret_cast<const uint32_t*>(img + L.i_n_contrib), dL_dpix, dL_ddepth, dL_dalpha, accum);
    return cudaGetLastError();
}

}  // namespace ghr
