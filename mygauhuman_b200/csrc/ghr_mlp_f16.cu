// LBS-weight offset network, second tensor-core kernel: FP16 operands (the same 11-bit significand as TF32, half the
// bytes), fp32 accumulation, TWO tiles of 128 Gaussians in flight per CTA.
//
// ghr_mlp.cu (TF32) spends a tile's life in a dependent chain -- MMA, commit, epilogue, fence, MMA -- with the tensor core
// idle during the epilogues, and its 198 KB of shared memory allow one CTA per SM.  With 2-byte operands a tile's A
// operand is 48 KB and a weight stage 16 KB, so one CTA holds two tiles, two accumulators (TMEM columns [0,128) and
// [128,256)) and a six-slot weight ring, and the roles are split:
//   warp 8, lane 0   weight producer (cp.async.bulk into the ring) and MMA issuer: layer L of tile X, then of tile Y
//   warps 0-3 / 4-7  tile X / tile Y: positional encoding, and per layer the epilogue (tcgen05.ld of the accumulator row,
//                    bias, ReLU, FP16 rounding, the next A operand)
// so the tensor core works on one tile while the CUDA cores finish the other; a weight stage is fetched once per PAIR of
// tiles.  Hand-offs are mbarriers: w_full[slot] (TMA bytes), a_ready[tile] (128 arrivals: this tile's next A operand is
// in shared memory and its accumulator has been read), acc_ready[tile] (tcgen05.commit: the layer's MMAs are done),
// layer_done (both tiles' MMAs of a layer are done: its weight slots can be refilled).
// Shared-memory operand layout as in ghr_mlp.cu with 8 FP16 values per 16-byte chunk (K = 16 per instruction).
#include <cuda_fp16.h>
#include <stdio.h>
#include "ghr_common.cuh"
#include "ghr_internal.h"
#include "ghr_umma.cuh"

namespace ghr {

namespace {

using namespace umma;

constexpr int kRows = 128;
constexpr int kHidden = 128;
constexpr int kEmb = 63;
constexpr int kOutPad = 32;
constexpr int kStageBytes = 16384;   // 64 K-columns x 128 rows x 2 B  (fc: 128 K-columns x 32 rows = 8 KB)
constexpr int kStagesPerPair = 9;    // L0: 1 | L1: 2 | L2: 2 | L3: 3 | fc: 1
constexpr int kSlots = 6;
constexpr int kBiasFloats = 4 * kHidden + kOutPad;
constexpr size_t kBlobBytes = 8 * (size_t)kStageBytes + 8192;
constexpr uint32_t kABytes = 24 * kRows * 16;  // per tile: chunks [0,16) activation, [16,24) encoding (+1 zero pad)
constexpr int kThreads = 288;
constexpr uint32_t kSmemBytes = 2 * kABytes + kSlots * kStageBytes + sizeof(float) * kBiasFloats + 128;

__host__ __device__ constexpr int stage_layer(int s) { return s == 0 ? 0 : s <= 2 ? 1 : s <= 4 ? 2 : s <= 7 ? 3 : 4; }
__host__ __device__ constexpr int stage_first(int layer) { return layer == 0 ? 0 : layer == 1 ? 1 : layer == 2 ? 3 : layer == 3 ? 5 : 8; }
__host__ __device__ constexpr uint32_t stage_bytes(int s) { return s == 8 ? 8192u : (uint32_t)kStageBytes; }

struct MlpWeights {
    const float *w0, *w1, *w2, *w3, *wf;
};

__global__ void __launch_bounds__(256) lbs_mlp_pack_f16_kernel(const MlpWeights w, __half* __restrict__ blob)
{
    const int total = (int)(kBlobBytes / 2);
    const int per_stage = kStageBytes / 2;
    for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
        const int s = i < 8 * per_stage ? i / per_stage : 8;
        const int j = i - s * per_stage;
        const int layer = stage_layer(s);
        const int R = layer == 4 ? kOutPad : kHidden;
        const int n = (j / 8) % R;
        const int k = (j / (8 * R)) * 8 + (j & 7) + (layer == 4 ? 0 : (s - stage_first(layer)) * 64);
        float v = 0.0f;
        if (layer == 0) v = k < kEmb ? w.w0[n * kEmb + k] : 0.0f;
        else if (layer == 1) v = w.w1[n * kHidden + k];
        else if (layer == 2) v = w.w2[n * kHidden + k];
        else if (layer == 3) v = k < kHidden ? w.w3[n * 191 + kEmb + k] : (k < kHidden + kEmb ? w.w3[n * 191 + (k - kHidden)] : 0.0f);
        else v = n < 24 ? w.wf[n * kHidden + k] : 0.0f;
        blob[i] = __float2half_rn(fminf(fmaxf(v, -65504.0f), 65504.0f));
    }
}

__device__ __forceinline__ uint32_t pack_h2(float a, float b)
{
    const __half2 h = __floats2half2_rn(fminf(a, 65504.0f), fminf(b, 65504.0f));
    return *reinterpret_cast<const uint32_t*>(&h);
}

__global__ void __launch_bounds__(kThreads, 1) lbs_mlp_forward_f16_kernel(int P, int n_pairs, const float* __restrict__ xyz,
                                                                          const unsigned char* __restrict__ blob,
                                                                          const float* __restrict__ biases,
                                                                          float* __restrict__ logits)
{
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* sA = smem;                                   // [2 tiles][24 chunks][128 rows] x 16 B
    unsigned char* sB = smem + 2 * kABytes;                     // kSlots x 16 KB
    float* sBias = reinterpret_cast<float*>(sB + kSlots * kStageBytes);
    uint64_t* w_full = reinterpret_cast<uint64_t*>(sBias + kBiasFloats);  // [kSlots]
    uint64_t* a_ready = w_full + kSlots;                                   // [2]
    uint64_t* acc_ready = a_ready + 2;                                     // [2]
    uint64_t* layer_done = acc_ready + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(layer_done + 1);

    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;

    if (t == 0) {
        for (int i = 0; i < kSlots; i++) mbar_init(&w_full[i], 1);
        for (int i = 0; i < 2; i++) { mbar_init(&a_ready[i], kRows); mbar_init(&acc_ready[i], 1); }
        mbar_init(layer_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {  // two accumulators: 128 lanes x 256 fp32 columns
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(256) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = t; i < kBiasFloats; i += kThreads) sBias[i] = biases[i];
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot;

    const int my_pairs = blockIdx.x < n_pairs ? (n_pairs - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

    if (warp == 8) {
        // ================= weight producer + MMA issuer (one thread) =================
        if (lane == 0) {
            const int total_q = my_pairs * kStagesPerPair;
            int fetched = 0;
            auto fetch = [&]() {
                const int s = fetched % kStagesPerPair, slot = fetched % kSlots;
                mbar_expect_tx(&w_full[slot], stage_bytes(s));
                bulk_load(sB + slot * kStageBytes, blob + (size_t)s * kStageBytes, stage_bytes(s), &w_full[slot]);
                fetched++;
            };
            while (fetched < kSlots && fetched < total_q) fetch();
            uint32_t a_par[2] = {0u, 0u}, l_par = 0u;
            int q = 0;  // stages consumed so far (over all pairs)
            for (int pr = 0; pr < my_pairs; pr++) {
#pragma unroll 1
                for (int layer = 0; layer < 5; layer++) {
                    const int n_st = layer == 0 ? 1 : layer == 3 ? 3 : layer == 4 ? 1 : 2;
                    const int Rb = layer == 4 ? kOutPad : kHidden;
                    const uint32_t idesc = idesc_f16(Rb);
                    const uint32_t a_chunk0 = layer == 0 ? 16u : 0u;
                    for (int i = 0; i < n_st; i++) mbar_wait(&w_full[(q + i) % kSlots], (uint32_t)(((q + i) / kSlots) & 1));
#pragma unroll 1
                    for (int tile = 0; tile < 2; tile++) {
                        mbar_wait(&a_ready[tile], a_par[tile]);
                        a_par[tile] ^= 1u;
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        uint32_t first = 1;
                        for (int i = 0; i < n_st; i++) {
                            const int slot = (q + i) % kSlots;
                            const int ksteps = layer == 4 ? 8 : 4;  // K = 16 per instruction; the fc stage holds all of K = 128
                            const uint32_t a_base = smem_u32(sA) + tile * kABytes + (a_chunk0 + 8u * i) * kRows * 16;
                            const uint32_t b_base = smem_u32(sB) + slot * kStageBytes;
                            for (int k = 0; k < ksteps; k++) {
                                const uint64_t ad = smem_desc(a_base + k * 2 * kRows * 16, kRows * 16, 128);
                                const uint64_t bd = smem_desc(b_base + k * 2 * Rb * 16, Rb * 16, 128);
                                umma_f16(tmem + tile * kHidden, ad, bd, idesc, first ? 0u : 1u);
                                first = 0;
                            }
                        }
                        umma_commit(&acc_ready[tile]);
                    }
                    umma_commit(layer_done);  // both tiles' MMAs of this layer: its weight slots are free afterwards
                    mbar_wait(layer_done, l_par);
                    l_par ^= 1u;
                    q += n_st;
                    for (int i = 0; i < n_st && fetched < total_q; i++) fetch();
                }
            }
        }
    } else {
        // ================= tile warps: encoding + epilogues =================
        const int tile = warp >> 2, row = t & (kRows - 1);
        unsigned char* myA = sA + tile * kABytes;
        const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16) + tile * kHidden;
        uint32_t c_par = 0u;
        for (int pr = 0; pr < my_pairs; pr++) {
            const int g = (2 * ((int)blockIdx.x + pr * (int)gridDim.x) + tile) * kRows + row;
            const bool live = g < P;
            {   // positional encoding -> chunks [16, 24) of this row (see ghr_mlp.cu for the angle-doubling argument)
                float e[64];
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    const float pc = live ? xyz[3 * (size_t)g + c] : 0.0f;
                    e[c] = pc;
                    float sn = 0.0f, cs = 1.0f;
#pragma unroll
                    for (int f = 0; f < 10; f++) {
                        if (f % 5 == 0) sincosf(pc * (float)(1 << f), &sn, &cs);
                        e[3 + 6 * f + c] = sn;
                        e[6 + 6 * f + c] = cs;
                        const float s2 = 2.0f * sn * cs, c2 = fmaf(-2.0f * sn, sn, 1.0f);
                        sn = s2; cs = c2;
                    }
                }
                e[63] = 0.0f;
                if (!live) {
#pragma unroll
                    for (int k = 0; k < 64; k++) e[k] = 0.0f;
                }
#pragma unroll
                for (int c = 0; c < 8; c++)
                    *reinterpret_cast<uint4*>(myA + ((16 + c) * kRows + row) * 16) =
                        make_uint4(pack_h2(e[8 * c], e[8 * c + 1]), pack_h2(e[8 * c + 2], e[8 * c + 3]),
                                   pack_h2(e[8 * c + 4], e[8 * c + 5]), pack_h2(e[8 * c + 6], e[8 * c + 7]));
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(&a_ready[tile]);

#pragma unroll 1
            for (int layer = 0; layer < 5; layer++) {
                mbar_wait(&acc_ready[tile], c_par);
                c_par ^= 1u;
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (layer < 4) {
                    const float4* bias4 = reinterpret_cast<const float4*>(sBias + layer * kHidden);
                    uint32_t v[128];
#pragma unroll
                    for (int c0 = 0; c0 < kHidden; c0 += 32) tmem_ld32(trow + c0, v + c0);
                    tmem_ld_wait();
#pragma unroll
                    for (int q8 = 0; q8 < 16; q8++) {
                        const float4 b0 = bias4[2 * q8], b1 = bias4[2 * q8 + 1];
                        const float* f = reinterpret_cast<const float*>(v + 8 * q8);
                        const uint4 o = make_uint4(pack_h2(fmaxf(f[0] + b0.x, 0.0f), fmaxf(f[1] + b0.y, 0.0f)),
                                                   pack_h2(fmaxf(f[2] + b0.z, 0.0f), fmaxf(f[3] + b0.w, 0.0f)),
                                                   pack_h2(fmaxf(f[4] + b1.x, 0.0f), fmaxf(f[5] + b1.y, 0.0f)),
                                                   pack_h2(fmaxf(f[6] + b1.z, 0.0f), fmaxf(f[7] + b1.w, 0.0f)));
                        *reinterpret_cast<uint4*>(myA + (q8 * kRows + row) * 16) = o;
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    mbar_arrive(&a_ready[tile]);
                } else {
                    uint32_t v[32];
                    tmem_ld32(trow, v);
                    tmem_ld_wait();
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    if (live) {
                        const float4* bias4 = reinterpret_cast<const float4*>(sBias + 4 * kHidden);
                        float4* out = reinterpret_cast<float4*>(logits + 24 * (size_t)g);
#pragma unroll
                        for (int q4 = 0; q4 < 6; q4++) {
                            const float4 bq = bias4[q4];
                            out[q4] = make_float4(__uint_as_float(v[4 * q4]) + bq.x, __uint_as_float(v[4 * q4 + 1]) + bq.y,
                                                  __uint_as_float(v[4 * q4 + 2]) + bq.z, __uint_as_float(v[4 * q4 + 3]) + bq.w);
                        }
                    }
                }
            }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        __syncwarp();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(256) : "memory");
    }
}

}  // namespace

size_t lbs_mlp_blob_f16_bytes() { return kBlobBytes; }

cudaError_t launch_lbs_mlp_pack_f16(const float* const* weights, void* blob_f16, cudaStream_t s)
{
    count_launch();
    MlpWeights w{weights[0], weights[1], weights[2], weights[3], weights[4]};
    lbs_mlp_pack_f16_kernel<<<64, 256, 0, s>>>(w, reinterpret_cast<__half*>(blob_f16));
    return cudaGetLastError();
}

cudaError_t launch_lbs_mlp_forward_f16(int P, const float* xyz, const void* blob_f16, const float* biases, float* logits,
                                       cudaStream_t s)
{
    static unsigned long long attr_set = 0ull;
    if (first_use_on_current_device(attr_set)) {
        cudaError_t e = cudaFuncSetAttribute(lbs_mlp_forward_f16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
        if (e != cudaSuccess) return e;
    }
    count_launch();
    const int n_tiles = (P + kRows - 1) / kRows;
    const int n_pairs = (n_tiles + 1) / 2;
    const int ctas = n_pairs < sm_count() ? n_pairs : sm_count();
    lbs_mlp_forward_f16_kernel<<<ctas, kThreads, kSmemBytes, s>>>(P, n_pairs, xyz, reinterpret_cast<const unsigned char*>(blob_f16),
                                                                  biases, logits);
    return cudaGetLastError();
}

}  // namespace ghr
