// LBS-weight offset network (reference nets/mlp_delta_weight_lbs.py:5-33 LBSOffsetDecoder, called from
// gaussian_renderer/__init__.py:105): per Gaussian
//     e = [x, sin(2^f x), cos(2^f x)]_{f=0..9}            63 values   (get_embedder(10), :36-83)
//     h = relu(W0 e + b0); h = relu(W1 h + b1); h = relu(W2 h + b2); h = relu(W3 [e, h] + b3);  logits = Wf h + bf   (24)
// as ONE kernel on the 5th-generation tensor cores: a CTA owns 128 Gaussians (the M = 128 rows of every layer),
// `tcgen05.mma.cta_group::1.kind::tf32` (128 x N x 8 per instruction, issued by one thread) accumulates in TMEM
// (128 lanes x 128 fp32 columns), the epilogue of a layer reads its accumulator row with `tcgen05.ld` (one thread per
// Gaussian), adds the bias, applies ReLU and writes the next layer's A operand back to shared memory; weights arrive from
// a prepacked blob (ghr_lbs_mlp_pack) by `cp.async.bulk` (the TMA engine's 1-D copy, completion on an mbarrier) through a
// three-slot ring that runs one layer ahead of the tensor core.  Nothing but xyz (12 B) is read and the 24 logits (96 B)
// written per Gaussian; the reference runs 5 cuDNN convolutions + 8 elementwise kernels with [P,128] activations in HBM.
//
// Arithmetic: TF32 operands (activations and weights rounded to nearest, `cvt.rna.tf32.f32`), fp32 accumulation -- what
// the reference's own Conv1d layers use on Ampere and later (torch.backends.cudnn.allow_tf32 defaults to True).
//
// Shared-memory operand layout (no swizzle, K-major, `SmemDescriptor` layout_type 0): 16-byte chunks of 4 consecutive K
// values; chunk c of row r at ((c * R + r) * 16) bytes.  Rows of an 8-row core matrix are 16 B apart, 8-row groups
// 128 B apart (SBO), the two chunks of one K = 8 instruction R * 16 B apart (LBO).
#include <stdio.h>
#include "ghr_common.cuh"
#include "ghr_internal.h"
#include "ghr_umma.cuh"

namespace ghr {

namespace {

constexpr int kRows = 128;            // Gaussians per CTA = UMMA M
constexpr int kThreads = 256;         // two threads per Gaussian: each takes half of the encoding and half of an accumulator row
constexpr int kHidden = 128;
constexpr int kEmb = 63;
constexpr int kOutPad = 32;           // 24 logits padded to a legal UMMA N
constexpr int kStageBytes = 32768;    // 64 K-columns x 128 rows x 4 B
// nine weight stages: L0: 1 | L1: 2 | L2: 2 | L3: 3 | fc: 1 (16 KB)
constexpr int kSlots = 3;
constexpr int kBiasFloats = 4 * kHidden + kOutPad;
constexpr size_t kBlobBytes = 8 * (size_t)kStageBytes + 16384 + sizeof(float) * kBiasFloats;

constexpr uint32_t kABytes = 48 * kRows * 16;  // 48 chunks: [0,32) current activation, [32,48) embedding (+1 zero pad)
constexpr uint32_t kSmemBytes = kABytes + kSlots * kStageBytes + sizeof(float) * kBiasFloats + 64;

__host__ __device__ constexpr int stage_layer(int s) { return s == 0 ? 0 : s <= 2 ? 1 : s <= 4 ? 2 : s <= 7 ? 3 : 4; }
__host__ __device__ constexpr int stage_first(int layer) { return layer == 0 ? 0 : layer == 1 ? 1 : layer == 2 ? 3 : layer == 3 ? 5 : 8; }
__host__ __device__ constexpr size_t stage_offset(int s) { return (size_t)s * kStageBytes; }
__host__ __device__ constexpr uint32_t stage_bytes(int s) { return s == 8 ? 16384u : (uint32_t)kStageBytes; }

__device__ __forceinline__ float to_tf32(float x)
{
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

// ---- prepack: the five weight matrices -> stage blobs in the shared-memory chunk layout, TF32-rounded; biases behind ----
struct MlpWeights {
    const float *w0, *w1, *w2, *w3, *wf;  // [128,63] [128,128] [128,128] [128,191] [24,128]   (Conv1d weight[out,in,1])
    const float *b0, *b1, *b2, *b3, *bf;
};

__global__ void __launch_bounds__(256) lbs_mlp_pack_kernel(const MlpWeights w, float* __restrict__ blob)
{
    const int total = (int)((kBlobBytes) / 4);
    for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
        const int wfloats = (8 * kStageBytes + 16384) / 4;
        if (i >= wfloats) {  // biases: b0 b1 b2 b3 | bf padded with zeros
            const int b = i - wfloats;
            const float* src[4] = {w.b0, w.b1, w.b2, w.b3};
            blob[i] = b < 4 * kHidden ? src[b / kHidden][b % kHidden] : (b - 4 * kHidden < 24 ? w.bf[b - 4 * kHidden] : 0.0f);
            continue;
        }
        const int s = i < 8 * (kStageBytes / 4) ? i / (kStageBytes / 4) : 8;
        const int j = i - s * (kStageBytes / 4);
        const int layer = stage_layer(s);
        const int R = layer == 4 ? kOutPad : kHidden;
        const int n = (j / 4) % R;
        const int k = (j / (4 * R)) * 4 + (j & 3) + (layer == 4 ? 0 : (s - stage_first(layer)) * 64);
        float v = 0.0f;
        if (layer == 0) v = k < kEmb ? w.w0[n * kEmb + k] : 0.0f;
        else if (layer == 1) v = w.w1[n * kHidden + k];
        else if (layer == 2) v = w.w2[n * kHidden + k];
        else if (layer == 3) v = k < kHidden ? w.w3[n * 191 + kEmb + k] : (k < kHidden + kEmb ? w.w3[n * 191 + (k - kHidden)] : 0.0f);
        else v = n < 24 ? w.wf[n * kHidden + k] : 0.0f;
        blob[i] = to_tf32(v);
    }
}

using namespace umma;

#ifdef GHR_MLP_TRACE  // tuning builds only (tools/build_variant.sh): phase timestamps of block 0, printed by its thread 0
#define MLP_T(i) do { if (blockIdx.x == 0 && t == 0) tr[i] = clock64(); } while (0)
#else
#define MLP_T(i) do { } while (0)
#endif

// ---- the network -----------------------------------------------------------------------------------------------------
// round to TF32 (nearest, ties away: what cvt.rna.tf32.f32 does) with two integer instructions: the conversion
// instruction runs at a quarter of the ALU rate and a layer's epilogue needs 128 of them per thread
__device__ __forceinline__ float tf32_rn(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u); }

__global__ void __launch_bounds__(kThreads, 1) lbs_mlp_forward_kernel(int P, int n_tiles, const float* __restrict__ xyz,
                                                                   const unsigned char* __restrict__ blob,
                                                                   float* __restrict__ logits)
{
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* sA = smem;                                  // [48 chunks][128 rows] x 16 B
    unsigned char* sB = smem + kABytes;                        // kSlots x 32 KB
    float* sBias = reinterpret_cast<float*>(sB + kSlots * kStageBytes);
    uint64_t* full = reinterpret_cast<uint64_t*>(sBias + kBiasFloats);  // [kSlots]
    uint64_t* acc_bar = full + kSlots;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);

    const int t = threadIdx.x, warp = t >> 5;
    const int row = t & (kRows - 1), half = t >> 7;  // warp w may touch TMEM lanes 32 (w % 4) ..: exactly its rows
#ifdef GHR_MLP_TRACE
    long long tr[24];
#endif
    MLP_T(0);

    if (t == 0) {
        for (int i = 0; i < kSlots; i++) mbar_init(&full[i], 1);
        mbar_init(acc_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {  // one warp allocates the accumulator: 128 lanes x 128 columns of fp32
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(128) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = t; i < kBiasFloats; i += kThreads) sBias[i] = reinterpret_cast<const float*>(blob + 8 * (size_t)kStageBytes + 16384)[i];
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot;
    MLP_T(1);

    // thread 0 is the weight producer and the MMA issuer.  Stage s of a tile lives in ring slot s % 3; every slot is
    // filled three times per tile, so the parity a wait needs is the number of fills already consumed from that slot.
    uint32_t consumed[kSlots] = {0u, 0u, 0u};
    auto fetch = [&](int s) {
        const int slot = s % kSlots;
        mbar_expect_tx(&full[slot], stage_bytes(s));
        bulk_load(sB + slot * kStageBytes, blob + stage_offset(s), stage_bytes(s), &full[slot]);
    };
    if (t == 0)
        for (int s = 0; s < kSlots; s++) fetch(s);  // on their way while the first encoding is computed

    uint32_t acc_parity = 0;
#pragma unroll 1
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int row0 = tile * kRows;
        const bool more = tile + (int)gridDim.x < n_tiles;

        // positional encoding of this thread's Gaussian -> chunks [32, 48) of row t (zero row beyond P).  sin / cos of
        // 2^0 x and 2^5 x are library calls; the four octaves after each follow by angle doubling (sin 2a = 2 sin a cos a,
        // cos 2a = 1 - 2 sin^2 a): the error grows by <= 2.9x per octave, <= 4e-6 after four -- 1/60 of the TF32 rounding
        // the value receives next -- for a fifth of the 30 sincosf calls (3 us of a 14 us tile).
        {
            const int g = row0 + row;
            const bool live = g < P;
            float* erow = reinterpret_cast<float*>(sA + (32 * kRows + row) * 16);  // value k at erow[(k / 4) * kRows * 4 + k % 4]
            auto put = [&](int k, float v) { erow[(k >> 2) * (kRows * 4) + (k & 3)] = tf32_rn(live ? v : 0.0f); };
#pragma unroll
            for (int c = 0; c < 3; c++) {
                const float pc = live ? xyz[3 * (size_t)g + c] : 0.0f;
                if (half == 0) put(c, pc);
                float sn, cs;
                sincosf(pc * (half ? 32.0f : 1.0f), &sn, &cs);  // half 0: octaves 0..4, half 1: octaves 5..9
#pragma unroll
                for (int f = 0; f < 5; f++) {
                    const int oct = 5 * half + f;
                    put(3 + 6 * oct + c, sn);
                    put(6 + 6 * oct + c, cs);
                    const float s2 = 2.0f * sn * cs, c2 = fmaf(-2.0f * sn, sn, 1.0f);
                    sn = s2; cs = c2;
                }
            }
            if (half) put(63, 0.0f);
        }
        MLP_T(2);

#ifdef GHR_MLP_TRACE
#pragma unroll
#else
#pragma unroll 1
#endif
        for (int layer = 0; layer < 5; layer++) {
            // the A operand of this layer was written with ordinary stores: make it visible to the tensor core (async
            // proxy); the previous accumulator has been read (tcgen05.wait::ld in the epilogue)
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();
            if (t == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const int s0 = stage_first(layer);
                const int n_st = layer == 0 ? 1 : layer == 3 ? 3 : layer == 4 ? 1 : 2;
                const int Rb = layer == 4 ? kOutPad : kHidden;
                const uint32_t idesc = idesc_tf32(Rb);
                const uint32_t a_chunk0 = layer == 0 ? 32u : 0u;
                uint32_t first = 1;
                for (int i = 0; i < n_st; i++) {
                    const int slot = (s0 + i) % kSlots;
                    mbar_wait(&full[slot], consumed[slot] & 1u);
                    consumed[slot]++;
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const int ksteps = layer == 4 ? 16 : 8;  // K = 8 per instruction; the fc stage holds all of K = 128
                    const uint32_t a_base = smem_u32(sA) + (a_chunk0 + 16u * i) * kRows * 16;
                    const uint32_t b_base = smem_u32(sB) + slot * kStageBytes;
                    for (int k = 0; k < ksteps; k++) {
                        const uint64_t ad = smem_desc(a_base + k * 2 * kRows * 16, kRows * 16, 128);
                        const uint64_t bd = smem_desc(b_base + k * 2 * Rb * 16, Rb * 16, 128);
                        umma_tf32(tmem, ad, bd, idesc, first ? 0u : 1u);
                        first = 0;
                    }
                }
                umma_commit(acc_bar);  // arrives when every MMA above has finished (accumulator complete, operands read)
            }
            MLP_T(3 + 3 * layer);
            mbar_wait(acc_bar, acc_parity);
            acc_parity ^= 1;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            MLP_T(4 + 3 * layer);

            // the slots this layer read are free again: fetch the stages of the layer after next (after the last two
            // layers: the first stages of this CTA's next tile)
            if (t == 0) {
                if (layer == 0) fetch(3);
                else if (layer == 1) { fetch(4); fetch(5); }
                else if (layer == 2) { fetch(6); fetch(7); }
                else if (layer == 3) { fetch(8); if (more) { fetch(0); fetch(1); } }
                else if (more) fetch(2);
            }

            // epilogue: thread t owns accumulator row t (TMEM lane t; a warp may only touch its own 32 lanes)
            const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);
            if (layer < 4) {
                const float4* bias4 = reinterpret_cast<const float4*>(sBias + layer * kHidden);
                // this thread's half of the row: both 32-column loads are in flight before the first value is used
                uint32_t v[64];
                tmem_ld32(trow + 64 * half, v);
                tmem_ld32(trow + 64 * half + 32, v + 32);
                tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 16; q++) {
                    const float4 bq = bias4[16 * half + q];
                    float4 o;
                    o.x = tf32_rn(fmaxf(__uint_as_float(v[4 * q]) + bq.x, 0.0f));
                    o.y = tf32_rn(fmaxf(__uint_as_float(v[4 * q + 1]) + bq.y, 0.0f));
                    o.z = tf32_rn(fmaxf(__uint_as_float(v[4 * q + 2]) + bq.z, 0.0f));
                    o.w = tf32_rn(fmaxf(__uint_as_float(v[4 * q + 3]) + bq.w, 0.0f));
                    *reinterpret_cast<float4*>(sA + ((16 * half + q) * kRows + row) * 16) = o;
                }
                MLP_T(5 + 3 * layer);
            } else {
                uint32_t v[16];
                tmem_ld16(trow + 16 * half, v);  // 24 logits: columns [0,16) here, [16,24) in the other half
                tmem_ld_wait();
                const int g = row0 + row;
                if (g < P) {
                    const float4* bias4 = reinterpret_cast<const float4*>(sBias + 4 * kHidden) + 4 * half;
                    float4* out = reinterpret_cast<float4*>(logits + 24 * (size_t)g) + 4 * half;
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        if (4 * half + q >= 6) break;
                        const float4 bq = bias4[q];
                        out[q] = make_float4(__uint_as_float(v[4 * q]) + bq.x, __uint_as_float(v[4 * q + 1]) + bq.y,
                                             __uint_as_float(v[4 * q + 2]) + bq.z, __uint_as_float(v[4 * q + 3]) + bq.w);
                    }
                }
            }
        }
        MLP_T(18);
#ifdef GHR_MLP_TRACE
        if (blockIdx.x == 0 && t == 0) {
            printf("mlp trace (cycles): setup %lld, embed %lld", tr[1] - tr[0], tr[2] - tr[1]);
            long long prev = tr[2];
            for (int l = 0; l < 5; l++) {
                printf(" | L%d issue %lld mma-wait %lld", l, tr[3 + 3 * l] - prev, tr[4 + 3 * l] - tr[3 + 3 * l]);
                prev = tr[4 + 3 * l];
                const long long end = l < 4 ? tr[5 + 3 * l] : tr[18];
                printf(" epi %lld", end - prev);
                prev = end;
            }
            printf(" | total %lld\n", tr[18] - tr[0]);
            tr[1] = tr[18];
        }
#endif
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        __syncwarp();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(128) : "memory");
    }
}

}  // namespace

// blob = [TF32 stages | biases (fp32)] [FP16 stages]: both kernels' weight formats, packed together
size_t lbs_mlp_blob_bytes() { return kBlobBytes + lbs_mlp_blob_f16_bytes(); }

// which tensor-core kernel serves ghr_lbs_mlp_forward: "f16" = two tiles in flight per CTA, FP16 operands (ghr_mlp_f16.cu,
// default), "tf32" = this file's kernel.  Same 11-bit significand either way.
static bool use_f16_kernel()
{
    static const int choice = [] { const char* e = getenv("GHR_MLP_KERNEL"); return e && e[0] == 't' ? 0 : 1; }();
    return choice == 1;
}

cudaError_t launch_lbs_mlp_pack(const float* const* weights, const float* const* biases, void* blob, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_MISC, s);
    count_launch();
    MlpWeights w{weights[0], weights[1], weights[2], weights[3], weights[4], biases[0], biases[1], biases[2], biases[3], biases[4]};
    lbs_mlp_pack_kernel<<<64, 256, 0, s>>>(w, reinterpret_cast<float*>(blob));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    return launch_lbs_mlp_pack_f16(weights, reinterpret_cast<char*>(blob) + kBlobBytes, s);
}

cudaError_t launch_lbs_mlp_forward(int P, const float* xyz, const void* blob, float* logits, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_MISC, s);
    if (use_f16_kernel())
        return launch_lbs_mlp_forward_f16(P, xyz, reinterpret_cast<const char*>(blob) + kBlobBytes,
                                          reinterpret_cast<const float*>(reinterpret_cast<const char*>(blob) + 8 * (size_t)kStageBytes + 16384),
                                          logits, s);
    static unsigned long long attr_set = 0ull;
    if (first_use_on_current_device(attr_set)) {
        cudaError_t e = cudaFuncSetAttribute(lbs_mlp_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
        if (e != cudaSuccess) return e;
    }
    count_launch();
    const int n_tiles = (P + kRows - 1) / kRows;
    const int ctas = n_tiles < sm_count() ? n_tiles : sm_count();  // persistent: one CTA per SM walks the tiles
    lbs_mlp_forward_kernel<<<ctas, kThreads, kSmemBytes, s>>>(P, n_tiles, xyz, reinterpret_cast<const unsigned char*>(blob), logits);
    return cudaGetLastError();
}

}  // namespace ghr
