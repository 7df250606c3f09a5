// Optimiser-side kernels of the training iteration:
//   * adam_multi_kernel  -- torch.optim.Adam(l, lr=0.0, eps=1e-15).step() of the reference
//     (scene/gaussian_model.py:283, train.py:418-419) for ALL parameter groups in one launch,
//     optionally zeroing the gradients in the same pass (optimizer.zero_grad / bucket reset);
//   * densify_stats_kernel -- the per-iteration densification statistics
//     (train.py:401-404, scene/gaussian_model.py:764-766).
// Both are pure HBM streams: 28 (32 with the gradient reset) bytes per parameter, 40 bytes per Gaussian.
#include <math.h>
#include <stdio.h>
#include <string.h>
#include "ghr_internal.h"

namespace ghr {
namespace {

constexpr int kAdamThreads = 256;
constexpr int kAdamVec = 2;                                   // float4 per array per thread
constexpr int kAdamChunk = kAdamThreads * kAdamVec * 4;       // floats per CTA (2 048)

struct AdamSeg {
    float* p;
    float* g;
    float* m;
    float* v;
    int64_t count;
    float neg_step_size;   // -lr / (1 - beta1^t)
    float inv_bc2_sqrt;    // 1 / sqrt(1 - beta2^t)
    uint32_t first_chunk;  // CTA index of the segment's first chunk
    uint32_t vec_ok;       // all four pointers 16-byte aligned
};
struct AdamTable {
    AdamSeg seg[GHR_ADAM_MAX_GROUPS];
    int n;
    float w1;      // 1 - beta1
    float beta2;
    float w2;      // 1 - beta2
    float eps;
    int zero_grads;
};

// One element of torch's _single_tensor_adam (torch/optim/adam.py): exp_avg.lerp_(grad, 1 - beta1);
// exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2); denom = (exp_avg_sq.sqrt() / bias_correction2_sqrt) + eps;
// param.addcdiv_(exp_avg, denom, value=-step_size).  The CUDA foreach path multiplies by the reciprocal of the scalar
// divisor, which is what inv_bc2_sqrt does.
__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamTable& t, const AdamSeg& s)
{
    m = __fmaf_rn(g - m, t.w1, m);
    v = __fmaf_rn(g * g, t.w2, v * t.beta2);
    const float denom = __fmaf_rn(sqrtf(v), s.inv_bc2_sqrt, t.eps);
    p = __fmaf_rn(s.neg_step_size, m / denom, p);
}

__global__ void __launch_bounds__(kAdamThreads) adam_multi_kernel(const __grid_constant__ AdamTable t)
{
    // which segment does this CTA work on?  (<= 16 entries, uniform per CTA)
    int k = 0;
#pragma unroll 1
    for (int i = 1; i < t.n; ++i)
        if (blockIdx.x >= t.seg[i].first_chunk) k = i;
    const AdamSeg& s = t.seg[k];
    const int64_t base = int64_t(blockIdx.x - s.first_chunk) * kAdamChunk;
    const int64_t left = s.count - base;
    if (s.vec_ok && left >= kAdamChunk) {
        float4 p[kAdamVec], g[kAdamVec], m[kAdamVec], v[kAdamVec];
#pragma unroll
        for (int j = 0; j < kAdamVec; ++j) {  // all loads first: 8 independent 16-byte requests per thread
            const int64_t i = base / 4 + j * kAdamThreads + threadIdx.x;
            p[j] = reinterpret_cast<const float4*>(s.p)[i];
            g[j] = reinterpret_cast<const float4*>(s.g)[i];
            m[j] = reinterpret_cast<const float4*>(s.m)[i];
            v[j] = reinterpret_cast<const float4*>(s.v)[i];
        }
#pragma unroll
        for (int j = 0; j < kAdamVec; ++j) {
            const int64_t i = base / 4 + j * kAdamThreads + threadIdx.x;
            adam_one(p[j].x, g[j].x, m[j].x, v[j].x, t, s);
            adam_one(p[j].y, g[j].y, m[j].y, v[j].y, t, s);
            adam_one(p[j].z, g[j].z, m[j].z, v[j].z, t, s);
            adam_one(p[j].w, g[j].w, m[j].w, v[j].w, t, s);
            reinterpret_cast<float4*>(s.p)[i] = p[j];
            reinterpret_cast<float4*>(s.m)[i] = m[j];
            reinterpret_cast<float4*>(s.v)[i] = v[j];
            if (t.zero_grads) reinterpret_cast<float4*>(s.g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        return;
    }
    // ragged tail of a segment, or unaligned views: scalar
    const int64_t n = left < kAdamChunk ? left : kAdamChunk;
    for (int64_t i = threadIdx.x; i < n; i += kAdamThreads) {
        float p = s.p[base + i], m = s.m[base + i], v = s.v[base + i];
        const float g = s.g[base + i];
        adam_one(p, g, m, v, t, s);
        s.p[base + i] = p;
        s.m[base + i] = m;
        s.v[base + i] = v;
        if (t.zero_grads) s.g[base + i] = 0.f;
    }
}

// train.py:401-404 + gaussian_model.py:764-766 for visibility_filter = radii > 0:
//   max_radii2D[f] = max(max_radii2D[f], radii[f]); xyz_gradient_accum[f] += ||viewspace.grad[f, :2]||; denom[f] += 1
__global__ void __launch_bounds__(256) densify_stats_kernel(int P, const int32_t* __restrict__ radii,
                                                            const float* __restrict__ dL_dmean2D, int stride,
                                                            float* __restrict__ max_radii2D,
                                                            float* __restrict__ grad_accum, float* __restrict__ denom)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r <= 0) return;
    if (max_radii2D) max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
    const float gx = dL_dmean2D[(int64_t)i * stride], gy = dL_dmean2D[(int64_t)i * stride + 1];
    if (grad_accum) grad_accum[i] += sqrtf(__fmaf_rn(gy, gy, gx * gx));
    if (denom) denom[i] += 1.f;
}


// ---- reduce-scatter + Adam + all-gather in ONE kernel over NVLink peer memory (view-sharded data parallelism) --------
// Every rank holds a replica of the flat parameter buffer and its own flat gradient bucket (the sum over ITS views).  Rank
// r owns the slice [lo, hi) of the buffers: it reads that slice of every rank's gradients straight from peer memory (P2P
// loads), sums them in rank order, applies Adam with ITS shard of the moments (optimizer state is sharded, ZeRO-1), and
// stores the new parameters into every replica (P2P stores).  Per rank and step: (W-1)/W of the bucket in and out over
// NVLink -- what a ring all-reduce moves -- but no separate all-reduce, no second pass over the gradients, 1/W of the
// Adam work and state, and bit-identical replicas (exactly one rank computes each element).
struct ShardSeg {
    int64_t begin, end;      // [begin, end) of the flat buffers this (group x shard) piece covers
    float neg_step_size, inv_bc2_sqrt;
};
struct ShardTable {
    const float* grads[GHR_MAX_PEERS];
    float* params[GHR_MAX_PEERS];
    float* m;                // [hi - lo] moments of the owned slice
    float* v;
    int64_t lo;
    ShardSeg seg[GHR_ADAM_MAX_GROUPS];
    uint32_t first_chunk[GHR_ADAM_MAX_GROUPS];
    int n_seg, world, rank;
    float w1, beta2, w2, eps;
    // optional in-kernel cross-rank barriers: flags[r] = rank r's flag words, mapped here (uint32 [kFlagWords]):
    // [ARRIVE + q] / [DONE + q] = last step for which rank q has arrived / finished, [COUNT] = local CTA counter
    uint32_t* flags[GHR_MAX_PEERS];
    uint32_t epoch;
    unsigned long long spin_budget_ns;  // a wait that lasts longer than this gives up (see wait_for_peers)
};
// [ERROR]: 0 or why this rank gave up a wait (kSpin*), latched until the host clears it; [ABORT]: non-zero = some rank
// gave up or its host raised (ghr_sharded_abort) -- every wait on every rank then returns at once
constexpr int kFlagArrive = 0, kFlagDone = GHR_MAX_PEERS, kFlagCount = 2 * GHR_MAX_PEERS, kFlagError = kFlagCount + 1,
              kFlagAbort = kFlagCount + 2;
constexpr uint32_t kSpinTimeoutArrive = 1, kSpinTimeoutDone = 2, kSpinAborted = 3;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long v;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(v));
    return v;
}

// Thread q < world waits until peer q has published `epoch` in this rank's word `base + q`.  The wait is BOUNDED: a peer
// that never launches (its host raised, it skipped a step, its step count differs) must not hang every GPU of the node.
// After spin_budget_ns, or as soon as somebody raised the abort word, the wait gives up: the reason is latched in this
// rank's error word, the abort word of every peer is raised (so they stop at once instead of timing out one by one) and
// false is returned -- the caller then skips its work.  The host sees the error at its next ShardedFusedAdam.check().
__device__ __forceinline__ bool wait_for_peers(const ShardTable& t, int base, uint32_t timeout_code)
{
    bool ok = true;
    if (threadIdx.x < t.world) {
        const uint32_t* word = t.flags[t.rank] + base + threadIdx.x;
        const uint32_t* abort_word = t.flags[t.rank] + kFlagAbort;
        const unsigned long long t0 = global_ns();
        uint32_t spins = 0;
        while ((int32_t)(ld_acquire_sys(word) - t.epoch) < 0) {
            if ((++spins & 63u) != 0) continue;
            uint32_t why = 0;
            if (ld_acquire_sys(abort_word) != 0) why = kSpinAborted;
            else if (global_ns() - t0 > t.spin_budget_ns) why = timeout_code;
            if (why) {
                atomicCAS(t.flags[t.rank] + kFlagError, 0u, why);
                for (int r = 0; r < t.world; ++r) st_release_sys(t.flags[r] + kFlagAbort, t.epoch);
                ok = false;
                break;
            }
        }
    }
    return __syncthreads_and(ok) != 0;
}

__global__ void __launch_bounds__(kAdamThreads) adam_sharded_kernel(const __grid_constant__ ShardTable t)
{
    const bool synced = t.flags[0] != nullptr;
    if (synced) {
        // Arrive: this kernel runs after this rank's backward passes (stream order), so its bucket is complete; tell
        // every peer, then wait until every peer has said the same.  Every CTA polls this rank's OWN flag words only.
        if (blockIdx.x == 0 && threadIdx.x < t.world)
            st_release_sys(t.flags[threadIdx.x] + kFlagArrive + t.rank, t.epoch);
        if (!wait_for_peers(t, kFlagArrive, kSpinTimeoutArrive)) return;  // gave up: no update, error latched
    }
    int k = 0;
#pragma unroll 1
    for (int i = 1; i < t.n_seg; ++i)
        if (blockIdx.x >= t.first_chunk[i]) k = i;
    const ShardSeg& sg = t.seg[k];
    AdamTable at;  // only the scalars adam_one reads
    at.w1 = t.w1; at.beta2 = t.beta2; at.w2 = t.w2; at.eps = t.eps;
    AdamSeg as;
    as.neg_step_size = sg.neg_step_size; as.inv_bc2_sqrt = sg.inv_bc2_sqrt;
    const int64_t base = sg.begin + int64_t(blockIdx.x - t.first_chunk[k]) * kAdamChunk;
    // pieces start and end on 4-float boundaries (the host aligns shard bounds and group offsets), so float4 throughout
    for (int64_t i = base + 4 * threadIdx.x; i < min(base + (int64_t)kAdamChunk, sg.end); i += 4 * kAdamThreads) {
        float4 g = *reinterpret_cast<const float4*>(t.grads[0] + i);
#pragma unroll 1
        for (int r = 1; r < t.world; ++r) {  // fixed rank order: every replica sees the same sum
            const float4 x = *reinterpret_cast<const float4*>(t.grads[r] + i);
            g.x += x.x; g.y += x.y; g.z += x.z; g.w += x.w;
        }
        float4 p = *reinterpret_cast<const float4*>(t.params[t.rank] + i);
        float4 m = *reinterpret_cast<const float4*>(t.m + (i - t.lo));
        float4 v = *reinterpret_cast<const float4*>(t.v + (i - t.lo));
        adam_one(p.x, g.x, m.x, v.x, at, as);
        adam_one(p.y, g.y, m.y, v.y, at, as);
        adam_one(p.z, g.z, m.z, v.z, at, as);
        adam_one(p.w, g.w, m.w, v.w, at, as);
        *reinterpret_cast<float4*>(t.m + (i - t.lo)) = m;
        *reinterpret_cast<float4*>(t.v + (i - t.lo)) = v;
#pragma unroll 1
        for (int r = 0; r < t.world; ++r) *reinterpret_cast<float4*>(t.params[r] + i) = p;
    }
    if (synced) {
        // Done: when the last CTA of this rank has finished (its peer stores fenced), tell every peer; the kernel only
        // retires once every peer has said the same -- then every replica is written and every bucket has been read.
        __shared__ bool s_last;
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) s_last = atomicAdd(t.flags[t.rank] + kFlagCount, 1u) == gridDim.x - 1;
        __syncthreads();
        if (s_last) {
            if (threadIdx.x == 0) t.flags[t.rank][kFlagCount] = 0;  // for the next launch
            __threadfence_system();
            if (threadIdx.x < t.world) st_release_sys(t.flags[threadIdx.x] + kFlagDone + t.rank, t.epoch);
            (void)wait_for_peers(t, kFlagDone, kSpinTimeoutDone);
        }
    }
}

__global__ void sharded_abort_kernel(ShardTable t)
{
    if ((int)threadIdx.x < t.world) st_release_sys(t.flags[threadIdx.x] + kFlagAbort, 0xFFFFFFFFu);
}
unsigned long long spin_budget_ns()
{
    static const unsigned long long ns = [] {
        const char* e = getenv("GHR_SHARDED_TIMEOUT_MS");
        const double ms = e ? atof(e) : 20000.0;
        return (unsigned long long)((ms > 0 ? ms : 20000.0) * 1e6);
    }();
    return ns;
}

inline bool misaligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) != 0; }

}  // namespace
}  // namespace ghr

using namespace ghr;

extern "C" {

int ghr_adam_step(int n_groups, const ghr_adam_group* groups, double beta1, double beta2, double eps, int zero_grads,
                  void* stream)
{
    if (n_groups < 0 || n_groups > GHR_ADAM_MAX_GROUPS) return GHR_ERR_BAD_ARG;
    if (n_groups == 0) return GHR_OK;
    if (!groups) return GHR_ERR_BAD_ARG;
    if (!(beta1 >= 0.0 && beta1 < 1.0 && beta2 >= 0.0 && beta2 < 1.0 && eps >= 0.0)) return GHR_ERR_BAD_ARG;
    AdamTable t;
    t.n = 0;
    t.w1 = (float)(1.0 - beta1);  // torch: the lerp weight is the Python double 1 - beta1, rounded to fp32 once
    t.beta2 = (float)beta2;
    t.w2 = (float)(1.0 - beta2);
    t.eps = (float)eps;
    t.zero_grads = zero_grads ? 1 : 0;
    uint64_t chunks = 0;
    for (int i = 0; i < n_groups; ++i) {
        const ghr_adam_group& g = groups[i];
        if (g.count < 0 || g.step < 1) return GHR_ERR_BAD_ARG;
        if (g.count == 0) continue;
        if (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq) return GHR_ERR_BAD_ARG;
        AdamSeg& s = t.seg[t.n++];
        s.p = g.param; s.g = g.grad; s.m = g.exp_avg; s.v = g.exp_avg_sq;
        s.count = g.count;
        // bias corrections in double, like the Python scalars of torch/optim/adam.py
        const double bc1 = 1.0 - pow(beta1, (double)g.step);
        const double bc2 = 1.0 - pow(beta2, (double)g.step);
        const double bc2_sqrt = sqrt(bc2);
        s.neg_step_size = (float)(-g.lr / bc1);
        s.inv_bc2_sqrt = (float)(1.0 / bc2_sqrt);
        s.first_chunk = (uint32_t)chunks;
        s.vec_ok = !(misaligned16(g.param) || misaligned16(g.grad) || misaligned16(g.exp_avg) || misaligned16(g.exp_avg_sq));
        chunks += (uint64_t)((g.count + kAdamChunk - 1) / kAdamChunk);
        if (chunks >= (uint64_t(1) << 31)) return GHR_ERR_BAD_ARG;
    }
    if (t.n == 0) return GHR_OK;
    cudaStream_t s = (cudaStream_t)stream;
    StageTimer timer(GHR_STAGE_OPTIMIZER, s);
    adam_multi_kernel<<<(unsigned)chunks, kAdamThreads, 0, s>>>(t);
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "[libghr] CUDA error in adam_step: %s\n", cudaGetErrorString(e));
        return GHR_ERR_CUDA;
    }
    return GHR_OK;
}

int ghr_enable_peer_access(int peer_device)
{
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess || peer_device < 0) return GHR_ERR_BAD_ARG;
    if (cur == peer_device) return GHR_OK;
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, cur, peer_device) != cudaSuccess || !can) return GHR_ERR_CUDA;
    const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { (void)cudaGetLastError(); return GHR_OK; }
    if (e != cudaSuccess) {
        fprintf(stderr, "[libghr] cudaDeviceEnablePeerAccess(%d): %s\n", peer_device, cudaGetErrorString(e));
        return GHR_ERR_CUDA;
    }
    return GHR_OK;
}

int ghr_ipc_open(const void* handle, void** base_out)
{
    if (!handle || !base_out) return GHR_ERR_BAD_ARG;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    const cudaError_t e = cudaIpcOpenMemHandle(base_out, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        fprintf(stderr, "[libghr] cudaIpcOpenMemHandle: %s\n", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return GHR_ERR_CUDA;
    }
    return GHR_OK;
}

int ghr_ipc_close(void* base)
{
    if (!base) return GHR_OK;
    return cudaIpcCloseMemHandle(base) == cudaSuccess ? GHR_OK : GHR_ERR_CUDA;
}

static int adam_sharded_impl(int world, int rank, const float* const* peer_grads, float* const* peer_params,
                             float* exp_avg_shard, float* exp_avg_sq_shard, int64_t shard_lo, int64_t shard_hi, int n_groups,
                             const ghr_shard_group* groups, double beta1, double beta2, double eps,
                             uint32_t* const* peer_flags, uint32_t epoch, void* stream);

int ghr_adam_step_sharded(int world, int rank, const float* const* peer_grads, float* const* peer_params, float* exp_avg_shard,
                          float* exp_avg_sq_shard, int64_t shard_lo, int64_t shard_hi, int n_groups,
                          const ghr_shard_group* groups, double beta1, double beta2, double eps, void* stream)
{
    return adam_sharded_impl(world, rank, peer_grads, peer_params, exp_avg_shard, exp_avg_sq_shard, shard_lo, shard_hi,
                             n_groups, groups, beta1, beta2, eps, nullptr, 0, stream);
}

int ghr_adam_step_sharded_synced(int world, int rank, const float* const* peer_grads, float* const* peer_params,
                                 float* exp_avg_shard, float* exp_avg_sq_shard, int64_t shard_lo, int64_t shard_hi,
                                 int n_groups, const ghr_shard_group* groups, double beta1, double beta2, double eps,
                                 uint32_t* const* peer_flags, uint32_t epoch, void* stream)
{
    if (!peer_flags || epoch == 0) return GHR_ERR_BAD_ARG;
    return adam_sharded_impl(world, rank, peer_grads, peer_params, exp_avg_shard, exp_avg_sq_shard, shard_lo, shard_hi,
                             n_groups, groups, beta1, beta2, eps, peer_flags, epoch, stream);
}

int ghr_sharded_flag_words(void) { return 2 * GHR_MAX_PEERS + 8; }
int ghr_sharded_error_word(void) { return kFlagError; }


int ghr_sharded_abort(int world, uint32_t* const* peer_flags, void* stream)
{
    if (world < 1 || world > GHR_MAX_PEERS || !peer_flags) return GHR_ERR_BAD_ARG;
    ShardTable t;
    for (int r = 0; r < GHR_MAX_PEERS; ++r) {
        t.flags[r] = r < world ? peer_flags[r] : nullptr;
        if (r < world && !peer_flags[r]) return GHR_ERR_BAD_ARG;
    }
    t.world = world;
    sharded_abort_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(t);
    count_launch();
    return cudaGetLastError() == cudaSuccess ? GHR_OK : GHR_ERR_CUDA;
}

static int adam_sharded_impl(int world, int rank, const float* const* peer_grads, float* const* peer_params,
                             float* exp_avg_shard, float* exp_avg_sq_shard, int64_t shard_lo, int64_t shard_hi, int n_groups,
                             const ghr_shard_group* groups, double beta1, double beta2, double eps,
                             uint32_t* const* peer_flags, uint32_t epoch, void* stream)
{
    if (world < 1 || world > GHR_MAX_PEERS || rank < 0 || rank >= world || n_groups < 0 || n_groups > GHR_ADAM_MAX_GROUPS)
        return GHR_ERR_BAD_ARG;
    if (shard_lo < 0 || shard_hi < shard_lo || (shard_lo & 3) || (shard_hi & 3)) return GHR_ERR_BAD_ARG;
    const bool synced = peer_flags != nullptr;  // a rank with nothing to step still takes part in the barriers
    if ((shard_hi == shard_lo || n_groups == 0) && !synced) return GHR_OK;
    if (!peer_grads || !peer_params || !groups || !exp_avg_shard || !exp_avg_sq_shard || misaligned16(exp_avg_shard) ||
        misaligned16(exp_avg_sq_shard))
        return GHR_ERR_BAD_ARG;
    if (!(beta1 >= 0.0 && beta1 < 1.0 && beta2 >= 0.0 && beta2 < 1.0 && eps >= 0.0)) return GHR_ERR_BAD_ARG;
    ShardTable t;
    for (int r = 0; r < world; ++r) {
        if (!peer_grads[r] || !peer_params[r] || misaligned16(peer_grads[r]) || misaligned16(peer_params[r])) return GHR_ERR_BAD_ARG;
        t.grads[r] = peer_grads[r];
        t.params[r] = peer_params[r];
        if (synced && (!peer_flags[r] || (reinterpret_cast<uintptr_t>(peer_flags[r]) & 3))) return GHR_ERR_BAD_ARG;
    }
    for (int r = 0; r < GHR_MAX_PEERS; ++r) t.flags[r] = (synced && r < world) ? peer_flags[r] : nullptr;
    t.epoch = epoch;
    t.spin_budget_ns = spin_budget_ns();
    t.m = exp_avg_shard; t.v = exp_avg_sq_shard; t.lo = shard_lo;
    t.world = world; t.rank = rank; t.n_seg = 0;
    t.w1 = (float)(1.0 - beta1); t.beta2 = (float)beta2; t.w2 = (float)(1.0 - beta2); t.eps = (float)eps;
    uint64_t chunks = 0;
    for (int i = 0; i < n_groups; ++i) {
        const ghr_shard_group& g = groups[i];
        if (g.offset < 0 || g.count < 0 || (g.offset & 3) || g.step < 1) return GHR_ERR_BAD_ARG;
        const int64_t b = g.offset > shard_lo ? g.offset : shard_lo;
        int64_t e = g.offset + g.count < shard_hi ? g.offset + g.count : shard_hi;
        if (e <= b) continue;
        e = (e + 3) / 4 * 4;  // a group's tail is padded to the next group's 4-float boundary: the padding is ours to touch
        ShardSeg& sg = t.seg[t.n_seg];
        sg.begin = b; sg.end = e;
        const double bc1 = 1.0 - pow(beta1, (double)g.step), bc2 = 1.0 - pow(beta2, (double)g.step);
        sg.neg_step_size = (float)(-g.lr / bc1);
        sg.inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
        t.first_chunk[t.n_seg++] = (uint32_t)chunks;
        chunks += (uint64_t)((e - b + kAdamChunk - 1) / kAdamChunk);
    }
    if (t.n_seg == 0) {
        if (!synced) return GHR_OK;
        t.seg[0].begin = t.seg[0].end = 0;  // one CTA, no elements: barriers only
        t.first_chunk[0] = 0;
        chunks = 1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    StageTimer timer(GHR_STAGE_OPTIMIZER, s);
    adam_sharded_kernel<<<(unsigned)chunks, kAdamThreads, 0, s>>>(t);
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "[libghr] CUDA error in adam_step_sharded: %s\n", cudaGetErrorString(e));
        return GHR_ERR_CUDA;
    }
    return GHR_OK;
}

int ghr_densify_stats(int32_t P, const int32_t* radii, const float* dL_dmean2D, int32_t grad_stride, float* max_radii2D,
                      float* xyz_gradient_accum, float* denom, void* stream)
{
    if (P < 0 || grad_stride < 2) return GHR_ERR_BAD_ARG;
    if (P == 0) return GHR_OK;
    if (!radii || !dL_dmean2D) return GHR_ERR_BAD_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    StageTimer timer(GHR_STAGE_OPTIMIZER, s);
    densify_stats_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, radii, dL_dmean2D, grad_stride, max_radii2D,
                                                         xyz_gradient_accum, denom);
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "[libghr] CUDA error in densify_stats: %s\n", cudaGetErrorString(e));
        return GHR_ERR_CUDA;
    }
    return GHR_OK;
}

}  // extern "C"
