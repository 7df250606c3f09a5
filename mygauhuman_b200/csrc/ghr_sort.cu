// Hand-written onesweep LSD radix sort of (uint64 key, uint32 value) pairs on
// key bits [0,bits) -- replaces cub::DeviceRadixSort::SortPairs at
// reference rasterizer_impl.cu:305-310.  Stable, so the output is the unique
// stably-sorted sequence (bit-identical to CUB's).
//
//   histogram kernel : one read of the keys, all passes' 256-bin histograms
//   plan kernel      : exclusive scans -> bin bases; marks passes whose digit
//                      is identical for every key as trivial (skipped: the
//                      permutation of such a pass is the identity); assigns
//                      ping-pong sources on the device (no host sync)
//   pass kernel (xK) : one sweep per non-trivial pass: per-tile ranking with
//                      warp match-any, chained-scan (decoupled look-back) of
//                      per-digit tile counts, scatter.
//
// Roofline: HBM stream, algorithmic bytes N*(8 + 24*passes).
// The pair count N is read from device memory (no D2H sync); the grid is sized
// for `capacity` and surplus tiles exit.
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

constexpr uint32_t kStAgg = 1u << 30;
constexpr uint32_t kStIncl = 2u << 30;
constexpr uint32_t kStMask = (1u << 30) - 1u;

__global__ void __launch_bounds__(256) sort_histogram_kernel(const uint64_t* __restrict__ keys,
                                                             const uint32_t* __restrict__ n_dev, int64_t capacity,
                                                             int bits, int passes, SortHeader* hdr)
{
    __shared__ uint32_t s_hist[kMaxPasses * kRadix];
    const int64_t N = min((int64_t)*n_dev, capacity);
    for (int i = threadIdx.x; i < passes * kRadix; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const uint32_t lastmask = (bits - (passes - 1) * kRadixBits) >= kRadixBits
                                  ? 0xFFu
                                  : ((1u << (bits - (passes - 1) * kRadixBits)) - 1u);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t k = keys[i];
        for (int p = 0; p < passes; p++) {
            uint32_t d = (uint32_t)(k >> (p * kRadixBits)) & 0xFFu;
            if (p == passes - 1) d &= lastmask;
            atomicAdd(&s_hist[p * kRadix + d], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * kRadix; i += blockDim.x) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&hdr->hist[i / kRadix][i % kRadix], c);
    }
}

// one block of 256 threads
__global__ void __launch_bounds__(256) sort_plan_kernel(const uint32_t* __restrict__ n_dev, int64_t capacity,
                                                        int passes, SortHeader* hdr)
{
    __shared__ uint32_t s_scan[kRadix];
    __shared__ uint32_t s_trivial[kMaxPasses];
    const uint32_t N = (uint32_t)min((int64_t)*n_dev, capacity);
    const int t = threadIdx.x;
    if (t < kMaxPasses) s_trivial[t] = 0;
    __syncthreads();
    for (int p = 0; p < passes; p++) {
        const uint32_t c = hdr->hist[p][t];
        if (c == N) s_trivial[p] = 1;  // N == 0 also lands here (every pass trivial)
        // exclusive scan over 256 bins (Hillis-Steele in smem)
        s_scan[t] = c;
        __syncthreads();
        for (int o = 1; o < kRadix; o <<= 1) {
            const uint32_t v = (t >= o) ? s_scan[t - o] : 0;
            __syncthreads();
            s_scan[t] += v;
            __syncthreads();
        }
        hdr->bin_base[p][t] = s_scan[t] - c;
        __syncthreads();
    }
    if (t == 0) {
        uint32_t sel = 0;
        for (int p = 0; p < passes; p++) {
            hdr->pass_trivial[p] = s_trivial[p];
            hdr->pass_src[p] = sel;
            if (!s_trivial[p]) sel ^= 1u;
        }
        hdr->sorted_sel = sel;
        hdr->num_passes = passes;
    }
}

template <int ITEMS>
__global__ void __launch_bounds__(kSortThreads) onesweep_pass_kernel(uint64_t* keys_a, uint64_t* keys_b,
                                                                     uint32_t* vals_a, uint32_t* vals_b,
                                                                     const uint32_t* __restrict__ n_dev,
                                                                     int64_t capacity, int pass, uint32_t digit_mask,
                                                                     SortHeader* hdr, uint32_t* status /*[tiles][256]*/)
{
    constexpr int TILE = kSortThreads * ITEMS;
    __shared__ uint32_t s_hist[8][kRadix];  // per-warp digit counts -> per-warp exclusive offsets
    __shared__ uint32_t s_base[kRadix];     // global output offset of this tile's first key of each digit
    __shared__ uint32_t s_tile;

    if (hdr->pass_trivial[pass]) return;
    const uint32_t src = hdr->pass_src[pass];
    const uint64_t* __restrict__ kin = src ? keys_b : keys_a;
    const uint32_t* __restrict__ vin = src ? vals_b : vals_a;
    uint64_t* __restrict__ kout = src ? keys_a : keys_b;
    uint32_t* __restrict__ vout = src ? vals_a : vals_b;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&hdr->ticket[pass], 1u);
    for (int i = tid; i < 8 * kRadix; i += kSortThreads) (&s_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const int64_t N = min((int64_t)*n_dev, capacity);
    const int64_t tile_start = (int64_t)tile * TILE;
    if (tile_start >= N) return;

    const int shift = pass * kRadixBits;
    uint64_t key[ITEMS];
    uint32_t val[ITEMS];
    uint32_t rank[ITEMS];  // low 16: rank inside the warp for its digit; high 16: digit (256 = invalid)
    const int64_t wbase = tile_start + (int64_t)warp * (32 * ITEMS) + lane;
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const int64_t idx = wbase + i * 32;
        if (idx < N) { key[i] = kin[idx]; val[i] = vin[idx]; }
        else { key[i] = ~0ull; val[i] = 0; }
    }
    const uint32_t lt = lanemask_lt();
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const bool valid = (wbase + i * 32) < N;
        const uint32_t d = valid ? ((uint32_t)(key[i] >> shift) & digit_mask) : 256u;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = s_hist[warp][d];
            s_hist[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[i] = (d << 16) | (old + __popc(peers & lt));
        __syncwarp();
    }
    __syncthreads();

    // thread t owns digit t: exclusive scan over the 8 warps, publish, look back
    {
        uint32_t sum = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            const uint32_t c = s_hist[w][tid];
            s_hist[w][tid] = sum;
            sum += c;
        }
        uint32_t* my = status + (size_t)tile * kRadix + tid;
        uint32_t excl = 0;
        if (tile == 0) {
            st_relaxed_u32(my, kStIncl | sum);
        } else {
            st_relaxed_u32(my, kStAgg | sum);
            int j = (int)tile - 1;
            while (true) {
                const uint32_t s = ld_relaxed_u32(status + (size_t)j * kRadix + tid);
                const uint32_t f = s & ~kStMask;
                if (f == 0) continue;
                excl += s & kStMask;
                if (f == kStIncl) break;
                --j;
            }
            st_relaxed_u32(my, kStIncl | (excl + sum));
        }
        s_base[tid] = hdr->bin_base[pass][tid] + excl;
    }
    __syncthreads();

#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const uint32_t d = rank[i] >> 16;
        if (d < 256u) {
            const uint32_t dst = s_base[d] + s_hist[warp][d] + (rank[i] & 0xFFFFu);
            kout[dst] = key[i];
            vout[dst] = val[i];
        }
    }
}

cudaError_t launch_sort(SortHeader* hdr, uint32_t* status, const uint32_t* n_dev, int64_t capacity, int bits,
                        uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, cudaStream_t s)
{
    if (capacity <= 0) return cudaSuccess;
    StageTimer timer(GHR_STAGE_SORT, s);
    const int passes = (bits + kRadixBits - 1) / kRadixBits;
    const int items = sort_items_per_thread(capacity);
    const int64_t tiles = sort_tiles(capacity);
    int64_t hg = (capacity + 256 * 8 - 1) / (256 * 8);
    const int hgrid = (int)(hg < 148 * 8 ? hg : 148 * 8);
    sort_histogram_kernel<<<hgrid, 256, 0, s>>>(keys_a, n_dev, capacity, bits, passes, hdr);
    sort_plan_kernel<<<1, 256, 0, s>>>(n_dev, capacity, passes, hdr);
    count_launch(2 + passes);
    for (int p = 0; p < passes; p++) {
        const int rem = bits - p * kRadixBits;
        const uint32_t mask = rem >= kRadixBits ? 0xFFu : ((1u << rem) - 1u);
        uint32_t* st = status + (size_t)p * tiles * kRadix;
        if (items == 4)
            onesweep_pass_kernel<4><<<(unsigned)tiles, kSortThreads, 0, s>>>(keys_a, keys_b, vals_a, vals_b, n_dev,
                                                                             capacity, p, mask, hdr, st);
        else
            onesweep_pass_kernel<16><<<(unsigned)tiles, kSortThreads, 0, s>>>(keys_a, keys_b, vals_a, vals_b, n_dev,
                                                                              capacity, p, mask, hdr, st);
    }
    return cudaGetLastError();
}

}  // namespace ghr
