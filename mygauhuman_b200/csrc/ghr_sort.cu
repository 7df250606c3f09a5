// Hand-written onesweep LSD radix sort of (uint64 key, uint32 value) pairs on
// key bits [0,bits) -- replaces cub::DeviceRadixSort::SortPairs at
// reference rasterizer_impl.cu:305-310.  Stable, so the output is the unique
// stably-sorted sequence (bit-identical to CUB's).
//
//   digits           : 9 bits (512 bins), so the 43..45-bit depth|tile keys of
//                      a 512^2 .. 1080p frame take 5 passes where an 8-bit
//                      sort (CUB's SortPairs) takes 6
//   histogram kernel : one read of the keys, all passes' histograms; the last
//                      CTA to finish also does the plan: exclusive scans ->
//                      bin bases, marks passes whose digit is identical for
//                      every key as trivial (skipped: such a pass is the
//                      identity permutation), assigns ping-pong sources on
//                      the device (no host sync)
//   pass kernel (xK) : one sweep per non-trivial pass: per-warp ranking --
//                      MATCH.ANY where the warp's digits come in runs (the
//                      keys of a frame), one ballot per digit bit where
//                      they are all different (MATCH.ANY's latency grows
//                      with the distinct values) --, per-digit chained scan
//                      over tiles (decoupled look-back, 4 predecessors per
//                      round trip), then keys and values are put in digit
//                      order in shared memory and written out so that
//                      consecutive threads write consecutive addresses of a
//                      digit's run.  2 .. 12 keys per thread by size
//                      (sort_items_per_thread): a tile's fixed work -- 512
//                      digit scans and look-backs -- is what a large sort
//                      spreads over more keys.  DESIGN.md 5.4b has the
//                      measurements behind each choice.
//
// Roofline: HBM stream, algorithmic bytes N*(8 + 24*passes).
// The pair count N is read from device memory (no D2H sync) when n_dev is
// given; the grid is sized for `capacity` and surplus tiles exit.
#include <type_traits>
#include "ghr_common.cuh"
#include "ghr_internal.h"
#include "ghr_sort.cuh"

#ifndef GHR_SORT_RANK
#define GHR_SORT_RANK 2
#endif
#ifndef GHR_SORT_MATCH_RUNS
#define GHR_SORT_MATCH_RUNS 12  // of 31 neighbour pairs
#endif

namespace ghr {

constexpr uint32_t kStAgg = 1u << 30;
constexpr uint32_t kStIncl = 2u << 30;
constexpr uint32_t kStMask = (1u << 30) - 1u;
constexpr int kSortWarps = kSortThreads / 32;
static_assert(kSortThreads == kRadix, "one thread per digit in the scans");

__device__ __forceinline__ int64_t sort_count(const uint32_t* n_dev, int64_t capacity)
{
    return n_dev ? min((int64_t)*n_dev, capacity) : capacity;
}

// Exclusive scan of one value per thread over the 512-thread block.  s_warp: 17 words; two barriers inside.
// Back-to-back calls must alternate between two s_warp buffers.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* s_warp)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = lane < kSortWarps ? s_warp[lane] : 0u;
        uint32_t winc = w;
#pragma unroll
        for (int o = 1; o < kSortWarps; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += n;
        }
        if (lane < kSortWarps) s_warp[lane] = winc - w;
    }
    __syncthreads();
    return s_warp[warp] + inc - v;
}

// PASSES is a template parameter: the per-key digit loop is straight-line code with constant shifts (the kernel is
// issue-bound: at 1.2e7 keys and 5 passes 100 warp instructions per 32 keys with a run-time loop).
template <int PASSES>
__global__ void __launch_bounds__(kSortThreads) sort_histogram_kernel(const uint64_t* __restrict__ keys,
                                                                      const uint32_t* __restrict__ n_dev,
                                                                      int64_t capacity, int bits,
                                                                      SortHeader* hdr, const uint32_t* __restrict__ sq)
{
    constexpr int passes = PASSES;
    __shared__ uint32_t s_hist[kMaxPasses * kRadix];
    __shared__ uint32_t s_warp[2][kSortWarps + 1];
    __shared__ uint32_t s_trivial[kMaxPasses];
    __shared__ uint32_t s_last;
    const int64_t N = sort_count(n_dev, capacity);
    const Squeeze squeeze = load_squeeze(sq, bits);
    const int t = threadIdx.x;
    for (int i = t; i < passes * kRadix; i += kSortThreads) s_hist[i] = 0;
    __syncthreads();
    const int lastbits = bits - (passes - 1) * kRadixBits;
    const uint32_t lastmask = (1u << lastbits) - 1u;  // lastbits in [1, kRadixBits]
    // four independent key loads in flight per thread
    const int64_t stride = (int64_t)gridDim.x * kSortThreads;
    for (int64_t i0 = (int64_t)blockIdx.x * kSortThreads + t; i0 < N; i0 += 4 * stride) {
        uint64_t k[4];
#pragma unroll
        for (int u = 0; u < 4; u++) k[u] = (i0 + u * stride < N) ? keys[i0 + u * stride] : 0ull;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            if (i0 + u * stride >= N) break;
            const uint64_t ks = squeezed(k[u], squeeze);
#pragma unroll
            for (int p = 0; p < passes; p++) {
                uint32_t d = (uint32_t)(ks >> (p * kRadixBits)) & (uint32_t)(kRadix - 1);
                if (p == passes - 1) d &= lastmask;
                atomicAdd(&s_hist[p * kRadix + hist_col(d)], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = t; i < passes * kRadix; i += kSortThreads) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&hdr->hist[i / kRadix][hist_col(i % kRadix)], c);  // hist_col is its own inverse
    }
    // the last CTA to get here owns the complete histograms and writes the plan
    __threadfence();
    __syncthreads();
    if (t == 0) s_last = (atomicAdd(&hdr->hist_done, 1u) == gridDim.x - 1u);
    if (t < kMaxPasses) s_trivial[t] = 0;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int p = 0; p < passes; p++) {
        const uint32_t c = __ldcg(&hdr->hist[p][t]);
        if (c == (uint32_t)N) s_trivial[p] = 1;  // N == 0 also lands here (every pass trivial)
        hdr->bin_base[p][t] = block_excl_scan(c, s_warp[p & 1]);
    }
    __syncthreads();
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
__global__ void __launch_bounds__(kSortThreads, 2) onesweep_pass_kernel(uint64_t* keys_a, uint64_t* keys_b,
                                                                        uint32_t* vals_a, uint32_t* vals_b,
                                                                        const uint32_t* __restrict__ n_dev,
                                                                        int64_t capacity, int pass, uint32_t digit_mask,
                                                                        SortHeader* hdr,
                                                                        uint32_t* status /*[tiles][kRadix]*/,
                                                                        const uint32_t* __restrict__ sq, int bits)
{
    constexpr int TILE = kSortThreads * ITEMS;
    // hist: per-warp digit counts, then (after the scans) the tile-local position of each warp's first key of each
    // digit.  Once every key knows its position the same storage stages the keys; the values have their own, so that
    // keys and values are staged together before the look-back and written out together after it (five block barriers
    // per tile instead of seven, and the two global stores of a pair issue back to back).
    union Stage {
        uint32_t hist[kSortWarps][kRadix];
        uint64_t keys[TILE];
    };
    extern __shared__ __align__(16) unsigned char sm_raw[];  // sized in launch_pass
    Stage& sm = *reinterpret_cast<Stage*>(sm_raw);
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(sm_raw + sizeof(Stage));
    __shared__ uint32_t s_base[kRadix];  // global offset of the tile's digit run minus its tile-local start
    __shared__ uint32_t s_warp[kSortWarps + 1];
    __shared__ uint32_t s_tile;

    // Everything the tile needs from global memory before its keys is asked for at once -- the ticket, the plan's two
    // words, the key count, the squeeze -- instead of one dependent round trip after the other (a tile is ~8 us long and
    // two are resident per SM: each round trip saved is throughput).
    // (Measured and not kept: persistent CTAs, two per SM, that ask for the next tile's ticket while ranking the
    // current one -- 0.961 vs 0.918 ms at 2^24 keys, 662 vs 642 us on the 3 M-Gaussian frame.)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&hdr->ticket[pass], 1u);
    const uint32_t trivial = hdr->pass_trivial[pass];
    const uint32_t src = hdr->pass_src[pass];
    const int64_t N = sort_count(n_dev, capacity);
    const Squeeze squeeze = load_squeeze(sq, bits);
    {
        static_assert(kSortWarps * kRadix % (4 * kSortThreads) == 0, "counters are cleared four at a time");
        uint4* z = reinterpret_cast<uint4*>(&sm.hist[0][0]);
#pragma unroll
        for (int i = 0; i < kSortWarps * kRadix / (4 * kSortThreads); i++) z[i * kSortThreads + tid] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
    if (trivial) return;
    const uint64_t* __restrict__ kin = src ? keys_b : keys_a;
    const uint32_t* __restrict__ vin = src ? vals_b : vals_a;
    uint64_t* __restrict__ kout = src ? keys_a : keys_b;
    uint32_t* __restrict__ vout = src ? vals_a : vals_b;
    const uint32_t tile = s_tile;
    const int64_t tile_start = (int64_t)tile * TILE;
    if (tile_start >= N) return;
    const int valid = (int)min((int64_t)TILE, N - tile_start);
    const int shift = pass * kRadixBits;
    // The tile's work, specialised for a full tile (all but the last one): no bounds test on any key.
    auto run = [&](auto full_tile) {
    constexpr bool FULL = decltype(full_tile)::value;
    uint64_t key[ITEMS];
    uint32_t val[ITEMS];
    // Per key: first its rank among the warp's keys of the same digit, then its tile-local position (0xFFFF: no key).
    // Both fit 16 bits (a tile holds at most 8192 keys), and two share a register: with the digit recomputed from the
    // key instead of carried along, 16 keys per thread fit in 64 registers without a spill.
    uint32_t pos2[(ITEMS + 1) / 2] = {};
    auto get_pos = [&](int i) { return (i & 1) ? (pos2[i >> 1] >> 16) : (pos2[i >> 1] & 0xFFFFu); };
    auto set_pos = [&](int i, uint32_t v, bool first) {  // first: nothing of the register is in use yet
        if (i & 1) pos2[i >> 1] = (pos2[i >> 1] & 0xFFFFu) | (v << 16);
        else pos2[i >> 1] = first ? v : ((pos2[i >> 1] & 0xFFFF0000u) | v);
    };
    const int wbase = warp * (32 * ITEMS) + lane;
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const int idx = wbase + i * 32;
        key[i] = (FULL || idx < valid) ? kin[tile_start + idx] : ~0ull;
    }
    const uint32_t lt = lanemask_lt();
    // Rank of a key among the warp's keys of the same digit.  MATCH.ANY's latency grows with the number of distinct
    // values in the warp.  A warp of random 9-bit digits holds ~31 of them, and one ballot per digit bit is faster
    // (2^24 random keys: 1.10 ms with MATCH.ANY, 0.96 ms with the ballots).  The keys of a frame are the opposite case:
    // a Gaussian's tiles are emitted together and share their depth, the stable passes keep them together, and a warp
    // holds a handful of distinct depth digits (3 M Gaussians at 1080p: 671 us with MATCH.ANY, 703 us with the ballots).
    // So a warp looks at its first 32 keys -- how many equal their left neighbour's digit -- and takes MATCH.ANY when
    // the digits come in runs.  GHR_SORT_RANK: 0 always MATCH.ANY, 1 always ballots, 2 (default) per warp.
    // (Also measured, both slower: all eight peer masks before the first counter update, 1.00 ms; that plus one shared
    // atomicAdd per leader instead of its load + store, 1.02 ms.)
    auto digit_of = [&](int i, bool ok) {
        return ok ? ((uint32_t)(squeezed(key[i], squeeze) >> shift) & digit_mask) : (uint32_t)kRadix;
    };
    auto ballot_peers = [&](uint32_t d, bool ok) {
        uint32_t peers = 0xffffffffu;
        if (!FULL) { peers = __ballot_sync(0xffffffffu, ok); if (!ok) peers = ~peers; }
#pragma unroll
        for (int b = 0; b < kRadixBits; b++) {
            uint32_t v;  // the lanes whose digit agrees with this lane's in bit b
            asm volatile("{\n\t.reg .pred p;\n\t.reg .b32 t;\n\t"
                         "and.b32 t, %1, %2;\n\tsetp.ne.u32 p, t, 0;\n\t"
                         "vote.sync.ballot.b32 %0, p, 0xffffffff;\n\t@!p not.b32 %0, %0;\n\t}"
                         : "=r"(v) : "r"(d), "r"(1u << b));
            peers &= v;
        }
        return peers;
    };
#if GHR_SORT_RANK == 2
    bool runs;
    {
        const uint32_t d0 = digit_of(0, FULL || wbase < valid);
        const uint32_t left = __shfl_up_sync(0xffffffffu, d0, 1);
        runs = __popc(__ballot_sync(0xffffffffu, lane != 0 && d0 == left)) >= GHR_SORT_MATCH_RUNS;
    }
#endif
    auto peers_of = [&](uint32_t d, bool ok) {
#if GHR_SORT_RANK == 0
        return __match_any_sync(0xffffffffu, d);
#elif GHR_SORT_RANK == 1
        return ballot_peers(d, ok);
#else
        return runs ? __match_any_sync(0xffffffffu, d) : ballot_peers(d, ok);
#endif
    };
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const bool ok = FULL || (wbase + i * 32) < valid;
        const uint32_t d = digit_of(i, ok);
        const uint32_t peers = peers_of(d, ok);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (ok && lane == leader) {
            old = sm.hist[warp][hist_col(d)];
            sm.hist[warp][hist_col(d)] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        set_pos(i, old + __popc(peers & lt), true);
        __syncwarp();
    }
    __syncthreads();

    // thread t owns digit t: tile count, publish, tile-local digit starts
    uint32_t sum = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; w++) sum += sm.hist[w][hist_col(tid)];
    uint32_t* my = status + (size_t)tile * kRadix + tid;
    st_relaxed_u32(my, (tile == 0 ? kStIncl : kStAgg) | sum);
    const uint32_t bin_base = hdr->bin_base[pass][tid];  // asked for here, used after the look-back
    const uint32_t tile_excl = block_excl_scan(sum, s_warp);
    {
        uint32_t run = tile_excl;
#pragma unroll
        for (int w = 0; w < kSortWarps; w++) {
            const uint32_t c = sm.hist[w][hist_col(tid)];
            sm.hist[w][hist_col(tid)] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const bool ok = FULL || (wbase + i * 32) < valid;
        const uint32_t d = digit_of(i, ok);
        set_pos(i, ok ? sm.hist[warp][hist_col(d)] + get_pos(i) : 0xFFFFu, false);
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; i++)
        if (FULL || get_pos(i) != 0xFFFFu) sm.keys[get_pos(i)] = key[i];
    // The values are asked for only now, with the keys' registers free again (more keys per thread fit in 64
    // registers), and staged after the look-back, which hides their latency.
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const int idx = wbase + i * 32;
        val[i] = (FULL || idx < valid) ? vin[tile_start + idx] : 0u;
    }

    // decoupled look-back over the predecessors' counts of digit t, four tiles per round trip
    {
        uint32_t excl = 0;
        if (tile != 0) {
            int j = (int)tile - 1;
            bool done = false;
            // Four predecessors per round trip.  (Measured: 16 per trip -- fewer dependent L2 round trips for the last tile
            // of a small sort, where all tiles of a pass publish their aggregates at once -- is SLOWER at both ends:
            // 64 vs 58 us at 1.8e5 keys, 764 vs 690 us at 1.2e7 keys; the extra words cost more than the trips save.)
            auto consume = [&](auto look, const uint32_t* s) {
                constexpr int kLook = decltype(look)::value;
                bool stall = false;
#pragma unroll
                for (int u = 0; u < kLook; u++) {
                    const uint32_t f = s[u] & ~kStMask;
                    if (!done && !stall) {
                        if (f == 0) stall = true;
                        else {
                            excl += s[u] & kStMask;
                            if (f == kStIncl) done = true; else --j;
                        }
                    }
                }
            };
            auto walk = [&](auto look) {
                constexpr int kLook = decltype(look)::value;
                uint32_t s[kLook];
#pragma unroll
                for (int u = 0; u < kLook; u++)
                    s[u] = (j - u) >= 0 ? ld_relaxed_u32(status + (size_t)(j - u) * kRadix + tid) : kStIncl;
                consume(look, s);
            };
#ifndef GHR_SORT_LOOK
#define GHR_SORT_LOOK 4
#endif
            walk(std::integral_constant<int, 4>{});
            while (!done) walk(std::integral_constant<int, GHR_SORT_LOOK>{});
            st_relaxed_u32(my, kStIncl | (excl + sum));
        }
        s_base[tid] = bin_base + excl - tile_excl;
    }
#pragma unroll
    for (int i = 0; i < ITEMS; i++)
        if (FULL || get_pos(i) != 0xFFFFu) s_vals[get_pos(i)] = val[i];
    __syncthreads();

#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        const int idx = k * kSortThreads + tid;
        if (FULL || idx < valid) {
            const uint64_t kk = sm.keys[idx];
            const uint32_t dst = s_base[(uint32_t)(squeezed(kk, squeeze) >> shift) & digit_mask] + idx;
            kout[dst] = kk;
            vout[dst] = s_vals[idx];
        }
    }
    };
    if (valid == TILE) run(std::true_type{});
    else run(std::false_type{});
}

template <int ITEMS>
static void launch_pass(int64_t tiles, uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b,
                        const uint32_t* n_dev, int64_t capacity, int p, uint32_t mask, SortHeader* hdr, uint32_t* st,
                        const uint32_t* sq, int bits, cudaStream_t s)
{
    constexpr size_t tile_keys = (size_t)kSortThreads * ITEMS;
    constexpr size_t hist_bytes = sizeof(uint32_t) * kSortWarps * kRadix;
    constexpr size_t smem = (tile_keys * 8 > hist_bytes ? tile_keys * 8 : hist_bytes) + tile_keys * 4;
    if (smem + 4096 > 48 * 1024) {  // with the kernel's static 2.1 KB: beyond what a launch gets without asking
        static unsigned long long attr_set = 0ull;
        if (first_use_on_current_device(attr_set))
            cudaFuncSetAttribute(onesweep_pass_kernel<ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    onesweep_pass_kernel<ITEMS><<<(unsigned)tiles, kSortThreads, smem, s>>>(keys_a, keys_b, vals_a, vals_b, n_dev, capacity,
                                                                            p, mask, hdr, st, sq, bits);
}

cudaError_t launch_sort(SortHeader* hdr, uint32_t* status, const uint32_t* n_dev, int64_t capacity, int bits,
                        uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, cudaStream_t s,
                        const uint32_t* squeeze, bool histogram_done)
{
    if (capacity <= 0) return cudaSuccess;
    StageTimer timer(GHR_STAGE_SORT, s);
    const int passes = sort_passes_for(bits);
    const int items = sort_items_per_thread(capacity);
    const int64_t tiles = sort_tiles(capacity);
    int64_t hg = (capacity + kSortThreads * 4 - 1) / (kSortThreads * 4);
    const int hgrid = (int)(hg < sm_count() * 4 ? hg : sm_count() * 4);
    if (!histogram_done) {
        switch (passes) {
#define GHR_HIST_CASE(n) \
    case n: sort_histogram_kernel<n><<<hgrid, kSortThreads, 0, s>>>(keys_a, n_dev, capacity, bits, hdr, squeeze); break;
            GHR_HIST_CASE(1) GHR_HIST_CASE(2) GHR_HIST_CASE(3) GHR_HIST_CASE(4)
            GHR_HIST_CASE(5) GHR_HIST_CASE(6) GHR_HIST_CASE(7) GHR_HIST_CASE(8)
#undef GHR_HIST_CASE
            default: return cudaErrorInvalidValue;
        }
    }
    count_launch((histogram_done ? 0 : 1) + passes);
    for (int p = 0; p < passes; p++) {
        const int rem = bits - p * kRadixBits;
        const uint32_t mask = rem >= kRadixBits ? (uint32_t)(kRadix - 1) : ((1u << rem) - 1u);
        uint32_t* st = status + (size_t)p * tiles * kRadix;
        switch (items) {
            case 2: launch_pass<2>(tiles, keys_a, keys_b, vals_a, vals_b, n_dev, capacity, p, mask, hdr, st, squeeze, bits, s); break;
            case 12: launch_pass<12>(tiles, keys_a, keys_b, vals_a, vals_b, n_dev, capacity, p, mask, hdr, st, squeeze, bits, s); break;
            case 16: launch_pass<16>(tiles, keys_a, keys_b, vals_a, vals_b, n_dev, capacity, p, mask, hdr, st, squeeze, bits, s); break;
            case 4: launch_pass<4>(tiles, keys_a, keys_b, vals_a, vals_b, n_dev, capacity, p, mask, hdr, st, squeeze, bits, s); break;
            default: launch_pass<8>(tiles, keys_a, keys_b, vals_a, vals_b, n_dev, capacity, p, mask, hdr, st, squeeze, bits, s); break;
        }
    }
    return cudaGetLastError();
}

}  // namespace ghr
