// Key emission (reference rasterizer_impl.cu:70-111 duplicateWithKeys) and
// tile range detection (rasterizer_impl.cu:116-138 identifyTileRanges).
// HBM-stream kernels: emission 8*P + 12*V + 12*N bytes, ranges 8*N + 8*tiles.
#include "ghr_common.cuh"
#include "ghr_internal.h"
#include "ghr_sort.cuh"

namespace ghr {

// One warp handles 32 Gaussians and writes their pairs -- one contiguous range of the list -- cooperatively (see the
// comment in the kernel).  The emitted order (y-major, then x, starting at point_offsets[idx-1]) is the reference's.
__global__ void __launch_bounds__(GHR_BLOCK) emit_keys_kernel(int P, int W, int H,
                                                              const uint32_t* __restrict__ tiles_touched,
                                                              const float4* __restrict__ rec,
                                                              const uint32_t* __restrict__ point_offsets,
                                                              const ghr_geom_header* __restrict__ hdr,
                                                              int64_t capacity, uint64_t* __restrict__ keys,
                                                              uint32_t* __restrict__ vals, int pack_masks,
                                                              SortHeader* __restrict__ sort_hdr, int bits, int passes)
{
    // sort_hdr != NULL: this kernel also does the work of sort_histogram_kernel (ghr_sort.cu) -- the digit histograms of
    // all passes over the keys it writes, and, in the last CTA to finish, the plan (bin bases, trivial passes, ping-pong
    // sources) -- so the frame's sort starts with its first pass (one launch and one read of the keys fewer).
    __shared__ uint32_t s_hist[kMaxPasses * kRadix];
    __shared__ uint32_t s_scan[9];
    __shared__ uint32_t s_trivial[kMaxPasses];
    __shared__ uint32_t s_last;
    if ((int64_t)hdr->num_rendered > capacity) return;  // overflow is latched by the ranges kernel
    Squeeze squeeze;
    squeeze.lo = 0u; squeeze.L = 32u;
    uint32_t lastmask = 0;
    if (sort_hdr) {
        for (int i = threadIdx.x; i < passes * kRadix; i += GHR_BLOCK) s_hist[i] = 0;
        squeeze = load_squeeze(&hdr->depth_hi, bits);
        lastmask = (1u << (bits - (passes - 1) * kRadixBits)) - 1u;
        __syncthreads();
    }
    const int idx = blockIdx.x * GHR_BLOCK + threadIdx.x;
    const uint32_t lane = lane_id();
    const uint32_t gx = (W + GHR_TILE - 1) / GHR_TILE, gy = (H + GHR_TILE - 1) / GHR_TILE;
    uint32_t minx = 0, miny = 0, maxx = 0, maxy = 0, dbits = 0, ext = 0;
    float cpx = 0.f, cpy = 0.f;
    // tiles_touched > 0  <=>  radii > 0 (preprocess stores both or neither)
    const bool visible = idx < P && tiles_touched[idx] > 0;
    if (visible) {
        const float4 r0 = rec[3 * (size_t)idx];
        get_rect(r0.x, r0.y, __float_as_int(r0.w), gx, gy, minx, miny, maxx, maxy);
        dbits = __float_as_uint(r0.z);
        cpx = r0.x; cpy = r0.y;
        if (pack_masks) ext = __float_as_uint(rec[3 * (size_t)idx + 2].w);
    }
    // where this Gaussian's pairs start: the exclusive scan of tiles_touched (monotone, so defined for culled ones too)
    const int last = min(idx, P) - 1;
    const uint32_t off = last < 0 ? 0u : point_offsets[last];
    const uint32_t w = maxx - minx;
    const uint32_t cnt = w * (maxy - miny);
    // k / w for the slots of this Gaussian as one multiply-high: magic = ceil(2^32 / w) is exact for k * w < 2^32
    // (k < the Gaussian's tile count).  w == 1 would need 33 bits: the loop takes k itself.
    const uint32_t magic = w >= 2u ? 0xFFFFFFFFu / w + 1u : 0u;
    const uint32_t minxy = minx | (miny << 16);
    // The 32 Gaussians of a warp own ONE contiguous range of the pair list.  The warp writes it slot by slot -- lane l takes
    // slots l, l + 32, ... -- so every store instruction covers 32 consecutive pairs (256 + 128 contiguous bytes) whatever
    // the Gaussians' tile counts are; the owner of a slot is found by a 5-step search over the lanes' start offsets.
    // (One thread per Gaussian writing its own pairs issued up to 31 scattered store pairs per thread: ncu at 12 M pairs,
    // lg_throttle 7 + long_scoreboard 16 warps per issued instruction.)  Emitted order: the reference's (y-major, then x).
    const uint32_t base = __shfl_sync(0xffffffffu, off, 0);
    const uint32_t rel = off - base;
    const uint32_t total = __shfl_sync(0xffffffffu, rel + cnt, 31);
    for (uint32_t j0 = 0; j0 < total; j0 += 32) {
        const uint32_t j = j0 + lane;
        uint32_t lo = 0, hi = 31;
#pragma unroll
        for (int it = 0; it < 5; it++) {
            const uint32_t mid = (lo + hi + 1) >> 1;
            const uint32_t v = __shfl_sync(0xffffffffu, rel, mid);
            if (v <= j) lo = mid; else hi = mid - 1;
        }
        const uint32_t o_rel = __shfl_sync(0xffffffffu, rel, lo), o_w = __shfl_sync(0xffffffffu, w, lo);
        const uint32_t o_magic = __shfl_sync(0xffffffffu, magic, lo), o_minxy = __shfl_sync(0xffffffffu, minxy, lo);
        const uint32_t o_d = __shfl_sync(0xffffffffu, dbits, lo);
        float o_px = 0.f, o_py = 0.f;
        uint32_t o_ext = 0;
        if (pack_masks) {
            o_px = __shfl_sync(0xffffffffu, cpx, lo); o_py = __shfl_sync(0xffffffffu, cpy, lo);
            o_ext = __shfl_sync(0xffffffffu, ext, lo);
        }
        if (j < total) {
            const uint32_t k = j - o_rel;
            const uint32_t row = o_w == 1u ? k : __umulhi(k, o_magic);
            const uint32_t ty = (o_minxy >> 16) + row, tx = (o_minxy & 0xFFFFu) + (k - row * o_w);
            const uint64_t key = ((uint64_t)(ty * gx + tx) << 32) | o_d;
            keys[base + j] = key;
            if (sort_hdr) {
                const uint64_t ks = squeezed(key, squeeze);
                for (int p = 0; p < passes; p++) {
                    uint32_t d = (uint32_t)(ks >> (p * kRadixBits)) & (uint32_t)(kRadix - 1);
                    if (p == passes - 1) d &= lastmask;
                    atomicAdd(&s_hist[p * kRadix + hist_col(d)], 1u);
                }
            }
            uint32_t v = (uint32_t)(idx - (int)lane) + lo;
            if (pack_masks) {
                // bit wb of the mask: the conservative bounding-box cull of block wb (8 wide, 4 high; wb = 2 * row + column)
                // of this tile -- the very test blend_forward_wide_kernel ran per pair and block warp
                float ex, ey;
                unpack_cull_extent(o_ext, ex, ey);
                // The same values in fewer instructions (this kernel is issue-bound at 1.2e7 pairs): the block edges are
                // the tile origin, converted once, plus small constants (exact in fp32), and
                // max(a, b, 0) > e  <=>  a > e || b > e  for the non-negative extents.
                const float fx = (float)(tx * GHR_TILE), fy = (float)(ty * GHR_TILE);
                uint32_t mx = 0, rows = 0;  // rows: bit 2r set when block row r passes
#pragma unroll
                for (uint32_t c = 0; c < 2; c++) {
                    const float rx0 = fx + (float)(c * 8u), rx1 = fx + (float)(c * 8u + 7u);
                    mx |= ((rx0 - o_px > ex) || (o_px - rx1 > ex)) ? 0u : (1u << c);
                }
#pragma unroll
                for (uint32_t r = 0; r < 4; r++) {
                    const float ry0 = fy + (float)(r * 4u), ry1 = fy + (float)(r * 4u + 3u);
                    rows |= ((ry0 - o_py > ey) || (o_py - ry1 > ey)) ? 0u : (1u << (2u * r));
                }
                v |= (rows * mx) << 24;  // mx in 0..3 lands in every passing row's two bits, no carries
            }
            vals[base + j] = v;
        }
    }
    if (!sort_hdr) return;
    __syncthreads();
    for (int i = threadIdx.x; i < passes * kRadix; i += GHR_BLOCK) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&sort_hdr->hist[i / kRadix][hist_col(i % kRadix)], c);  // hist_col is its own inverse
    }
    // the last CTA to get here owns the complete histograms and writes the plan (as sort_histogram_kernel does)
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(&sort_hdr->hist_done, 1u) == gridDim.x - 1u);
    if (threadIdx.x < kMaxPasses) s_trivial[threadIdx.x] = 0;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const uint32_t N = hdr->num_rendered;
    const int t = threadIdx.x;  // owns digits 2t and 2t + 1
    for (int p = 0; p < passes; p++) {
        const uint32_t c0 = __ldcg(&sort_hdr->hist[p][2 * t]), c1 = __ldcg(&sort_hdr->hist[p][2 * t + 1]);
        if (c0 == N || c1 == N) s_trivial[p] = 1;  // N == 0 also lands here (every pass trivial)
        // exclusive scan of c0 + c1 over the 256 threads
        uint32_t inc = c0 + c1;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
            if ((int)lane >= o) inc += n;
        }
        if (lane == 31) s_scan[threadIdx.x >> 5] = inc;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t run = 0;
            for (int w = 0; w < GHR_BLOCK / 32; w++) { const uint32_t x = s_scan[w]; s_scan[w] = run; run += x; }
        }
        __syncthreads();
        const uint32_t excl = s_scan[threadIdx.x >> 5] + inc - (c0 + c1);
        sort_hdr->bin_base[p][2 * t] = excl;
        sort_hdr->bin_base[p][2 * t + 1] = excl + c0;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        uint32_t sel = 0;
        for (int p = 0; p < passes; p++) {
            sort_hdr->pass_trivial[p] = s_trivial[p];
            sort_hdr->pass_src[p] = sel;
            if (!s_trivial[p]) sel ^= 1u;
        }
        sort_hdr->sorted_sel = sel;
        sort_hdr->num_passes = passes;
    }
}

// Tile ranges (reference identifyTileRanges, rasterizer_impl.cu:126-152) and, per list entry, the blend kernels'
// staging word {Gaussian id, block mask}: bit (4*by + bx) of the mask is set when the Gaussian may contribute to
// some pixel of the tile's 4x4-pixel block (bx, by) -- the conservative bounding-box test followed by the exact
// ellipse-vs-block test (ghr_common.cuh).  The blend kernels (sixteen or eight work items per tile) would otherwise
// repeat this test once per item; here it is done once per entry.  The words go into the key buffer the sort left
// unused (8 bytes per entry, exactly).
__global__ void __launch_bounds__(GHR_BLOCK) tile_ranges_kernel(ghr_geom_header* hdr, int64_t capacity, int gx,
                                                                const SortHeader* __restrict__ sh,
                                                                uint64_t* keys_a, uint64_t* keys_b,
                                                                const uint32_t* __restrict__ vals_a,
                                                                const uint32_t* __restrict__ vals_b,
                                                                const float4* __restrict__ rec, uint2* ranges,
                                                                int with_masks)
{
    const int64_t L = hdr->num_rendered;
    if (L > capacity) {
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicMax(&hdr->error, (uint32_t)(-GHR_ERR_CAPACITY));
        return;
    }
    // a render stage re-run with a large enough buffer (optimistic launch, _C.py) clears the latch the first try left
    if (blockIdx.x == 0 && threadIdx.x == 0 && hdr->error == (uint32_t)(-GHR_ERR_CAPACITY)) hdr->error = 0;
    const uint32_t sel = sh->sorted_sel;
    const uint64_t* __restrict__ keys = sel ? keys_b : keys_a;
    const uint32_t* __restrict__ vals = sel ? vals_b : vals_a;
    uint2* __restrict__ entries = reinterpret_cast<uint2*>(sel ? keys_a : keys_b);
    for (int64_t idx = (int64_t)blockIdx.x * GHR_BLOCK + threadIdx.x; idx < L; idx += (int64_t)gridDim.x * GHR_BLOCK) {
        const uint32_t cur = (uint32_t)(keys[idx] >> 32);
        if (idx == 0) ranges[cur].x = 0;
        else {
            const uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
            if (cur != prev) { ranges[prev].y = (uint32_t)idx; ranges[cur].x = (uint32_t)idx; }
        }
        if (idx == L - 1) ranges[cur].y = (uint32_t)L;

        if (!with_masks) continue;  // throughput regime: the wide blend kernels cull inside (ghr_blend_fwd_wide.cu)
        const uint32_t id = vals[idx];
        const float4 r0 = rec[3 * (size_t)id], r1 = rec[3 * (size_t)id + 1];
        float ex, ey;
        unpack_cull_extent(__float_as_uint(rec[3 * (size_t)id + 2].w), ex, ey);
        const float tx0 = (float)((cur % (uint32_t)gx) * GHR_TILE), ty0 = (float)((cur / (uint32_t)gx) * GHR_TILE);
        const CullQuad quad = cull_quad(r1.x, r1.y, r1.z, r1.w);
        uint32_t mask = 0;
#pragma unroll
        for (int by = 0; by < 4; by++) {
            const float ry0 = ty0 + 4.0f * by, ry1 = ry0 + 3.0f;
            const float ddy = fmaxf(fmaxf(ry0 - r0.y, r0.y - ry1), 0.0f);
            if (ddy > ey) continue;
#pragma unroll
            for (int bx = 0; bx < 4; bx++) {
                const float rx0 = tx0 + 4.0f * bx, rx1 = rx0 + 3.0f;
                const float ddx = fmaxf(fmaxf(rx0 - r0.x, r0.x - rx1), 0.0f);
                if (ddx > ex) continue;
                if (cull_keep_exact(quad, r0.x, r0.y, rx0, rx1, ry0, ry1)) mask |= 1u << (4 * by + bx);
            }
        }
        entries[idx] = make_uint2(id, mask);
    }
}

// Throughput regime (the wide blend kernels cull inside, no block masks): tile boundaries only -- a pure stream over the
// sorted keys.  Four consecutive keys per thread per step (two 16-byte loads in flight + the predecessor's key): the
// one-key-per-iteration loop above sat on long_scoreboard (35 warps per issued instruction at 12 M pairs, 1.7 TB/s).
__global__ void __launch_bounds__(GHR_BLOCK) tile_ranges_only_kernel(ghr_geom_header* hdr, int64_t capacity,
                                                                     const SortHeader* __restrict__ sh,
                                                                     const uint64_t* __restrict__ keys_a,
                                                                     const uint64_t* __restrict__ keys_b, uint2* ranges)
{
    const int64_t L = hdr->num_rendered;
    if (L > capacity) {
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicMax(&hdr->error, (uint32_t)(-GHR_ERR_CAPACITY));
        return;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && hdr->error == (uint32_t)(-GHR_ERR_CAPACITY)) hdr->error = 0;
    const uint64_t* __restrict__ keys = sh->sorted_sel ? keys_b : keys_a;
    const int64_t stride = (int64_t)gridDim.x * GHR_BLOCK * 4;
    for (int64_t i0 = ((int64_t)blockIdx.x * GHR_BLOCK + threadIdx.x) * 4; i0 < L; i0 += stride) {
        uint32_t t[5];  // tile ids of entries i0-1 .. i0+3
        if (i0 + 4 <= L) {
            const ulonglong2 a = *reinterpret_cast<const ulonglong2*>(keys + i0);
            const ulonglong2 b = *reinterpret_cast<const ulonglong2*>(keys + i0 + 2);
            t[1] = (uint32_t)(a.x >> 32); t[2] = (uint32_t)(a.y >> 32); t[3] = (uint32_t)(b.x >> 32); t[4] = (uint32_t)(b.y >> 32);
        } else {
#pragma unroll
            for (int u = 0; u < 4; u++) t[1 + u] = i0 + u < L ? (uint32_t)(keys[i0 + u] >> 32) : 0u;
        }
        t[0] = i0 ? (uint32_t)(keys[i0 - 1] >> 32) : 0u;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int64_t idx = i0 + u;
            if (idx >= L) break;
            const uint32_t cur = t[1 + u];
            if (idx == 0) ranges[cur].x = 0;
            else if (cur != t[u]) { ranges[t[u]].y = (uint32_t)idx; ranges[cur].x = (uint32_t)idx; }
            if (idx == L - 1) ranges[cur].y = (uint32_t)L;
        }
    }
}

// (Tried: running this ordering in the last CTA of the tile-range kernel instead of as a launch of its own.  Measured
// slower on B200, 7 095 vs 7 240 views/s in the 8-lane step: the one-CTA kernel overlaps with the other views' work,
// the merged tail does not.)
// Tile ids ordered by decreasing list length (power-of-two buckets): the blend kernels pull (tile, block)
// work items from a queue in this order, so the few heavy tiles of a human-shaped scene start first and
// spread over all SMs (longest-processing-time-first) instead of landing wherever blockIdx order puts them.
__global__ void __launch_bounds__(1024) tile_order_kernel(int tiles, const uint2* __restrict__ ranges,
                                                          const ghr_geom_header* __restrict__ hdr, int64_t capacity,
                                                          uint32_t* __restrict__ order)
{
    __shared__ uint32_t s_hist[33], s_base[33];
    const bool overflow = (int64_t)hdr->num_rendered > capacity;
    if (threadIdx.x < 33) s_hist[threadIdx.x] = 0;
    // Up to 8192 tiles (1080p has 8160) every thread keeps its tiles' buckets in registers: the ranges are read once,
    // all loads in flight together, instead of two passes of one dependent load each per 1024 tiles.
    constexpr int kHeld = 8;
    const bool held = tiles <= kHeld * 1024;
    uint32_t bucket[kHeld];
    if (held) {
        uint2 r[kHeld];
#pragma unroll
        for (int k = 0; k < kHeld; k++) {
            const int t = k * 1024 + threadIdx.x;
            r[k] = t < tiles ? ranges[t] : make_uint2(0u, 0u);  // (not waiting for the overflow word)
        }
#pragma unroll
        for (int k = 0; k < kHeld; k++) {
            const uint32_t n = overflow ? 0u : r[k].y - r[k].x;
            bucket[k] = n ? 32 - __clz(n) : 0;  // bucket b holds lengths in [2^(b-1), 2^b)
        }
    }
    __syncthreads();
    // A frame has a dozen distinct buckets: one shared-memory atomic per tile serialises 32-fold on the same word (ncu at
    // 8160 tiles: 10 us, short_scoreboard).  The lanes of a warp that share a bucket are counted by their leader.
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    if (held) {
#pragma unroll
        for (int k = 0; k < kHeld; k++) {
            if (k * 1024 + (int)(threadIdx.x & ~31u) >= tiles) break;  // whole warp past the end
            const bool in = k * 1024 + (int)threadIdx.x < tiles;
            const uint32_t key = in ? bucket[k] : 33u;
            const uint32_t peers = __match_any_sync(0xffffffffu, key);
            if (in && (peers & lt) == 0u) atomicAdd(&s_hist[key], (uint32_t)__popc(peers));
        }
    } else {
        for (int t = threadIdx.x; t < tiles; t += 1024) {
            const uint32_t n = overflow ? 0u : ranges[t].y - ranges[t].x;
            atomicAdd(&s_hist[n ? 32 - __clz(n) : 0], 1u);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t acc = 0;
        for (int b = 32; b >= 0; b--) { s_base[b] = acc; acc += s_hist[b]; }
    }
    __syncthreads();
    if (held) {
#pragma unroll
        for (int k = 0; k < kHeld; k++) {
            if (k * 1024 + (int)(threadIdx.x & ~31u) >= tiles) break;
            const int t = k * 1024 + threadIdx.x;
            const bool in = t < tiles;
            const uint32_t key = in ? bucket[k] : 33u;
            const uint32_t peers = __match_any_sync(0xffffffffu, key);
            const int leader = __ffs(peers) - 1;
            uint32_t base = 0;
            if (in && (int)lane == leader) base = atomicAdd(&s_base[key], (uint32_t)__popc(peers));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (in) order[base + __popc(peers & lt)] = (uint32_t)t;
        }
    } else {
        for (int t = threadIdx.x; t < tiles; t += 1024) {
            const uint32_t n = overflow ? 0u : ranges[t].y - ranges[t].x;
            order[atomicAdd(&s_base[n ? 32 - __clz(n) : 0], 1u)] = (uint32_t)t;
        }
    }
}

__global__ void __launch_bounds__(256) recolor_kernel(int P, const float* __restrict__ colors, float4* __restrict__ rec)
{
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= P) return;
    float4 r2 = rec[3 * (size_t)idx + 2];
    r2.x = colors[3 * (size_t)idx]; r2.y = colors[3 * (size_t)idx + 1]; r2.z = colors[3 * (size_t)idx + 2];
    rec[3 * (size_t)idx + 2] = r2;  // .w (cull extent) kept
}

// packed[idx][attr_offset(n,k) + c] <-> attribute k, colour c (AttrLayout, ghr_internal.h); padding floats are zeroed
__global__ void __launch_bounds__(256) pack_attributes_kernel(int P, int n, int stride, AttrPtrs ptrs, float* __restrict__ packed)
{
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= P) return;
    float* row = packed + (size_t)idx * stride;
    const int half = stride / 2, used = 3 * n / 2;
    for (int c = used; c < half; c++) { row[c] = 0.f; row[half + c] = 0.f; }
    for (int k = 0; k < n; k++) {
        const float* src = ptrs.src[k] + 3 * (size_t)idx;
        float* dst = row + (k / (n / 2)) * half + 3 * (k % (n / 2));
        dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
    }
}

__global__ void __launch_bounds__(256) unpack_attributes_kernel(int P, int n, int stride, const float* __restrict__ packed,
                                                                AttrPtrs ptrs)
{
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= P) return;
    const float* row = packed + (size_t)idx * stride;
    const int half = stride / 2;
    for (int k = 0; k < n; k++) {
        if (!ptrs.dst[k]) continue;
        const float* src = row + (k / (n / 2)) * half + 3 * (k % (n / 2));
        float* dst = ptrs.dst[k] + 3 * (size_t)idx;
        dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
    }
}

cudaError_t launch_pack_attributes(int P, int n, const AttrPtrs& ptrs, float* packed, cudaStream_t s)
{
    if (P <= 0) return cudaSuccess;
    count_launch();
    pack_attributes_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, n, attr_stride(n), ptrs, packed);
    return cudaGetLastError();
}

cudaError_t launch_unpack_attributes(int P, int n, const float* packed, const AttrPtrs& ptrs, cudaStream_t s)
{
    if (P <= 0) return cudaSuccess;
    count_launch();
    unpack_attributes_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, n, attr_stride(n), packed, ptrs);
    return cudaGetLastError();
}

cudaError_t launch_recolor(int P, const float* colors, char* geom, const ghr_layout& L, cudaStream_t s)
{
    if (P <= 0) return cudaSuccess;
    count_launch();
    recolor_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, colors, reinterpret_cast<float4*>(geom + L.g_rec));
    return cudaGetLastError();
}

cudaError_t launch_emit_keys(const ghr_raster_params& p, char* geom, char* binning,
                             const ghr_layout& L, int64_t capacity, cudaStream_t s)
{
    if (p.P == 0) return cudaSuccess;
    StageTimer timer(GHR_STAGE_EMIT_KEYS, s);
    count_launch();
    emit_keys_kernel<<<(p.P + GHR_BLOCK - 1) / GHR_BLOCK, GHR_BLOCK, 0, s>>>(
        p.P, p.W, p.H, reinterpret_cast<const uint32_t*>(geom + L.g_tiles_touched),
        reinterpret_cast<const float4*>(geom + L.g_rec),
        reinterpret_cast<const uint32_t*>(geom + L.g_point_offsets),
        reinterpret_cast<const ghr_geom_header*>(geom + L.g_header), capacity,
        reinterpret_cast<uint64_t*>(binning + L.b_keys_a), reinterpret_cast<uint32_t*>(binning + L.b_vals_a),
        wide_masks_packed(p.P, capacity) ? 1 : 0,
        sort_histogram_in_emit(capacity) ? reinterpret_cast<SortHeader*>(binning + L.b_header) : nullptr, L.sort_bits,
        L.sort_passes);
    return cudaGetLastError();
}

cudaError_t launch_tile_ranges(const ghr_raster_params& p, const char* geom, char* binning, char* img,
                               const ghr_layout& L, int64_t capacity, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_TILE_RANGES, s);
    count_launch();
    const int tiles = ((p.W + GHR_TILE - 1) / GHR_TILE) * ((p.H + GHR_TILE - 1) / GHR_TILE);
    cudaError_t e = cudaMemsetAsync(img + L.i_ranges, 0, sizeof(uint2) * (size_t)tiles, s);
    if (e != cudaSuccess) return e;
    int64_t grid = (capacity + GHR_BLOCK - 1) / GHR_BLOCK;
    if (grid > sm_count() * 16) grid = sm_count() * 16;
    if (grid < 1) grid = 1;
    if (blend_throughput_regime(capacity)) {
        int64_t g4 = (capacity / 4 + GHR_BLOCK - 1) / GHR_BLOCK;
        if (g4 > sm_count() * 16) g4 = sm_count() * 16;
        if (g4 < 1) g4 = 1;
        tile_ranges_only_kernel<<<(unsigned)g4, GHR_BLOCK, 0, s>>>(
            reinterpret_cast<ghr_geom_header*>(const_cast<char*>(geom) + L.g_header), capacity,
            reinterpret_cast<const SortHeader*>(binning + L.b_header),
            reinterpret_cast<const uint64_t*>(binning + L.b_keys_a), reinterpret_cast<const uint64_t*>(binning + L.b_keys_b),
            reinterpret_cast<uint2*>(img + L.i_ranges));
    } else
    tile_ranges_kernel<<<(unsigned)grid, GHR_BLOCK, 0, s>>>(
        reinterpret_cast<ghr_geom_header*>(const_cast<char*>(geom) + L.g_header), capacity,
        (p.W + GHR_TILE - 1) / GHR_TILE, reinterpret_cast<const SortHeader*>(binning + L.b_header),
        reinterpret_cast<uint64_t*>(binning + L.b_keys_a), reinterpret_cast<uint64_t*>(binning + L.b_keys_b),
        reinterpret_cast<const uint32_t*>(binning + L.b_vals_a), reinterpret_cast<const uint32_t*>(binning + L.b_vals_b),
        reinterpret_cast<const float4*>(geom + L.g_rec), reinterpret_cast<uint2*>(img + L.i_ranges),
        blend_throughput_regime(capacity) ? 0 : 1);
    count_launch();
    tile_order_kernel<<<1, 1024, 0, s>>>(tiles, reinterpret_cast<const uint2*>(img + L.i_ranges),
                                         reinterpret_cast<const ghr_geom_header*>(geom + L.g_header), capacity,
                                         reinterpret_cast<uint32_t*>(img + L.i_order));
    return cudaGetLastError();
}

}  // namespace ghr
