// Pieces of the onesweep sort (ghr_sort.cu) that the key emission (ghr_binning.cu) shares: it accumulates the digit
// histograms of the keys it writes and its last CTA writes the sort plan, so that a frame needs no histogram kernel.
#pragma once
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

// Squeezed keys.  A depth|tile key is (tile << 32) | bits(depth) and the depths of one frame span a narrow range (a
// body 2.4 .. 3.6 m from the camera: sign, exponent and nothing else are constant), so the key's low word minus the
// frame's smallest depth needs only L = bit-length(max - min) bits, and the digits are taken from
//     sq(key) = (low32(key) - lo)  |  (high32(key) << L)        (L + tile bits instead of 32 + tile bits)
// Order is preserved (lo <= every depth), the passes still move the ORIGINAL keys, and the passes left without any
// varying bit become "all keys share this digit" -- the trivial passes the plan already skips.  A 512^2 .. 1080p body
// frame goes from 5 radix passes to 4.  `sq` = {depth_hi, ~depth_lo} written by the preprocess kernel, or NULL
// (generic sort: identity).
struct Squeeze {
    uint32_t lo, L;
};
__device__ __forceinline__ Squeeze load_squeeze(const uint32_t* sq, int bits)
{
    Squeeze q;
    q.lo = 0u; q.L = 32u;
    if (sq) {
        const uint32_t hi = sq[0], lo = ~sq[1];
        if (hi >= lo) {  // hi < lo: no visible Gaussian, no key
            const uint32_t L = 32u - (uint32_t)__clz(hi - lo);
            // only when it saves a pass: otherwise the plain digits are the better balanced ones (3 M Gaussians over
            // 1 .. 11 m: 38 squeezed bits still take 5 passes, the last over 2 bits -- measured 9 % slower)
            const int squeezed_bits = bits - 32 + (int)L;
            if ((squeezed_bits + kRadixBits - 1) / kRadixBits < (bits + kRadixBits - 1) / kRadixBits) { q.lo = lo; q.L = L; }
        }
    }
    return q;
}
__device__ __forceinline__ uint64_t squeezed(uint64_t key, Squeeze q)
{
    return (uint64_t)((uint32_t)key - q.lo) | ((key >> 32) << q.L);
}

// Column of digit d in a warp's row of shared-memory counters.  A digit that straddles the depth exponent and the tile id
// (bits 27..35 of a 1080p key) takes values t * 32 + c with two or three c only: unswizzled, every such digit of a warp
// lands in the same two or three banks and the leaders' read-modify-writes serialise (ncu: 2.5x the shared-memory
// wavefronts of the other passes).  XOR-ing the upper digit bits into the bank bits keeps the map a bijection and keeps
// "thread t owns digit t" conflict-free (a warp's 32 threads share t >> 5).
__device__ __forceinline__ uint32_t hist_col(uint32_t d) { return d ^ (d >> 5); }

}  // namespace ghr
