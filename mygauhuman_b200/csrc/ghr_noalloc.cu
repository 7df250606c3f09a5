// ghr_sort_pairs_u64_static: the (key, value) radix sort of ghr_sort_pairs_u64 for callers that cannot allocate (CUDA
// graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The only difference from ghr_sort_pairs_u64 is where the scratch lives: instead of a caller
// workspace, the SortHeader (digit histograms, tickets) and the onesweep look-back words sit in ONE file-scope device
// array, g_static_sort_scratch, that every GPU loading libghr.so holds until exit.
//
//  * limit: N <= GHR_STATIC_SORT_MAX_KEYS (2^20 pairs; the array is ~4.2 MB);  larger N -> GHR_ERR_CAPACITY + message
//  * the entry point calls no cudaMalloc / cudaFree / allocating library routine: two async memsets/copies on `stream`
//    and the same kernels as ghr_sort_pairs_u64 (launch_sort, ghr_sort.cu)
//  * every call shares the array and finds the previous call's contents in it: the header and the look-back words the
//    call will use are cleared on `stream` first, exactly as ghr_sort_pairs_u64 clears its workspace
//  * calls that overlap on the GPU (two streams, two host threads) RACE on the array: serialise them
#include <stdio.h>
#include "ghr_common.cuh"
#include "ghr_internal.h"

using namespace ghr;

namespace {

constexpr int64_t kStaticMaxKeys = int64_t(1) << 20;  // == GHR_STATIC_SORT_MAX_KEYS (include/ghr.h)
static_assert(kStaticMaxKeys == GHR_STATIC_SORT_MAX_KEYS, "header and kernel file disagree on the limit");

// The largest workspace ghr_sort_workspace_bytes() asks for at any N <= kStaticMaxKeys.  sort_tiles() is not monotone in
// N (keys per tile grow 1024 -> 2048 -> 4096 -> 6144 at 2^17, 3*2^17 and 5*2^17 keys), so take a bound on the maximum
// over the regimes (128, 192, 160 and 171 tiles): 256 tiles.
constexpr size_t kHeaderBytes = (sizeof(SortHeader) + 255) / 256 * 256;
constexpr size_t kStatusBytes = sizeof(uint32_t) * (size_t)kMaxPasses * 256 * (size_t)kRadix;
constexpr size_t kScratchBytes = kHeaderBytes + kStatusBytes + 256;

// In host code this name is a host-side stand-in, NOT a device address: never hand it to a kernel or a memset.  The
// address comes from cudaGetSymbolAddress (scratch_for_current_device).
static __device__ __align__(256) unsigned char g_static_sort_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
char* g_scratch_addr[kMaxDevices];  // per device; written once (idempotent: the same value from every thread)

// Device address of the array on the CURRENT device, resolved once per device with that device current.
int scratch_for_current_device(char** out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        fprintf(stderr, "[libghr] ghr_sort_pairs_u64_static: cudaGetDevice: %s (device %d)\n", cudaGetErrorString(e), dev);
        return GHR_ERR_CUDA;
    }
    char* p = __atomic_load_n(&g_scratch_addr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        void* sym = nullptr;
        e = cudaGetSymbolAddress(&sym, g_static_sort_scratch);
        if (e != cudaSuccess || !sym) {
            fprintf(stderr, "[libghr] ghr_sort_pairs_u64_static: cudaGetSymbolAddress on device %d: %s\n", dev,
                    cudaGetErrorString(e));
            return GHR_ERR_CUDA;
        }
        p = static_cast<char*>(sym);
        __atomic_store_n(&g_scratch_addr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return GHR_OK;
}

int cuda_ok(cudaError_t e, const char* what)
{
    if (e == cudaSuccess) return GHR_OK;
    fprintf(stderr, "[libghr] CUDA error in %s: %s\n", what, cudaGetErrorString(e));
    return GHR_ERR_CUDA;
}

}  // namespace

extern "C" {

int64_t ghr_sort_static_max_keys(void) { return kStaticMaxKeys; }

int ghr_sort_pairs_u64_static(int64_t N, int32_t bits, uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a,
                              uint32_t* vals_b, uint32_t* sorted_sel_dev, void* stream)
{
    if (N < 0 || bits <= 0 || bits > 64) return GHR_ERR_BAD_ARG;
    if (N > kStaticMaxKeys) {
        fprintf(stderr, "[libghr] ghr_sort_pairs_u64_static: N = %lld exceeds the static scratch limit of %lld pairs; "
                        "use ghr_sort_pairs_u64 with a caller workspace\n", (long long)N, (long long)kStaticMaxKeys);
        return GHR_ERR_CAPACITY;
    }
    if (N > 0 && (!keys_a || !keys_b || !vals_a || !vals_b)) return GHR_ERR_BAD_ARG;
    // the tile size is tunable through GHR_SORT_ITEMS (tools only): recheck against what this call really needs
    const size_t used = kHeaderBytes + sort_status_bytes(sort_passes_for(bits), N);
    if (used + 256 > kScratchBytes) {
        fprintf(stderr, "[libghr] ghr_sort_pairs_u64_static: %zu scratch bytes needed, %zu held\n", used + 256, kScratchBytes);
        return GHR_ERR_BUFFER;
    }
    char* w = nullptr;
    int r = scratch_for_current_device(&w);
    if (r != GHR_OK) return r;
    cudaStream_t s = (cudaStream_t)stream;
    SortHeader* hdr = reinterpret_cast<SortHeader*>(w);
    uint32_t* status = reinterpret_cast<uint32_t*>(w + kHeaderBytes);
    // the array keeps the last call's histograms, tickets and look-back words: clear what this call will read
    r = cuda_ok(cudaMemsetAsync(w, 0, used, s), "static sort memset");
    if (r != GHR_OK) return r;
    r = cuda_ok(launch_sort(hdr, status, nullptr, N, bits, keys_a, keys_b, vals_a, vals_b, s), "static sort");
    if (r != GHR_OK) return r;
    if (sorted_sel_dev)
        r = cuda_ok(cudaMemcpyAsync(sorted_sel_dev, &hdr->sorted_sel, sizeof(uint32_t), cudaMemcpyDeviceToDevice, s),
                    "static sort sel copy");
    return r;
}

}  // extern "C"
