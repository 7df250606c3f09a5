// Measurement hooks of libghr (include/ghr.h "measurement" section): a
// process-wide kernel-launch counter and optional per-stage CUDA-event timers
// recorded on the caller's stream (so the figures are device time of the
// stage on the stream it is launched on, not host time).  Off by default;
// when off, a stage boundary costs one relaxed atomic load.
#include <atomic>
#include <mutex>
#include <vector>
#include "ghr_internal.h"

namespace ghr {

std::atomic<uint64_t> g_launches{0};
std::atomic<int> g_profile_on{0};
std::atomic<int> g_shared_gpu{0};

namespace {
struct Span {
    cudaEvent_t beg, end;
    int stage;
};
std::mutex g_mu;
std::vector<Span> g_spans;        // recorded, not yet read
std::vector<cudaEvent_t> g_pool;  // recycled events
constexpr size_t kMaxSpans = 1 << 16;

cudaEvent_t get_event()
{
    if (!g_pool.empty()) {
        cudaEvent_t e = g_pool.back();
        g_pool.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}
}  // namespace

int sm_count()
{
    static std::atomic<int> cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int n = cached[dev].load(std::memory_order_relaxed);
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

StageTimer::StageTimer(int stage, cudaStream_t s) : stage_(stage), stream_(s), beg_(nullptr)
{
    if (!g_profile_on.load(std::memory_order_relaxed)) return;
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_spans.size() >= kMaxSpans) return;
    beg_ = get_event();
    if (beg_) cudaEventRecord(beg_, stream_);
}

StageTimer::~StageTimer()
{
    if (!beg_) return;
    std::lock_guard<std::mutex> lk(g_mu);
    cudaEvent_t end = get_event();
    if (!end) { g_pool.push_back(beg_); return; }
    cudaEventRecord(end, stream_);
    g_spans.push_back({beg_, end, stage_});
}

}  // namespace ghr

using namespace ghr;

extern "C" {

int ghr_set_shared_gpu(int on)
{
    g_shared_gpu.store(on ? 1 : 0, std::memory_order_relaxed);
    return GHR_OK;
}

int ghr_shared_gpu(void) { return g_shared_gpu.load(std::memory_order_relaxed); }

uint64_t ghr_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }


int ghr_profile_enable(int on)
{
    g_profile_on.store(on ? 1 : 0, std::memory_order_relaxed);
    return GHR_OK;
}

int ghr_profile_read(double* ms_out, int64_t* count_out)
{
    if (!ms_out || !count_out) return GHR_ERR_BAD_ARG;
    for (int i = 0; i < GHR_STAGE_COUNT; i++) { ms_out[i] = 0.0; count_out[i] = 0; }
    std::lock_guard<std::mutex> lk(g_mu);
    int rc = GHR_OK;
    for (const Span& sp : g_spans) {
        float ms = 0.f;
        if (cudaEventSynchronize(sp.end) != cudaSuccess || cudaEventElapsedTime(&ms, sp.beg, sp.end) != cudaSuccess) {
            rc = GHR_ERR_CUDA;
        } else if (sp.stage >= 0 && sp.stage < GHR_STAGE_COUNT) {
            ms_out[sp.stage] += ms;
            count_out[sp.stage] += 1;
        }
        g_pool.push_back(sp.beg);
        g_pool.push_back(sp.end);
    }
    g_spans.clear();
    return rc;
}

}  // extern "C"
