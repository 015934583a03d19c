// Launch accounting and optional per-kernel CUDA-event timing (measurement aid for bench.py; off by default).
#include "../../include/riggs_b200.h"

#include <atomic>
#include <cstring>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <utility>
#include <vector>

#include "common.cuh"

namespace rg {

static std::atomic<long long> g_launches{0};
static std::atomic<int> g_prof_on{0};
static std::mutex g_prof_mu;
struct EvPair { cudaEvent_t a, b; };
static std::map<std::string, std::vector<EvPair>> g_events;   // recorded, not yet read
static std::vector<cudaEvent_t> g_pool;

static cudaEvent_t get_event() {
    if (!g_pool.empty()) { cudaEvent_t e = g_pool.back(); g_pool.pop_back(); return e; }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}

LaunchScope::LaunchScope(const char* name, cudaStream_t stream, int n_kernels) : name_(name), stream_(stream), on_(false) {
    g_launches.fetch_add(n_kernels, std::memory_order_relaxed);
    if (g_prof_on.load(std::memory_order_relaxed)) {
        std::lock_guard<std::mutex> l(g_prof_mu);
        on_ = true;
        start_ = get_event();
        cudaEventRecord(start_, stream_);
    }
}

LaunchScope::~LaunchScope() {
    if (on_) {
        std::lock_guard<std::mutex> l(g_prof_mu);
        cudaEvent_t stop = get_event();
        cudaEventRecord(stop, stream_);
        g_events[name_].push_back({start_, stop});
    }
}

SideStream* side_stream() {
    thread_local std::map<int, SideStream> streams;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    auto it = streams.find(dev);
    if (it != streams.end()) return &it->second;
    SideStream s = {};
    if (cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s.join_a, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s.join_b, cudaEventDisableTiming) != cudaSuccess)
        return nullptr;
    return &streams.emplace(dev, s).first->second;
}

int ensure_dyn_smem(const void* func, int bytes) {
    static std::mutex mu;
    static std::set<std::pair<const void*, int>> done;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_error("ensure_dyn_smem: no current device"); return -1; }
    std::lock_guard<std::mutex> l(mu);
    if (done.count({func, dev})) return 0;
    cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) {
        set_error(std::string("cudaFuncSetAttribute(MaxDynamicSharedMemorySize): ") + cudaGetErrorString(e));
        return -1;
    }
    done.insert({func, dev});
    return 0;
}

}  // namespace rg

namespace rg {
void set_fwd_tile_profile(uint32_t* p);
void set_bwd_tile_profile(uint32_t* p);
}

using namespace rg;

extern "C" {

int rg_debug_tile_profile(uint32_t* fwd, uint32_t* bwd) {
    set_fwd_tile_profile(fwd);
    set_bwd_tile_profile(bwd);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

long long rg_launch_count(void) { return g_launches.load(); }

void rg_profile_enable(int on) { g_prof_on.store(on ? 1 : 0); }

void rg_profile_reset(void) {
    std::lock_guard<std::mutex> l(g_prof_mu);
    for (auto& kv : g_events)
        for (auto& p : kv.second) { g_pool.push_back(p.a); g_pool.push_back(p.b); }
    g_events.clear();
}

int rg_profile_count(void) {
    std::lock_guard<std::mutex> l(g_prof_mu);
    return (int)g_events.size();
}

int rg_profile_get(int index, char* name_out, int name_cap, float* total_ms, int* launches) {
    std::lock_guard<std::mutex> l(g_prof_mu);
    if (index < 0 || index >= (int)g_events.size()) return -1;
    auto it = g_events.begin();
    std::advance(it, index);
    float total = 0.f;
    for (auto& p : it->second) {
        cudaEventSynchronize(p.b);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) total += ms;
    }
    if (name_out && name_cap > 0) {
        strncpy(name_out, it->first.c_str(), name_cap - 1);
        name_out[name_cap - 1] = 0;
    }
    if (total_ms) *total_ms = total;
    if (launches) *launches = (int)it->second.size();
    return 0;
}

}  // extern "C"
