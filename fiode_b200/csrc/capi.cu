// fiode_b200/csrc/capi.cu -- C-ABI plumbing shared by every entry point (error string, launch counter).
#include "common.cuh"
#include <cstdarg>
#include <atomic>
#include <cstdlib>
#include <mutex>

namespace fiode {
static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap; va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return -100 - (int)e;
}
// Per-device caches: one process may drive several GPUs (cudaFuncSetAttribute and the SM count are per device).
int current_device() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return 0;
    return dev;
}
int sm_count() {
    static std::atomic<int> n[kMaxDevices];
    const int dev = current_device();
    int v = n[dev].load(std::memory_order_relaxed);
    if (v == 0) {
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        n[dev].store(v, std::memory_order_relaxed);
    }
    return v;
}
// true exactly once per (flag array, device): the caller then sets its per-device function attributes
bool first_use_on_device(std::atomic<bool>* flags) {
    return !flags[current_device()].exchange(true, std::memory_order_acq_rel);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---- fork / join of an internal side stream (common.cuh) ------------------------------------------------------------
constexpr int kSideSlots = 8;          // callers on different streams mostly get different branches (hashed by stream handle)
struct SideSlot { cudaStream_t s = nullptr; cudaEvent_t fork = nullptr, join = nullptr; };
static SideSlot g_side[kMaxDevices][kSideSlots];
static std::mutex g_side_mu;
static std::atomic<int> g_fork{-1};     // -1: take FIODE_FORK on first use
static bool fork_enabled() {
    int v = g_fork.load(std::memory_order_relaxed);
    if (v < 0) {
        const char* e = getenv("FIODE_FORK");
        v = !(e && e[0] == '0');
        g_fork.store(v, std::memory_order_relaxed);
    }
    return v != 0;
}
static SideSlot* side_slot(cudaStream_t st) {
    const int dev = current_device();
    const size_t h = reinterpret_cast<size_t>(st);
    SideSlot* sl = &g_side[dev][((h >> 4) ^ (h >> 9) ^ (h >> 14)) % kSideSlots];
    if (sl->s) return sl;
    std::lock_guard<std::mutex> lk(g_side_mu);
    if (sl->s) return sl;
    cudaStream_t s = nullptr;
    int lo = 0, hi = 0;                      // branch kernels are on their layer's way to the next layer: highest priority
    if (cudaDeviceGetStreamPriorityRange(&lo, &hi) != cudaSuccess) { lo = hi = 0; (void)cudaGetLastError(); }
    static const bool prio = [] { const char* e = getenv("FIODE_GRAPH_PRIO"); return !(e && e[0] == '0'); }();
    if (cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking, prio ? hi : 0) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&sl->fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&sl->join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    sl->s = s;
    return sl;
}
cudaStream_t fork_side(cudaStream_t st) {
    if (!fork_enabled()) return st;
    SideSlot* sl = side_slot(st);
    if (!sl) { (void)cudaGetLastError(); return st; }          // no branch available: run in line
    // record + wait as one step: a wait refers to the event's latest record, and another host thread whose stream shares this
    // branch must not re-record the event in between
    std::lock_guard<std::mutex> lk(g_side_mu);
    if (cudaEventRecord(sl->fork, st) != cudaSuccess || cudaStreamWaitEvent(sl->s, sl->fork, 0) != cudaSuccess) {
        (void)cudaGetLastError();
        return st;
    }
    return sl->s;
}
int join_side(cudaStream_t st) {
    if (!fork_enabled()) return 0;
    SideSlot* sl = side_slot(st);
    if (!sl) return 0;
    std::lock_guard<std::mutex> lk(g_side_mu);
    cudaError_t e = cudaEventRecord(sl->join, sl->s);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(st, sl->join, 0);
    if (e != cudaSuccess) return cuda_fail(e, "join_side");
    return 0;
}
}  // namespace fiode

extern "C" const char* fiode_last_error(void) { return fiode::g_err; }
extern "C" int fiode_version(void) { return 100; }
extern "C" long long fiode_launch_count(void) { return fiode::g_launches.load(); }
extern "C" void fiode_reset_launch_count(void) { fiode::g_launches.store(0); }
extern "C" void fiode_set_side_streams(int on) { fiode::g_fork.store(on ? 1 : 0); }
