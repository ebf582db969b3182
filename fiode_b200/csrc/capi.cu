// fiode_b200/csrc/capi.cu -- C-ABI plumbing shared by every entry point (error string, launch counter).
#include "common.cuh"
#include <cstdarg>
#include <atomic>

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
}  // namespace fiode

extern "C" const char* fiode_last_error(void) { return fiode::g_err; }
extern "C" int fiode_version(void) { return 100; }
extern "C" long long fiode_launch_count(void) { return fiode::g_launches.load(); }
extern "C" void fiode_reset_launch_count(void) { fiode::g_launches.store(0); }
