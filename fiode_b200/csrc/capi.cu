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
int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return 148;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace fiode

extern "C" const char* fiode_last_error(void) { return fiode::g_err; }
extern "C" int fiode_version(void) { return 100; }
extern "C" long long fiode_launch_count(void) { return fiode::g_launches.load(); }
extern "C" void fiode_reset_launch_count(void) { fiode::g_launches.store(0); }
