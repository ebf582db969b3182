// fiode_b200/csrc/grid.cu -- direct enumeration of the decision-boundary grid.
//
// Replaces robustness/eval_utils.py:31-61 sample_decision_boundary (an O(N^2) numpy
// concatenate-in-a-loop dynamic programme; minutes and >10 GB at T=40) and :72-89
// count_samples_decision_boundary.  The set is exactly
//     { z in Z^n_{>=0} : sum z = T, z_0 = max_{i>=1} z_i },   eta = z / T,
// enumerated here by UNRANKING: a small table of bounded-composition counts is built on the host
// and each thread turns its row index into the row's numerators.  Output is the canonical label-0
// grid as uint8 numerators (413 MB at n=10, T=40 instead of 1.65 GB of fp32 per label); the
// per-label column swap of eval_utils.py:64-69 is applied by the consumer on load.
#include "common.cuh"
#include <vector>

namespace fiode {

// cnt[((m * n + slots) * (T + 1) + rem) * 2 + hit] = #{x in [0,m]^slots : sum x = rem, (hit or some x_i == m)}
static void build_tables(int n, int T, std::vector<unsigned long long>& cnt, std::vector<unsigned long long>& prefix) {
    cnt.assign((size_t)(T + 1) * n * (T + 1) * 2, 0ull);
    prefix.assign(T + 2, 0ull);
    auto at = [&](int m, int slots, int rem, int hit) -> unsigned long long& {
        return cnt[(((size_t)m * n + slots) * (T + 1) + rem) * 2 + hit];
    };
    for (int m = 0; m <= T; ++m) {
        at(m, 0, 0, 1) = 1;
        for (int slots = 1; slots < n; ++slots)
            for (int rem = 0; rem <= T; ++rem)
                for (int hit = 0; hit < 2; ++hit) {
                    unsigned long long c = 0;
                    for (int z = 0; z <= m && z <= rem; ++z) c += at(m, slots - 1, rem - z, hit | (z == m));
                    at(m, slots, rem, hit) = c;
                }
        prefix[m + 1] = prefix[m] + (T - m >= 0 ? at(m, n - 1, T - m, 0) : 0ull);
    }
}

__global__ void __launch_bounds__(256) grid_unrank_kernel(const unsigned long long* __restrict__ cnt,
                                                          const unsigned long long* __restrict__ prefix, int n,
                                                          int T, long long first, long long count,
                                                          uint8_t* __restrict__ out) {
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < count; i += (long long)gridDim.x * 256) {
        unsigned long long r = (unsigned long long)(first + i);
        int m = 0;
        while (prefix[m + 1] <= r) ++m;
        r -= prefix[m];
        uint8_t* row = out + i * n;
        row[0] = (uint8_t)m;
        int rem = T - m, hit = 0;
        for (int s = 1; s < n; ++s) {
            const int slots = n - 1 - s;
            int z = 0;
            for (; z < m && z < rem; ++z) {
                const unsigned long long c = cnt[(((size_t)m * n + slots) * (T + 1) + (rem - z)) * 2 + (hit | (z == m))];
                if (r < c) break;
                r -= c;
            }
            row[s] = (uint8_t)z;
            rem -= z;
            hit |= (z == m);
        }
    }
}

}  // namespace fiode

using namespace fiode;

extern "C" long long fiode_grid_count(int n, int T) {
    if (n < 2 || n > 16 || T < 1 || T > 255) { set_error("fiode_grid_count: need 2 <= n <= 16, 1 <= T <= 255"); return -1; }
    std::vector<unsigned long long> cnt, prefix;
    build_tables(n, T, cnt, prefix);
    return (long long)prefix[T + 1];
}

extern "C" size_t fiode_grid_workspace_bytes(int n, int T) {
    return ((size_t)(T + 1) * n * (T + 1) * 2 + (T + 2)) * sizeof(unsigned long long) + 256;
}

extern "C" int fiode_grid_enumerate(int n, int T, long long first, long long count, uint8_t* out, void* ws,
                                    size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n < 2 || n > 16 || T < 1 || T > 255) { set_error("fiode_grid_enumerate: need 2 <= n <= 16, 1 <= T <= 255"); return -1; }
    if (!out || !ws) { set_error("fiode_grid_enumerate: null pointer"); return -1; }
    if (ws_bytes < fiode_grid_workspace_bytes(n, T)) { set_error("fiode_grid_enumerate: workspace too small"); return -1; }
    std::vector<unsigned long long> cnt, prefix;
    build_tables(n, T, cnt, prefix);
    const long long total = (long long)prefix[T + 1];
    if (first < 0 || count < 0 || first + count > total) { set_error("fiode_grid_enumerate: rows [%lld, %lld) outside [0, %lld)", first, first + count, total); return -1; }
    if (count == 0) return 0;
    unsigned long long* d_cnt = reinterpret_cast<unsigned long long*>(ws);
    unsigned long long* d_prefix = d_cnt + cnt.size();
    // pageable host vectors: synchronous staging copies are fine here (one-off table upload)
    FIODE_CUDA(cudaMemcpyAsync(d_cnt, cnt.data(), cnt.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice, st));
    FIODE_CUDA(cudaMemcpyAsync(d_prefix, prefix.data(), prefix.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice, st));
    FIODE_CUDA(cudaStreamSynchronize(st));
    const long long want = (count + 255) / 256;
    const int grid = (int)(want < 16ll * sm_count() ? want : 16ll * sm_count());
    grid_unrank_kernel<<<grid, 256, 0, st>>>(d_cnt, d_prefix, n, T, first, count, out);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
