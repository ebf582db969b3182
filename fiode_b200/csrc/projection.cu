// fiode_b200/csrc/projection.cu -- the barrier QP projection as a stand-alone op (autograd seam).
//
// Replaces barrier_projection/barrier_projection.py:
//   * FastBarrierProjectionNoUpper.forward :232-269  / FastBarrierProjection.forward :123-160
//       argmin 1/2 ||v - nominal||^2  s.t.  sum v = 0, v >= lower (, v <= upper)
//     bisection on mu, v = clamp(nominal - mu); the loop stops for the WHOLE batch at the first
//     midpoint where max_rows |sum v| < tol (:247).  Pass 1 records, per row, at which midpoints it
//     is not converged and ORs the words over the batch; pass 2 replays to that iteration.
//   * FastBarrierProjectionNoUpper.backward :272-311 in closed form.
#include "common.cuh"

namespace fiode {

template <bool UPPER>
__device__ __forceinline__ void proj_bounds(const float (&f)[NH], const float (&lo)[NH], const float (&up)[NH],
                                            float& hi, float& lw) {
    hi = f[0] - lo[0];
    lw = UPPER ? f[0] - up[0] : f[0];
#pragma unroll
    for (int j = 1; j < NH; ++j) {
        hi = fmaxf(hi, f[j] - lo[j]);
        lw = fminf(lw, UPPER ? f[j] - up[j] : f[j]);
    }
}
template <bool UPPER>
__device__ __forceinline__ float proj_sum(const float (&f)[NH], const float (&lo)[NH], const float (&up)[NH], float mu) {
    float e = 0.f;
#pragma unroll
    for (int j = 0; j < NH; ++j) {
        float v = fmaxf(f[j] - mu, lo[j]);
        if (UPPER) v = fminf(v, up[j]);
        e += v;
    }
    return e;
}

template <bool UPPER>
__global__ void __launch_bounds__(256) project_scan_kernel(const float* __restrict__ lower,
                                                           const float* __restrict__ upper,
                                                           const float* __restrict__ nominal, long long N,
                                                           int max_iter, float tol, unsigned* __restrict__ notconv) {
    unsigned nc = 0;
    for (long long r = blockIdx.x * 256ll + threadIdx.x; r < N; r += (long long)gridDim.x * 256) {
        float f[NH], lo[NH], up[NH];
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            f[j] = nominal[r * NH + j]; lo[j] = lower[r * NH + j]; up[j] = UPPER ? upper[r * NH + j] : 0.f;
        }
        float hi, lw;
        proj_bounds<UPPER>(f, lo, up, hi, lw);
        for (int k = 0; k < max_iter; ++k) {
            const float mu = (hi - lw) / 2.f + lw;
            const float e = proj_sum<UPPER>(f, lo, up, mu);
            if (!(fabsf(e) < tol)) nc |= 1u << k;
            if (e > 0.f) lw = mu;
            if (e < 0.f) hi = mu;
        }
    }
    nc = __reduce_or_sync(0xffffffffu, nc);
    if ((threadIdx.x & 31) == 0 && nc) atomicOr(notconv, nc);
}

template <bool UPPER>
__global__ void __launch_bounds__(256) project_replay_kernel(const float* __restrict__ lower,
                                                             const float* __restrict__ upper,
                                                             const float* __restrict__ nominal, long long N,
                                                             int max_iter, const unsigned* __restrict__ notconv,
                                                             float* __restrict__ v, float* __restrict__ mu_out,
                                                             int* __restrict__ n_iter) {
    const int K = stop_iter(*notconv, max_iter);
    if (blockIdx.x == 0 && threadIdx.x == 0 && n_iter) *n_iter = K + 1;
    for (long long r = blockIdx.x * 256ll + threadIdx.x; r < N; r += (long long)gridDim.x * 256) {
        float f[NH], lo[NH], up[NH];
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            f[j] = nominal[r * NH + j]; lo[j] = lower[r * NH + j]; up[j] = UPPER ? upper[r * NH + j] : 0.f;
        }
        float hi, lw, mu = 0.f;
        proj_bounds<UPPER>(f, lo, up, hi, lw);
        for (int k = 0; k <= K; ++k) {
            mu = (hi - lw) / 2.f + lw;
            const float e = proj_sum<UPPER>(f, lo, up, mu);
            if (e > 0.f) lw = mu;
            if (e < 0.f) hi = mu;
        }
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            float x = fmaxf(f[j] - mu, lo[j]);
            if (UPPER) x = fminf(x, up[j]);
            v[r * NH + j] = x;
        }
        if (mu_out) mu_out[r] = mu;
    }
}

// backward, barrier_projection.py:288-311: L = {v - nominal + mu > 0} (lower bound active), F = complement,
// d_nominal_j = g_j - mean_F g (j in F), d_lower_j = g_j - mean_F g (j in L); |F| = 0 -> d_lower = g.
__global__ void __launch_bounds__(256) project_bwd_kernel(const float* __restrict__ g, const float* __restrict__ v,
                                                          const float* __restrict__ mu,
                                                          const float* __restrict__ nominal, long long N,
                                                          float* __restrict__ d_lower, float* __restrict__ d_nominal) {
    for (long long r = blockIdx.x * 256ll + threadIdx.x; r < N; r += (long long)gridDim.x * 256) {
        const float m = mu[r];
        float gj[NH]; bool act[NH];
        float sum_free = 0.f; int n_free = 0;
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            gj[j] = g[r * NH + j];
            act[j] = (v[r * NH + j] - nominal[r * NH + j] + m) > 0.f;
            if (!act[j]) { sum_free += gj[j]; ++n_free; }
        }
        const float mean = n_free > 0 ? sum_free / (float)n_free : 0.f;
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            const float d = gj[j] - mean;
            d_lower[r * NH + j] = act[j] ? d : 0.f;
            d_nominal[r * NH + j] = act[j] ? 0.f : d;
        }
    }
}

// backward of the upper-bounded projection, barrier_projection.py:164-214, as written there: lambda = v - nominal + mu,
// L = {lambda > 0}, U = {lambda < 0}, F = neither;  d_nominal_j = g_j - mean_F g (j in F),  d_lower_j = g_j - mean_{not L} g
// (j in L),  d_upper_j = g_j - mean_{not U} g (j in U)  -- the two bound Jacobians average over the COMPLEMENT of their own
// active set (which contains the other bound's active entries), not over F; an empty complement leaves d = g.
__global__ void __launch_bounds__(256) project_bwd_upper_kernel(const float* __restrict__ g, const float* __restrict__ v,
                                                                const float* __restrict__ mu, const float* __restrict__ nominal,
                                                                long long N, float* __restrict__ d_lower, float* __restrict__ d_upper,
                                                                float* __restrict__ d_nominal) {
    for (long long r = blockIdx.x * 256ll + threadIdx.x; r < N; r += (long long)gridDim.x * 256) {
        const float m = mu[r];
        float gj[NH]; bool lo[NH], up[NH];
        float s_free = 0.f, s_notl = 0.f, s_notu = 0.f; int n_free = 0, n_notl = 0, n_notu = 0;
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            gj[j] = g[r * NH + j];
            const float lam = v[r * NH + j] - nominal[r * NH + j] + m;
            lo[j] = lam > 0.f; up[j] = lam < 0.f;
            if (!lo[j]) { s_notl += gj[j]; ++n_notl; }
            if (!up[j]) { s_notu += gj[j]; ++n_notu; }
            if (!lo[j] && !up[j]) { s_free += gj[j]; ++n_free; }
        }
        const float mf = n_free > 0 ? s_free / (float)n_free : 0.f;
        const float ml = n_notl > 0 ? s_notl / (float)n_notl : 0.f;
        const float mu_ = n_notu > 0 ? s_notu / (float)n_notu : 0.f;
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            d_lower[r * NH + j] = lo[j] ? gj[j] - ml : 0.f;
            d_upper[r * NH + j] = up[j] ? gj[j] - mu_ : 0.f;
            d_nominal[r * NH + j] = (!lo[j] && !up[j]) ? gj[j] - mf : 0.f;
        }
    }
}

}  // namespace fiode

using namespace fiode;

extern "C" int fiode_barrier_project_bwd_upper(const float* grad_v, const float* v, const float* mu, const float* nominal,
                                               long long N, float* d_lower, float* d_upper, float* d_nominal, void* stream) {
    if (!grad_v || !v || !mu || !nominal || !d_lower || !d_upper || !d_nominal) { set_error("fiode_barrier_project_bwd_upper: null pointer"); return -1; }
    if (N <= 0) { set_error("fiode_barrier_project_bwd_upper: N=%lld", N); return -1; }
    const long long want = (N + 255) / 256;
    const int grid = (int)(want < 8ll * sm_count() ? want : 8ll * sm_count());
    project_bwd_upper_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(grad_v, v, mu, nominal, N, d_lower, d_upper, d_nominal);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}

extern "C" int fiode_barrier_project_fwd(const float* lower, const float* upper, const float* nominal, long long N,
                                         int max_iter, float tol, float* v, float* mu, int* n_iter, void* ws,
                                         size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!lower || !nominal || !v || !ws) { set_error("fiode_barrier_project_fwd: null pointer"); return -1; }
    if (N <= 0) { set_error("fiode_barrier_project_fwd: N=%lld", N); return -1; }
    if (max_iter < 1 || max_iter > 32) { set_error("fiode_barrier_project_fwd: max_iter must be in [1,32]"); return -1; }
    if (ws_bytes < sizeof(unsigned)) { set_error("fiode_barrier_project_fwd: workspace too small"); return -1; }
    unsigned* nc = reinterpret_cast<unsigned*>(ws);
    FIODE_CUDA(cudaMemsetAsync(nc, 0, sizeof(unsigned), st));
    const long long want = (N + 255) / 256;
    const int grid = (int)(want < 8ll * sm_count() ? want : 8ll * sm_count());
    if (upper) {
        project_scan_kernel<true><<<grid, 256, 0, st>>>(lower, upper, nominal, N, max_iter, tol, nc);
        project_replay_kernel<true><<<grid, 256, 0, st>>>(lower, upper, nominal, N, max_iter, nc, v, mu, n_iter);
    } else {
        project_scan_kernel<false><<<grid, 256, 0, st>>>(lower, nullptr, nominal, N, max_iter, tol, nc);
        project_replay_kernel<false><<<grid, 256, 0, st>>>(lower, nullptr, nominal, N, max_iter, nc, v, mu, n_iter);
    }
    FIODE_CUDA(cudaGetLastError());
    count_launch(2);
    return 0;
}

extern "C" int fiode_barrier_project_bwd(const float* grad_v, const float* v, const float* mu, const float* nominal,
                                         long long N, float* d_lower, float* d_nominal, void* stream) {
    if (!grad_v || !v || !mu || !nominal || !d_lower || !d_nominal) { set_error("fiode_barrier_project_bwd: null pointer"); return -1; }
    if (N <= 0) { set_error("fiode_barrier_project_bwd: N=%lld", N); return -1; }
    const long long want = (N + 255) / 256;
    const int grid = (int)(want < 8ll * sm_count() ? want : 8ll * sm_count());
    project_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(grad_v, v, mu, nominal, N, d_lower, d_nominal);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
