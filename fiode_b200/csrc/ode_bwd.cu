// fiode_b200/csrc/ode_bwd.cu -- vector-Jacobian products of the ODE right-hand side, for the differentiable solve.
//
// The reference differentiates THROUGH the solver (use_adjoint=False, pl_modules.py:303): torchdiffeq's dopri5 is plain
// torch arithmetic, so loss.backward() (train_ode, pl_modules.py:490-500; PGD in validation_step :189-199) back-propagates
// through every accepted Runge-Kutta stage with the step sizes held constant (its step-size controller runs under
// torch.no_grad).  The host side (fiode_b200/odeint.py) replays the accepted steps the device solver logged and walks
// them backwards; the three entry points here are what one stage needs:
//   fiode_rhs_fwd_train   f(h, s) with the activations kept (z1, z2 post-ReLU, ftilde, mu)
//   fiode_rhs_bwd_data    g_f -> g_h, g_s and the per-layer deltas (dft, d2, d1) of ONE evaluation
//   fiode_rhs_bwd_params  all evaluations stacked: dQ1, dQx, dQ2, dQ3, db1 (= dbx), db2, db3 in one set of GEMMs
// following dynamics/classification.py:96-115 and the NoUpper projection adjoint barrier_projection.py:288-311.
#include "common.cuh"
#include "rows.cuh"

namespace fiode {

int gemm_nt(const float* A, int lda, const float* Bm, int ldb, int M, int N, int K, const float* bias, float alpha,
            int group_sort, float* y, int ldy, float* ws, size_t ws_bytes, cudaStream_t st);
size_t gemm_nt_workspace(int M, int N, int K);
int gemm_tn(const float* A, int lda, const float* Bm, int ldb, int M, int N, long long K, float alpha, float* y, int ldy,
            float* ws, size_t ws_bytes, cudaStream_t st);
size_t gemm_tn_workspace(int M, int N, long long K);
int launch_rhs_finish(const float* ft, const float* h, long long N, const fiode_dyn_weights* w, const unsigned* notconv,
                      float* f, float* mu, int* n_iter, cudaStream_t st);

static size_t al_(size_t x) { return (x + 255) & ~(size_t)255; }

// g_f -> projection adjoint -> (dft: gradient w.r.t. the raw MLP output, gh: the direct h terms through the barrier
// lower bound and, with scale_nominal, through upper / lower inside the sigmoid rescale)
__global__ void __launch_bounds__(256) rhs_seed_kernel(const float* __restrict__ h, long long N, fiode_dyn_weights w,
                                                       const float* __restrict__ g, const float* __restrict__ f,
                                                       const float* __restrict__ ft, const float* __restrict__ mu,
                                                       float* __restrict__ dft, float* __restrict__ gh) {
    for (long long r = blockIdx.x * 256ll + threadIdx.x; r < N; r += (long long)gridDim.x * 256) {
        const float m = mu[r];
        float gj[NH]; bool act[NH];
        float sum_free = 0.f; int n_free = 0;
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            gj[j] = g[r * NH + j];
            act[j] = (f[r * NH + j] - ft[r * NH + j] + m) > 0.f;
            if (!act[j]) { sum_free += gj[j]; ++n_free; }
        }
        const float mean = n_free > 0 ? sum_free / (float)n_free : 0.f;
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            const float d = gj[j] - mean;
            const float d_lower = act[j] ? d : 0.f, d_nom = act[j] ? 0.f : d;
            const float hj = h[r * NH + j];
            const float dlo = -w.alpha_1 * w.sigma_1 * exp_cr(w.sigma_1 * hj);      // d lower / d h
            float dx = d_nom, dh = d_lower * dlo;
            if (w.scale_nominal) {   // nominal = (up - lo) sig(x) + lo, up = alpha_2 (1 - h)
                const float lo = barrier_lower(w.alpha_1, w.sigma_1, hj), up = w.alpha_2 * (1.f - hj);
                const float nom = ft[r * NH + j];
                const float sg = (nom - lo) / (up - lo);
                dx = d_nom * (nom - lo) * (up - nom) / (up - lo);
                dh += d_nom * ((-w.alpha_2 - dlo) * sg + dlo);
            }
            dft[r * NH + j] = dx;
            gh[r * NH + j] = dh;
        }
    }
}
__global__ void __launch_bounds__(256) relu_bwd_kernel(float* __restrict__ dz, const float* __restrict__ z, size_t n) {
    for (size_t e = blockIdx.x * 256ull + threadIdx.x; e < n; e += (size_t)gridDim.x * 256)
        if (!(z[e] > 0.f)) dz[e] = 0.f;
}
__global__ void __launch_bounds__(256) add_inplace_kernel(float* __restrict__ y, const float* __restrict__ x, size_t n) {
    for (size_t e = blockIdx.x * 256ull + threadIdx.x; e < n; e += (size_t)gridDim.x * 256) y[e] += x[e];
}
__global__ void tr_kernel(const float* __restrict__ x, int rows, int cols, float* __restrict__ y) {
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < rows * cols; e += gridDim.x * blockDim.x) {
        const int r = e / cols, c = e - r * cols;
        y[c * rows + r] = x[e];
    }
}
__global__ void fill_ones_kernel(float* __restrict__ x, size_t n) {
    for (size_t e = blockIdx.x * 256ull + threadIdx.x; e < n; e += (size_t)gridDim.x * 256) x[e] = 1.f;
}

struct LinComb { const float* x[8]; float c[8]; int n; };
// out = sum_i c_i x_i  (out may alias any x_i)
__global__ void __launch_bounds__(256) lincomb_kernel(LinComb a, float* out, size_t n) {
    for (size_t e = blockIdx.x * 256ull + threadIdx.x; e < n; e += (size_t)gridDim.x * 256) {
        float acc = 0.f;
        for (int i = 0; i < a.n; ++i) acc = fmaf(a.c[i], a.x[i][e], acc);
        out[e] = acc;
    }
}

static unsigned gridn(size_t total) {
    const size_t b = (total + 255) / 256;
    return (unsigned)(b < 1 ? 1 : (b < 4096 ? b : 4096));
}

}  // namespace fiode

using namespace fiode;

extern "C" size_t fiode_rhs_train_workspace_bytes(long long N) {
    const size_t g1 = gemm_nt_workspace((int)N, NM, NM), g2 = gemm_nt_workspace((int)N, NH, NM);
    return 256 + al_((size_t)N * NH * 4) + al_(g1 > g2 ? g1 : g2) + 256;
}

extern "C" int fiode_rhs_fwd_train(const fiode_dyn_weights* w, const float* h, const float* s, long long N, float* f,
                                   float* ftilde, float* mu, float* z1, float* z2, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!w || !h || !s || !f || !ftilde || !mu || !z1 || !z2 || !ws) { set_error("fiode_rhs_fwd_train: null pointer"); return -1; }
    if (N <= 0 || ws_bytes < 256) { set_error("fiode_rhs_fwd_train: bad arguments"); return -1; }
    unsigned* notconv = reinterpret_cast<unsigned*>(ws);
    int rc = launch_rows_store(w, h, s, N, N, 1, ftilde, notconv, nullptr, nullptr, z1, z2, st);
    if (rc) return rc;
    return launch_rhs_finish(ftilde, h, N, w, notconv, f, mu, nullptr, st);
}

// Q1T (10,128), QxT (10,128), Q2T (128,128), Q3T (128,10): transposes used as the B operand of the data-gradient GEMMs
extern "C" int fiode_dyn_transposes(const fiode_dyn_weights* w, float* Q1T, float* QxT, float* Q2T, float* Q3T, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!w || !Q1T || !QxT || !Q2T || !Q3T) { set_error("fiode_dyn_transposes: null pointer"); return -1; }
    tr_kernel<<<8, 256, 0, st>>>(w->Q1, NM, NH, Q1T);
    tr_kernel<<<8, 256, 0, st>>>(w->Qx, NM, NH, QxT);
    tr_kernel<<<64, 256, 0, st>>>(w->Q2, NM, NM, Q2T);
    tr_kernel<<<8, 256, 0, st>>>(w->Q3, NH, NM, Q3T);
    FIODE_CUDA(cudaGetLastError());
    count_launch(4);
    return 0;
}

extern "C" int fiode_rhs_bwd_data(const fiode_dyn_weights* w, const float* h, long long N, const float* g_f, const float* f,
                                  const float* ftilde, const float* mu, const float* z1, const float* z2, const float* Q1T,
                                  const float* QxT, const float* Q2T, const float* Q3T, float* dft, float* d2, float* d1,
                                  float* g_h, float* g_s, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!w || !h || !g_f || !f || !ftilde || !mu || !z1 || !z2 || !Q1T || !QxT || !Q2T || !Q3T || !dft || !d2 || !d1 || !g_h || !ws) {
        set_error("fiode_rhs_bwd_data: null pointer"); return -1; }
    if (N <= 0 || N > 0x7fffffffLL / NM) { set_error("fiode_rhs_bwd_data: N=%lld", N); return -1; }
    if (ws_bytes < fiode_rhs_train_workspace_bytes(N)) { set_error("fiode_rhs_bwd_data: workspace too small"); return -1; }
    char* p = reinterpret_cast<char*>(ws) + 256;
    float* gh_direct = reinterpret_cast<float*>(p); p += al_((size_t)N * NH * 4);
    float* gws = reinterpret_cast<float*>(p);
    const size_t gwb = ws_bytes - (size_t)(p - reinterpret_cast<char*>(ws));
    const int Ni = (int)N;
    int rc;
    rhs_seed_kernel<<<gridn((size_t)N), 256, 0, st>>>(h, N, *w, g_f, f, ftilde, mu, dft, gh_direct);
    count_launch(1);
    // Q3T here is the (128,10) transpose of Q3 (10,128): d2 = dft Q3 = dft (Q3T)^T
    if ((rc = gemm_nt(dft, NH, Q3T, NH, Ni, NM, NH, nullptr, 1.f, 0, d2, NM, gws, gwb, st))) return rc;
    relu_bwd_kernel<<<gridn((size_t)N * NM), 256, 0, st>>>(d2, z2, (size_t)N * NM);
    if ((rc = gemm_nt(d2, NM, Q2T, NM, Ni, NM, NM, nullptr, 1.f, 0, d1, NM, gws, gwb, st))) return rc;
    relu_bwd_kernel<<<gridn((size_t)N * NM), 256, 0, st>>>(d1, z1, (size_t)N * NM);
    if ((rc = gemm_nt(d1, NM, Q1T, NM, Ni, NH, NM, nullptr, 1.f, 0, g_h, NH, gws, gwb, st))) return rc;
    add_inplace_kernel<<<gridn((size_t)N * NH), 256, 0, st>>>(g_h, gh_direct, (size_t)N * NH);
    count_launch(3);
    if (g_s && (rc = gemm_nt(d1, NM, QxT, NM, Ni, NH, NM, nullptr, 1.f, 0, g_s, NH, gws, gwb, st))) return rc;
    FIODE_CUDA(cudaGetLastError());
    return 0;
}

extern "C" size_t fiode_rhs_bwd_params_workspace_bytes(long long M) {
    return al_((size_t)M * 4) + al_(gemm_tn_workspace(NM, NM, M)) + 256;
}

// stacked evaluations (M rows in total): dQ3 = dft^T z2, dQ2 = d2^T z1, dQ1 = d1^T h, dQx = d1^T s, biases = column sums
extern "C" int fiode_rhs_bwd_params(const float* h, const float* s, const float* z1, const float* z2, const float* dft,
                                    const float* d2, const float* d1, long long M, float* dQ1, float* dQx, float* dQ2,
                                    float* dQ3, float* db1, float* db2, float* db3, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!h || !s || !z1 || !z2 || !dft || !d2 || !d1 || !dQ1 || !dQx || !dQ2 || !dQ3 || !db1 || !db2 || !db3 || !ws) {
        set_error("fiode_rhs_bwd_params: null pointer"); return -1; }
    if (M <= 0 || ws_bytes < fiode_rhs_bwd_params_workspace_bytes(M)) { set_error("fiode_rhs_bwd_params: bad arguments"); return -1; }
    float* ones = reinterpret_cast<float*>(ws);
    float* gws = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + al_((size_t)M * 4));
    const size_t gwb = ws_bytes - al_((size_t)M * 4);
    fill_ones_kernel<<<gridn((size_t)M), 256, 0, st>>>(ones, (size_t)M);
    count_launch(1);
    int rc;
    if ((rc = gemm_tn(dft, NH, ones, 1, NH, 1, M, 1.f, db3, 1, gws, gwb, st))) return rc;
    if ((rc = gemm_tn(dft, NH, z2, NM, NH, NM, M, 1.f, dQ3, NM, gws, gwb, st))) return rc;
    if ((rc = gemm_tn(d2, NM, ones, 1, NM, 1, M, 1.f, db2, 1, gws, gwb, st))) return rc;
    if ((rc = gemm_tn(d2, NM, z1, NM, NM, NM, M, 1.f, dQ2, NM, gws, gwb, st))) return rc;
    if ((rc = gemm_tn(d1, NM, ones, 1, NM, 1, M, 1.f, db1, 1, gws, gwb, st))) return rc;
    if ((rc = gemm_tn(d1, NM, h, NH, NM, NH, M, 1.f, dQ1, NH, gws, gwb, st))) return rc;
    if ((rc = gemm_tn(d1, NM, s, NH, NM, NH, M, 1.f, dQx, NH, gws, gwb, st))) return rc;
    FIODE_CUDA(cudaGetLastError());
    return 0;
}

// out = sum_i c[i] * x[i]  (n_terms <= 8, n elements; out may alias an input)
extern "C" int fiode_lincomb(float* out, long long n, int n_terms, const float* const* x, const float* c, void* stream) {
    if (!out || !x || !c || n <= 0 || n_terms < 1 || n_terms > 8) { set_error("fiode_lincomb: bad arguments"); return -1; }
    LinComb a{};
    a.n = n_terms;
    for (int i = 0; i < n_terms; ++i) { if (!x[i]) { set_error("fiode_lincomb: null term"); return -1; } a.x[i] = x[i]; a.c[i] = c[i]; }
    lincomb_kernel<<<gridn((size_t)n), 256, 0, (cudaStream_t)stream>>>(a, out, (size_t)n);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
