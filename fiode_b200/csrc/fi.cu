// fiode_b200/csrc/fi.cu -- the forward-invariance training workload: simplex samplers, the
// DecisionBoundary Lyapunov candidate, the FI loss and its hand-written backward.
//
// Replaces
//   * sampling/sampler.py:24-38 UniformSimplexSampling, :104-128 CorrectConeSampling,
//     :130-153 DecisionBoundarySampling, :195-216 CompositeSampler (composition of the parts)
//   * lya_cands.py:72-94 DecisionBoundary (on_simplex=True):  V = 1 + max_{j!=y} h_j - h_y
//   * pl_modules.py:390-484 LyapunovLearning.compute_loss (order 1, act relu):
//       vdot = jvp(V, h, f) == f[argmax_{j!=y} h_j] - f[y]   
//       loss = mean relu(vdot + kappa V); effective_batch_size; mean_active_constraints (:474-483)
//   * the autograd graph behind loss.backward() down to the Cayley matrices, the biases and the static
//     feature s (barrier_projection.py:272-311, classification.py:96-115).
#include "common.cuh"
#include "rows.cuh"

namespace fiode {

int gemm_nt(const float* A, int lda, const float* Bm, int ldb, int M, int N, int K, const float* bias, float alpha,
            int group_sort, float* y, int ldy, float* ws, size_t ws_bytes, cudaStream_t st);
size_t gemm_nt_workspace(int M, int N, int K);
int gemm_tn(const float* A, int lda, const float* Bm, int ldb, int M, int N, long long K, float alpha, float* y, int ldy,
            float* ws, size_t ws_bytes, cudaStream_t st);
size_t gemm_tn_workspace(int M, int N, long long K);

// ---- samplers: pure functions of pre-drawn Exp(1) variates ------------------------------------------
// F.normalize(e, p=1): e / max(sum |e|, 1e-12)
__device__ __forceinline__ void l1_normalize(float (&v)[NH]) {
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < NH; ++j) sum += fabsf(v[j]);
    const float d = fmaxf(sum, 1e-12f);
#pragma unroll
    for (int j = 0; j < NH; ++j) v[j] = v[j] / d;
}

// kind 0: uniform simplex, e (n,10) shared by every image (sampler.py:34-38, :209-210)
// kind 1: correct cone, e (B,n,10): row maximum swapped into the label slot (sampler.py:113-125)
// kind 2: decision boundary, e (B,n,9): z1 = max(zs); [z1, zs] / sum; label slot <- z1 (sampler.py:139-150)
__global__ void __launch_bounds__(256) sampler_fill_kernel(int kind, const float* __restrict__ e,
                                                           const int* __restrict__ y, int B, int n, int S, int s0,
                                                           float* __restrict__ h) {
    const long long total = (long long)B * n;
    for (long long idx = blockIdx.x * 256ll + threadIdx.x; idx < total; idx += (long long)gridDim.x * 256) {
        const int b = (int)(idx / n), i = (int)(idx - (long long)b * n);
        float v[NH];
        if (kind == 0) {
#pragma unroll
            for (int j = 0; j < NH; ++j) v[j] = e[(size_t)i * NH + j];
            l1_normalize(v);
        } else if (kind == 1) {
#pragma unroll
            for (int j = 0; j < NH; ++j) v[j] = e[((size_t)b * n + i) * NH + j];
            l1_normalize(v);
            const int lab = y[b];
            int im = 0; float hm = v[0];
#pragma unroll
            for (int j = 1; j < NH; ++j) if (v[j] > hm) { hm = v[j]; im = j; }
            float hl = 0.f;
#pragma unroll
            for (int j = 0; j < NH; ++j) if (j == lab) hl = v[j];
#pragma unroll
            for (int j = 0; j < NH; ++j) {       // scatter label first, then the arg-max slot
                if (j == lab) v[j] = hm;
                if (j == im) v[j] = hl;
            }
            if (im == lab) {
#pragma unroll
                for (int j = 0; j < NH; ++j) if (j == lab) v[j] = hm;
            }
        } else {
            float raw[NH];
            float z1 = -INFINITY;
#pragma unroll
            for (int j = 0; j < NH - 1; ++j) { raw[j + 1] = e[((size_t)b * n + i) * (NH - 1) + j]; z1 = fmaxf(z1, raw[j + 1]); }
            raw[0] = z1;
            l1_normalize(raw);
            const int lab = y[b];
            int src = 1;
#pragma unroll
            for (int j = 0; j < NH; ++j) {
                if (j == lab) v[j] = raw[0];
                else { float x = 0.f;
#pragma unroll
                    for (int q = 1; q < NH; ++q) if (q == src) x = raw[q];
                    v[j] = x; ++src; }
            }
        }
        float* dst = h + ((size_t)b * S + s0 + i) * NH;
#pragma unroll
        for (int j = 0; j < NH; ++j) dst[j] = v[j];
    }
}

// ---- FI loss forward: finish the projection, V, Vdot, violations, block partials ---------------------
struct FiPart { double loss; unsigned long long eff; unsigned long long mac; };

__global__ void __launch_bounds__(256) fi_finish_kernel(const float* __restrict__ ft, const float* __restrict__ h,
                                                        const int* __restrict__ y, long long N, int S,
                                                        fiode_dyn_weights w, float kappa_h,
                                                        const float* __restrict__ kappa_d,
                                                        const unsigned* __restrict__ notconv, float* __restrict__ f,
                                                        float* __restrict__ mu_out, float* __restrict__ V,
                                                        float* __restrict__ Vdot, FiPart* __restrict__ part) {
    __shared__ double sl[8];
    __shared__ unsigned long long se[8], sm[8];
    const int K = stop_iter(*notconv, w.max_iter);
    const float kappa = kappa_d ? *kappa_d : kappa_h;
    double loss = 0; unsigned long long eff = 0, mac = 0;
    for (long long r = blockIdx.x * 256ll + threadIdx.x; r < N; r += (long long)gridDim.x * 256) {
        float v[NH], lo[NH], hh[NH];
#pragma unroll
        for (int i = 0; i < NH; ++i) {
            hh[i] = h[r * NH + i];
            v[i] = ft[r * NH + i];
            lo[i] = barrier_lower(w.alpha_1, w.sigma_1, hh[i]);
        }
        float hi = v[0] - lo[0], lw = v[0], mu = 0.f;
#pragma unroll
        for (int j = 1; j < NH; ++j) { hi = fmaxf(hi, v[j] - lo[j]); lw = fminf(lw, v[j]); }
        for (int k = 0; k <= K; ++k) {
            mu = (hi - lw) / 2.f + lw;
            float e = 0.f;
#pragma unroll
            for (int j = 0; j < NH; ++j) e += fmaxf(v[j] - mu, lo[j]);
            if (e > 0.f) lw = mu;
            if (e < 0.f) hi = mu;
        }
        const int lab = y[r / S];
        int wi = -1; float hw = -INFINITY, hy = 0.f, fw = 0.f, fy = 0.f;
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            const float fj = fmaxf(v[j] - mu, lo[j]);
            f[r * NH + j] = fj;
            if (j == lab) { hy = hh[j]; fy = fj; }
            else if (hh[j] > hw) { hw = hh[j]; fw = fj; wi = j; }
            // pl_modules.py:478-481: the logged statistic uses the LINEAR lower bound -alpha_1 h
            const float lin = -w.alpha_1 * hh[j], up = w.alpha_2 * (1.f - hh[j]);
            if (fabsf(fj - lin) <= 1e-6f || fabsf(fj - up) <= 1e-6f) ++mac;
        }
        (void)wi;
        const float Vr = 1.f + hw - hy, Vd = fw - fy;
        const float viol = fmaxf(Vd + kappa * Vr, 0.f);
        mu_out[r] = mu; V[r] = Vr; Vdot[r] = Vd;
        loss += (double)viol;
        if (viol > 0.f) ++eff;
    }
    loss = warp_sum(loss);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { eff += __shfl_xor_sync(0xffffffffu, eff, o); mac += __shfl_xor_sync(0xffffffffu, mac, o); }
    const int wp = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { sl[wp] = loss; se[wp] = eff; sm[wp] = mac; }
    __syncthreads();
    if (threadIdx.x == 0) {
        FiPart p{0, 0, 0};
        for (int i = 0; i < 8; ++i) { p.loss += sl[i]; p.eff += se[i]; p.mac += sm[i]; }
        part[blockIdx.x] = p;
    }
}
// out[0] = loss (mean), out[1] = effective_batch_size, out[2] = mean_active_constraints, out[3] = bisection iterations
__global__ void fi_reduce_kernel(const FiPart* __restrict__ part, int np, long long N, const unsigned* __restrict__ notconv,
                                 int max_iter, float* __restrict__ out, unsigned long long* __restrict__ drop_step) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        if (drop_step) *drop_step += 1ull;      // next call (or next replay of a captured graph) draws fresh dropout masks
        double loss = 0; unsigned long long eff = 0, mac = 0;
        for (int i = 0; i < np; ++i) { loss += part[i].loss; eff += part[i].eff; mac += part[i].mac; }
        out[0] = (float)(loss / (double)N);
        out[1] = (float)eff;
        out[2] = (float)((double)mac / ((double)N * NH));
        out[3] = (float)(stop_iter(*notconv, max_iter) + 1);
    }
}

// ---- backward ---------------------------------------------------------------------------------------------
// d loss / d ftilde_raw for every row (App. C-4, C-5): loss seed -> projection Jacobian -> sigmoid rescale
__global__ void __launch_bounds__(256) fi_seed_kernel(const float* __restrict__ h, const int* __restrict__ y,
                                                      long long N, int S, fiode_dyn_weights w, float kappa_h,
                                                      const float* __restrict__ kappa_d,
                                                      float gscale, const float* __restrict__ f,
                                                      const float* __restrict__ ft, const float* __restrict__ mu,
                                                      const float* __restrict__ V, const float* __restrict__ Vdot,
                                                      float* __restrict__ dft, float* __restrict__ ones) {
    const float kappa = kappa_d ? *kappa_d : kappa_h;
    for (long long r = blockIdx.x * 256ll + threadIdx.x; r < N; r += (long long)gridDim.x * 256) {
        ones[r] = 1.f;
        const bool on = (Vdot[r] + kappa * V[r]) > 0.f;
        const float g = on ? gscale / (float)N : 0.f;
        const int lab = y[r / S];
        float hh[NH]; int wi = 0; float hw = -INFINITY;
#pragma unroll
        for (int j = 0; j < NH; ++j) { hh[j] = h[r * NH + j]; if (j != lab && hh[j] > hw) { hw = hh[j]; wi = j; } }
        const float m = mu[r];
        float gj[NH]; bool act[NH]; float sf = 0.f; int nf = 0;
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            gj[j] = (j == wi ? g : 0.f) - (j == lab ? g : 0.f);
            act[j] = (f[r * NH + j] - ft[r * NH + j] + m) > 0.f;
            if (!act[j]) { sf += gj[j]; ++nf; }
        }
        const float mean = nf > 0 ? sf / (float)nf : 0.f;
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            float d = act[j] ? 0.f : gj[j] - mean;
            if (w.scale_nominal) {   // f' = (up - lo) sigmoid(x) + lo  =>  df'/dx = (f' - lo)(up - f') / (up - lo)
                const float lo = barrier_lower(w.alpha_1, w.sigma_1, hh[j]), up = w.alpha_2 * (1.f - hh[j]);
                const float fp = ft[r * NH + j];
                d *= (fp - lo) * (up - fp) / (up - lo);
            }
            dft[r * NH + j] = d;
        }
    }
}
// da = dz * 1[z > 0] * mask (in place on dz)
// (in-kernel dropout: z > 0 implies the unit was kept, so its mask value is the constant 1 / (1 - p) = `scale`)
__global__ void __launch_bounds__(256) relu_mask_bwd_kernel(float* __restrict__ dz, const float* __restrict__ z,
                                                            const float* __restrict__ mask, float scale, size_t n) {
    for (size_t e = blockIdx.x * 256ull + threadIdx.x; e < n; e += (size_t)gridDim.x * 256) {
        float d = z[e] > 0.f ? dz[e] * scale : 0.f;
        if (mask) d *= mask[e];
        dz[e] = d;
    }
}
// D[b][c] = sum over the S rows of image b of da1[row][c]  (sequential, deterministic)
__global__ void __launch_bounds__(128) segment_sum_kernel(const float* __restrict__ da1, int B, int S,
                                                          float* __restrict__ D) {
    const int b = blockIdx.x, c = threadIdx.x;
    if (b >= B) return;
    const float* src = da1 + (size_t)b * S * NM + c;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    int r = 0;
    for (; r + 3 < S; r += 4) { a0 += src[(size_t)r * NM]; a1 += src[(size_t)(r + 1) * NM]; a2 += src[(size_t)(r + 2) * NM]; a3 += src[(size_t)(r + 3) * NM]; }
    for (; r < S; ++r) a0 += src[(size_t)r * NM];
    D[(size_t)b * NM + c] = (a0 + a1) + (a2 + a3);
}
__global__ void small_transpose_kernel(const float* __restrict__ x, int rows, int cols, float* __restrict__ y) {
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < rows * cols; e += gridDim.x * blockDim.x) {
        const int r = e / cols, c = e - r * cols;
        y[c * rows + r] = x[e];
    }
}

static size_t al(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr int kFiBlocks = 592;

// lya_cands.py:72-94 DecisionBoundary.forward stand-alone: V = 1 + max_{j != y} p_j - p_y (log_mode: log of it), p = the state
// itself (on_simplex) or its softmax
__global__ void __launch_bounds__(256) decision_boundary_kernel(const float* __restrict__ h, const int* __restrict__ y, long long N,
                                                                int on_simplex, int log_mode, float* __restrict__ V) {
    for (long long r = blockIdx.x * 256ll + threadIdx.x; r < N; r += (long long)gridDim.x * 256) {
        float p[NH];
#pragma unroll
        for (int j = 0; j < NH; ++j) p[j] = h[r * NH + j];
        if (!on_simplex) {
            float mx = p[0];
#pragma unroll
            for (int j = 1; j < NH; ++j) mx = fmaxf(mx, p[j]);
            float sum = 0.f;
#pragma unroll
            for (int j = 0; j < NH; ++j) { p[j] = expf(p[j] - mx); sum += p[j]; }
#pragma unroll
            for (int j = 0; j < NH; ++j) p[j] = p[j] / sum;
        }
        const int yy = y[r];
        float py = 0.f, mw = -INFINITY;
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            if (j == yy) py = p[j];
            else mw = fmaxf(mw, p[j]);
        }
        const float v = 1.f + mw - py;
        V[r] = log_mode ? logf(v) : v;
    }
}
}  // namespace fiode


using namespace fiode;

extern "C" int fiode_sampler_fill(int kind, const float* e, const int* y, int B, int n, int S, int s0, float* h,
                                  void* stream) {
    if (!e || !h || (kind != 0 && !y)) { set_error("fiode_sampler_fill: null pointer"); return -1; }
    if (kind < 0 || kind > 2 || B <= 0 || n < 0 || s0 < 0 || s0 + n > S) { set_error("fiode_sampler_fill: bad arguments"); return -1; }
    if (n == 0) return 0;
    const long long want = ((long long)B * n + 255) / 256;
    sampler_fill_kernel<<<(int)(want < 2048 ? want : 2048), 256, 0, (cudaStream_t)stream>>>(kind, e, y, B, n, S, s0, h);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}

extern "C" size_t fiode_fi_loss_workspace_bytes(long long N, int B) {
    size_t b = 256 + al(kFiBlocks * sizeof(FiPart));
    b += al((size_t)N * NH * 4) + al((size_t)N * 4);             // dft, ones
    b += 2 * al((size_t)N * NM * 4);                             // dz2/da2, dz1/da1
    b += al((size_t)B * NM * 4);                                 // D
    b += 3 * al((size_t)NM * NM * 4);                            // Q3^T, Q2^T, Qx^T
    // the data-gradient products and the weight-gradient products run side by side: one split-K workspace each
    return b + al(gemm_nt_workspace((int)N, NM, NM)) + al(gemm_tn_workspace(NM, NM, N));
}

extern "C" int fiode_fi_loss_fwd(const fiode_dyn_weights* w, const float* h, const float* s, const int* y, int B, int S,
                                 float kappa, const float* m1, const float* m2, float* f, float* ftilde, float* mu,
                                 float* z1, float* z2, float* V, float* Vdot, float* out4, void* ws, size_t ws_bytes,
                                 void* stream) {
    return fiode_fi_loss_fwd_opt(w, h, s, y, B, S, kappa, m1, m2, nullptr, f, ftilde, mu, z1, z2, V, Vdot, out4, ws, ws_bytes, stream);
}

static int check_fi_options(const char* who, const fiode_fi_options* o, const float* m1, const float* m2) {
    if (!o) return 0;
    if (!(o->dropout_p >= 0.f) || o->dropout_p >= 1.f) { set_error("%s: dropout_p must be in [0, 1)", who); return -1; }
    if (o->dropout_p > 0.f && (m1 || m2)) { set_error("%s: pass either dropout masks or dropout_p, not both", who); return -1; }
    return 0;
}

extern "C" int fiode_fi_loss_fwd_opt(const fiode_dyn_weights* w, const float* h, const float* s, const int* y, int B, int S,
                                     float kappa, const float* m1, const float* m2, const fiode_fi_options* opt, float* f,
                                     float* ftilde, float* mu, float* z1, float* z2, float* V, float* Vdot, float* out4,
                                     void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (check_fi_options("fiode_fi_loss_fwd_opt", opt, m1, m2)) return -1;
    const float* kappa_d = opt ? opt->kappa_dev : nullptr;
    const float drop_p = opt ? opt->dropout_p : 0.f;
    unsigned long long* step = (opt && drop_p > 0.f) ? opt->step_dev : nullptr;
    if (!w || !h || !s || !y || !f || !ftilde || !mu || !V || !Vdot || !out4 || !ws) { set_error("fiode_fi_loss_fwd: null pointer"); return -1; }
    if (B <= 0 || S <= 0) { set_error("fiode_fi_loss_fwd: B=%d S=%d", B, S); return -1; }
    const long long N = (long long)B * S;
    if (ws_bytes < fiode_fi_loss_workspace_bytes(N, B)) { set_error("fiode_fi_loss_fwd: workspace too small"); return -1; }
    unsigned* notconv = reinterpret_cast<unsigned*>(ws);
    FiPart* part = reinterpret_cast<FiPart*>(reinterpret_cast<char*>(ws) + 256);
    int rc = launch_rows_store(w, h, s, N, B, S, ftilde, notconv, m1, m2, z1, z2, st, drop_p, opt ? opt->seed : 0ull, step);
    if (rc) return rc;
    const long long want = (N + 255) / 256;
    const int grid = (int)(want < kFiBlocks ? want : kFiBlocks);
    fi_finish_kernel<<<grid, 256, 0, st>>>(ftilde, h, y, N, S, *w, kappa, kappa_d, notconv, f, mu, V, Vdot, part);
    fi_reduce_kernel<<<1, 32, 0, st>>>(part, grid, N, notconv, w->max_iter, out4, step);
    FIODE_CUDA(cudaGetLastError());
    count_launch(2);
    return 0;
}

extern "C" int fiode_fi_loss_bwd(const fiode_dyn_weights* w, const float* h, const float* s, const int* y, int B, int S,
                                 float kappa, float grad_loss, const float* m1, const float* m2, const float* f,
                                 const float* ftilde, const float* mu, const float* z1, const float* z2, const float* V,
                                 const float* Vdot, float* dQ1, float* dQx, float* dQ2, float* dQ3, float* db1, float* db2,
                                 float* db3, float* ds, void* ws, size_t ws_bytes, void* stream) {
    return fiode_fi_loss_bwd_opt(w, h, s, y, B, S, kappa, grad_loss, m1, m2, nullptr, f, ftilde, mu, z1, z2, V, Vdot, dQ1, dQx,
                                 dQ2, dQ3, db1, db2, db3, ds, ws, ws_bytes, stream);
}

extern "C" int fiode_fi_loss_bwd_opt(const fiode_dyn_weights* w, const float* h, const float* s, const int* y, int B, int S,
                                     float kappa, float grad_loss, const float* m1, const float* m2,
                                     const fiode_fi_options* opt, const float* f, const float* ftilde, const float* mu,
                                     const float* z1, const float* z2, const float* V, const float* Vdot, float* dQ1,
                                     float* dQx, float* dQ2, float* dQ3, float* db1, float* db2, float* db3, float* ds,
                                     void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (check_fi_options("fiode_fi_loss_bwd_opt", opt, m1, m2)) return -1;
    const float* kappa_d = opt ? opt->kappa_dev : nullptr;
    const float dscale = (opt && opt->dropout_p > 0.f) ? 1.f / (1.f - opt->dropout_p) : 1.f;
    if (!w || !h || !s || !y || !f || !ftilde || !mu || !z1 || !z2 || !V || !Vdot || !dQ1 || !dQx || !dQ2 || !dQ3 || !db1 ||
        !db2 || !db3 || !ds || !ws) { set_error("fiode_fi_loss_bwd: null pointer"); return -1; }
    const long long N = (long long)B * S;
    if (N > 0x7fffffffLL / NM) { set_error("fiode_fi_loss_bwd: N too large"); return -1; }
    if (ws_bytes < fiode_fi_loss_workspace_bytes(N, B)) { set_error("fiode_fi_loss_bwd: workspace too small"); return -1; }
    char* p = reinterpret_cast<char*>(ws) + 256 + al(kFiBlocks * sizeof(FiPart));
    auto take = [&](size_t bytes) { void* r = p; p += al(bytes); return reinterpret_cast<float*>(r); };
    float* dft = take((size_t)N * NH * 4);
    float* ones = take((size_t)N * 4);
    float* d2 = take((size_t)N * NM * 4);
    float* d1 = take((size_t)N * NM * 4);
    float* D = take((size_t)B * NM * 4);
    float* Q3T = take((size_t)NM * NM * 4);
    float* Q2T = take((size_t)NM * NM * 4);
    float* QxT = take((size_t)NM * NM * 4);
    float* gws = reinterpret_cast<float*>(p);                       // data-gradient chain (caller's stream)
    const size_t gwb = al(gemm_nt_workspace((int)N, NM, NM));
    float* gws2 = reinterpret_cast<float*>(p + gwb);                // weight / bias gradients (branch)
    const size_t gwb2 = ws_bytes - (size_t)(p + gwb - reinterpret_cast<char*>(ws));
    const int Ni = (int)N;
    const long long want = (N + 255) / 256;
    const int grid = (int)(want < 4096 ? want : 4096);
    int rc;
    fi_seed_kernel<<<grid, 256, 0, st>>>(h, y, N, S, *w, kappa, kappa_d, grad_loss, f, ftilde, mu, V, Vdot, dft, ones);
    small_transpose_kernel<<<8, 256, 0, st>>>(w->Q3, NH, NM, Q3T);     // (10,128) -> (128,10)
    small_transpose_kernel<<<64, 256, 0, st>>>(w->Q2, NM, NM, Q2T);
    small_transpose_kernel<<<8, 256, 0, st>>>(w->Qx, NM, NH, QxT);     // (128,10) -> (10,128)
    count_launch(4);
    // The chain dft -> d2 -> d1 -> D -> ds stays on the caller's stream; each layer's weight and bias gradients (reductions over
    // all N rows, leaves of the backward pass) go to a branch that re-synchronises with the chain whenever its next input is ready.
    // layer 3
    cudaStream_t sd = fork_side(st);
    if ((rc = gemm_tn(dft, NH, ones, 1, NH, 1, N, 1.f, db3, 1, gws2, gwb2, sd))) return rc;
    if ((rc = gemm_tn(dft, NH, z2, NM, NH, NM, N, 1.f, dQ3, NM, gws2, gwb2, sd))) return rc;
    if ((rc = gemm_nt(dft, NH, Q3T, NH, Ni, NM, NH, nullptr, 1.f, 0, d2, NM, gws, gwb, st))) return rc;
    relu_mask_bwd_kernel<<<4096, 256, 0, st>>>(d2, z2, m2, dscale, (size_t)N * NM);
    // layer 2
    if (sd != st) sd = fork_side(st);
    if ((rc = gemm_tn(d2, NM, ones, 1, NM, 1, N, 1.f, db2, 1, gws2, gwb2, sd))) return rc;
    if ((rc = gemm_tn(d2, NM, z1, NM, NM, NM, N, 1.f, dQ2, NM, gws2, gwb2, sd))) return rc;
    if ((rc = gemm_nt(d2, NM, Q2T, NM, Ni, NM, NM, nullptr, 1.f, 0, d1, NM, gws, gwb, st))) return rc;
    relu_mask_bwd_kernel<<<4096, 256, 0, st>>>(d1, z1, m1, dscale, (size_t)N * NM);
    // layer 1: z = Q1 h + b1 + Qx s + bx
    if (sd != st) sd = fork_side(st);
    if ((rc = gemm_tn(d1, NM, h, NH, NM, NH, N, 1.f, dQ1, NH, gws2, gwb2, sd))) return rc;
    segment_sum_kernel<<<B, NM, 0, st>>>(d1, B, S, D);
    count_launch(3);
    if (sd != st) sd = fork_side(st);
    if ((rc = gemm_tn(D, NM, ones, 1, NM, 1, B, 1.f, db1, 1, gws2, gwb2, sd))) return rc;
    if ((rc = gemm_tn(D, NM, s, NH, NM, NH, B, 1.f, dQx, NH, gws2, gwb2, sd))) return rc;
    if ((rc = gemm_nt(D, NM, QxT, NM, B, NH, NM, nullptr, 1.f, 0, ds, NH, gws, gwb, st))) return rc;
    if (sd != st && (rc = join_side(st))) return rc;
    FIODE_CUDA(cudaGetLastError());
    return 0;
}


// The masks the forward draws for (seed, step): mask (N,128) of layer 0 (after hidden_to_mlp + U_x) or 1 (after mlp_to_mlp),
// values 0 or 1 / (1 - p).  Lets a caller (and the parity tests) replay an in-kernel-dropout step with explicit masks.
namespace fiode {
__global__ void __launch_bounds__(256) dropout_mask_kernel(DropCfg d, int layer, long long N, float* __restrict__ out) {
    for (long long e = blockIdx.x * 256ll + threadIdx.x; e < N * (NM / 4); e += (long long)gridDim.x * 256) {
        const long long row = e / (NM / 4); const int k4 = (int)(e - row * (NM / 4));
        *reinterpret_cast<float4*>(out + row * NM + k4 * 4) = dropout_mask4(d, row, k4, layer);
    }
}
}  // namespace fiode
extern "C" int fiode_dropout_mask(float dropout_p, unsigned long long seed, const unsigned long long* step_dev, int layer,
                                  long long N, float* mask, void* stream) {
    if (!mask || !step_dev || N <= 0 || (layer != 0 && layer != 1) || !(dropout_p > 0.f) || dropout_p >= 1.f) {
        set_error("fiode_dropout_mask: bad arguments"); return -1; }
    unsigned long long step = 0;
    FIODE_CUDA(cudaMemcpyAsync(&step, step_dev, sizeof(step), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    FIODE_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    fiode::DropCfg d{(unsigned)fminf(dropout_p * 4294967296.f, 4294967040.f), 1.f / (1.f - dropout_p), seed, step};
    const long long want = (N * (NM / 4) + 255) / 256;
    fiode::dropout_mask_kernel<<<(unsigned)(want < 4096 ? want : 4096), 256, 0, (cudaStream_t)stream>>>(d, layer, N, mask);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}

extern "C" int fiode_decision_boundary(const float* h, const int* y, long long N, int on_simplex, int log_mode, float* V,
                                       void* stream) {
    if (!h || !y || !V) { set_error("fiode_decision_boundary: null pointer"); return -1; }
    if (N <= 0) { set_error("fiode_decision_boundary: N=%lld", N); return -1; }
    const long long want = (N + 255) / 256;
    decision_boundary_kernel<<<(int)(want < 4096 ? want : 4096), 256, 0, (cudaStream_t)stream>>>(h, y, N, on_simplex, log_mode, V);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
