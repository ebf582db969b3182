// fiode_b200/csrc/crown.cu -- the interval step that follows the CROWN bound in robustness/certify_crown.py:139-150.
//
// auto_LiRPA's CROWN pass itself stays on the reference (north_star); what it returns -- element-wise bounds lb <= ftilde <= ub
// of the raw MLP output over the box [eta - eps, eta + eps] -- is turned into bounds on the projected dynamics f here:
//   * ibp_sigmoid     dynamics/classification.py:175-181   (scale_nominal models)
//   * ibp_cbf_qp      dynamics/classification.py:208-242   (upper=False, the variant certify_crown.py:143 calls)
//   * perturbed_vdot  robustness/certify_crown.py:29-34
// ibp_cbf_qp builds, for every grid point n and coordinate i, one projection problem whose coordinate i takes the
// worst case (lower bound of ftilde_i, barrier at eta_i + eps) and every other coordinate the opposite extreme, solves the
// 10 N problems with FastBarrierProjectionNoUpper (ONE batch-global stop iteration per call, barrier_projection.py:247) and
// keeps coordinate i of problem (n, i).  The reference materialises the (10 N, 10) operands three times; here a thread
// owns problem (n, i) in registers: pass 1 finds the not-converged masks of both calls, pass 2 replays the bisection to the
// global stop iteration and writes f_lb[n][i], f_ub[n][i].
#include "common.cuh"

namespace fiode {

struct IbpRow { float lo[NH], ft[NH]; };

// problem (n, i) of the lower-bound call (which = 0) or of the upper-bound call (which = 1)
__device__ __forceinline__ IbpRow ibp_row(const float* __restrict__ h, const float* __restrict__ lb, const float* __restrict__ ub,
                                          long long n, int i, float eps, float alpha_1, float sigma_1, int which) {
    IbpRow r;
#pragma unroll
    for (int j = 0; j < NH; ++j) {
        const float hj = h[n * NH + j];
        const float hm = hj - eps, hp = hj + eps;                 // h_lower / h_upper before the diagonal swap (:210-217)
        float hv, fv;
        if (which == 0) { hv = (j == i) ? hp : hm; fv = (j == i) ? lb[n * NH + j] : ub[n * NH + j]; }
        else            { hv = (j == i) ? hm : hp; fv = (j == i) ? ub[n * NH + j] : lb[n * NH + j]; }
        r.lo[j] = barrier_lower(alpha_1, sigma_1, hv);             // :224-225
        r.ft[j] = fv;                                              // :226-231
    }
    return r;
}

__global__ void __launch_bounds__(256) ibp_pass1_kernel(const float* __restrict__ h, const float* __restrict__ lb,
                                                        const float* __restrict__ ub, long long N, float eps, float alpha_1,
                                                        float sigma_1, int max_iter, float tol, unsigned* __restrict__ notconv2) {
    unsigned w0 = 0, w1 = 0;
    for (long long e = blockIdx.x * 256ll + threadIdx.x; e < N * NH; e += (long long)gridDim.x * 256) {
        const long long n = e / NH;
        const int i = (int)(e - n * NH);
#pragma unroll
        for (int which = 0; which < 2; ++which) {
            const IbpRow r = ibp_row(h, lb, ub, n, i, eps, alpha_1, sigma_1, which);
            float hi = r.ft[0] - r.lo[0], lw = r.ft[0];
#pragma unroll
            for (int j = 1; j < NH; ++j) { hi = fmaxf(hi, r.ft[j] - r.lo[j]); lw = fminf(lw, r.ft[j]); }
            unsigned nc = 0;
            for (int k = 0; k < max_iter; ++k) {
                const float mu = (hi - lw) / 2.f + lw;
                float s = 0.f;
#pragma unroll
                for (int j = 0; j < NH; ++j) s += fmaxf(r.ft[j] - mu, r.lo[j]);
                if (!(fabsf(s) < tol)) nc |= (1u << k);
                if (s > 0.f) lw = mu;
                if (s < 0.f) hi = mu;
            }
            if (which == 0) w0 |= nc; else w1 |= nc;
        }
    }
    w0 = __reduce_or_sync(0xffffffffu, w0);
    w1 = __reduce_or_sync(0xffffffffu, w1);
    if ((threadIdx.x & 31) == 0) {
        if (w0) atomicOr(notconv2, w0);
        if (w1) atomicOr(notconv2 + 1, w1);
    }
}

__global__ void __launch_bounds__(256) ibp_pass2_kernel(const float* __restrict__ h, const float* __restrict__ lb,
                                                        const float* __restrict__ ub, long long N, float eps, float alpha_1,
                                                        float sigma_1, int max_iter, const unsigned* __restrict__ notconv2,
                                                        float* __restrict__ f_lb, float* __restrict__ f_ub,
                                                        int* __restrict__ iters2) {
    const int K0 = stop_iter(notconv2[0], max_iter), K1 = stop_iter(notconv2[1], max_iter);
    if (iters2 && blockIdx.x == 0 && threadIdx.x == 0) { iters2[0] = K0 + 1; iters2[1] = K1 + 1; }
    for (long long e = blockIdx.x * 256ll + threadIdx.x; e < N * NH; e += (long long)gridDim.x * 256) {
        const long long n = e / NH;
        const int i = (int)(e - n * NH);
#pragma unroll
        for (int which = 0; which < 2; ++which) {
            const IbpRow r = ibp_row(h, lb, ub, n, i, eps, alpha_1, sigma_1, which);
            float hi = r.ft[0] - r.lo[0], lw = r.ft[0], mu = 0.f;
#pragma unroll
            for (int j = 1; j < NH; ++j) { hi = fmaxf(hi, r.ft[j] - r.lo[j]); lw = fminf(lw, r.ft[j]); }
            const int K = which ? K1 : K0;
            for (int k = 0; k <= K; ++k) {
                mu = (hi - lw) / 2.f + lw;
                float s = 0.f;
#pragma unroll
                for (int j = 0; j < NH; ++j) s += fmaxf(r.ft[j] - mu, r.lo[j]);
                if (s > 0.f) lw = mu;
                if (s < 0.f) hi = mu;
            }
            float vi = 0.f;
#pragma unroll
            for (int j = 0; j < NH; ++j) if (j == i) vi = fmaxf(r.ft[j] - mu, r.lo[j]);
            (which ? f_ub : f_lb)[e] = vi;
        }
    }
}

// ibp_sigmoid (classification.py:175-181): bounds of (upper - lower) sigmoid(ftilde) + lower over the box; separate roundings
// like the reference's separate torch ops
__global__ void ibp_sigmoid_kernel(const float* __restrict__ lb, const float* __restrict__ ub, const float* __restrict__ h,
                                   long long total, float eps, float alpha_1, float alpha_2, float sigma_1,
                                   float* __restrict__ out_lb, float* __restrict__ out_ub) {
    for (long long e = blockIdx.x * 256ll + threadIdx.x; e < total; e += (long long)gridDim.x * 256) {
        const float hl = h[e] - eps, hu = h[e] + eps;
        const float lower_lb = barrier_lower(alpha_1, sigma_1, hu), lower_ub = barrier_lower(alpha_1, sigma_1, hl);
        out_lb[e] = __fadd_rn(__fmul_rn(alpha_2 * (1.f - hu) - lower_lb, sigmoidf_(lb[e])), lower_lb);
        out_ub[e] = __fadd_rn(__fmul_rn(alpha_2 * (1.f - hl) - lower_ub, sigmoidf_(ub[e])), lower_ub);
    }
}

// perturbed_vdot (certify_crown.py:29-34): -f_lb[label] + max over the runner-up candidates of f_ub; also the running maximum
__global__ void __launch_bounds__(256) perturbed_vdot_kernel(const float* __restrict__ f_lb, const float* __restrict__ f_ub,
                                                             const int* __restrict__ label, int label_stride,
                                                             const uint8_t* __restrict__ runner_up, long long N,
                                                             float* __restrict__ vdot, unsigned* __restrict__ vmax_ord) {
    float m = -INFINITY;
    for (long long n = blockIdx.x * 256ll + threadIdx.x; n < N; n += (long long)gridDim.x * 256) {
        const int y = label[n * label_stride];
        float fw = -INFINITY;
#pragma unroll
        for (int j = 0; j < NH; ++j) if (runner_up[n * NH + j]) fw = fmaxf(fw, f_ub[n * NH + j]);
        const float v = -f_lb[n * NH + y] + fw;
        vdot[n] = v;
        m = fmaxf(m, v);
    }
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0 && vmax_ord) atomicMax(vmax_ord, f2ord(m));
}

// certify_crown.py:146-148: candidates that can become the runner-up inside the box: eta_j >= max_j eta - margin, label excluded
__global__ void runner_up_mask_kernel(const float* __restrict__ eta, long long N, int label, float margin,
                                      uint8_t* __restrict__ mask) {
    for (long long n = blockIdx.x * 256ll + threadIdx.x; n < N; n += (long long)gridDim.x * 256) {
        float m = eta[n * NH];
#pragma unroll
        for (int j = 1; j < NH; ++j) m = fmaxf(m, eta[n * NH + j]);
        const float thr = m - margin;
#pragma unroll
        for (int j = 0; j < NH; ++j) mask[n * NH + j] = (j != label && eta[n * NH + j] >= thr) ? 1 : 0;
    }
}

__global__ void ord_to_float_kernel(const unsigned* __restrict__ o, float* __restrict__ out) { out[0] = ord2f(o[0]); }

}  // namespace fiode

using namespace fiode;

static unsigned grid_for(long long total) {
    const long long b = (total + 255) / 256;
    return (unsigned)(b < 1 ? 1 : (b < 148 * 16 ? b : 148 * 16));
}

extern "C" size_t fiode_ibp_workspace_bytes(void) { return 256; }

extern "C" int fiode_ibp_cbf_qp(const float* h, const float* lb, const float* ub, long long N, float eps, float alpha_1,
                                float sigma_1, int max_iter, float tol, float* f_lb, float* f_ub, int* iters2, void* ws,
                                size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!h || !lb || !ub || !f_lb || !f_ub || !ws) { set_error("fiode_ibp_cbf_qp: null pointer"); return -1; }
    if (N <= 0 || max_iter < 1 || max_iter > 32 || ws_bytes < 8) { set_error("fiode_ibp_cbf_qp: bad arguments"); return -1; }
    unsigned* nc = reinterpret_cast<unsigned*>(ws);
    FIODE_CUDA(cudaMemsetAsync(nc, 0, 8, st));
    const unsigned g = grid_for(N * NH);
    ibp_pass1_kernel<<<g, 256, 0, st>>>(h, lb, ub, N, eps, alpha_1, sigma_1, max_iter, tol, nc);
    ibp_pass2_kernel<<<g, 256, 0, st>>>(h, lb, ub, N, eps, alpha_1, sigma_1, max_iter, nc, f_lb, f_ub, iters2);
    FIODE_CUDA(cudaGetLastError());
    count_launch(2);
    return 0;
}

extern "C" int fiode_ibp_sigmoid(const float* lb, const float* ub, const float* h, long long N, float eps, float alpha_1,
                                 float alpha_2, float sigma_1, float* out_lb, float* out_ub, void* stream) {
    if (!lb || !ub || !h || !out_lb || !out_ub || N <= 0) { set_error("fiode_ibp_sigmoid: bad arguments"); return -1; }
    ibp_sigmoid_kernel<<<grid_for(N * NH), 256, 0, (cudaStream_t)stream>>>(lb, ub, h, N * NH, eps, alpha_1, alpha_2, sigma_1,
                                                                           out_lb, out_ub);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}

extern "C" int fiode_perturbed_vdot(const float* f_lb, const float* f_ub, const int* label, int label_rows,
                                    const uint8_t* runner_up, long long N, float* vdot, float* vmax, void* ws, size_t ws_bytes,
                                    void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!f_lb || !f_ub || !label || !runner_up || !vdot || N <= 0) { set_error("fiode_perturbed_vdot: bad arguments"); return -1; }
    if (label_rows != 1 && label_rows != N) { set_error("fiode_perturbed_vdot: label_rows must be 1 or N"); return -1; }
    unsigned* ord = nullptr;
    if (vmax) {
        if (!ws || ws_bytes < 4) { set_error("fiode_perturbed_vdot: workspace too small"); return -1; }
        ord = reinterpret_cast<unsigned*>(ws);
        FIODE_CUDA(cudaMemsetAsync(ord, 0, 4, st));
    }
    perturbed_vdot_kernel<<<grid_for(N), 256, 0, st>>>(f_lb, f_ub, label, label_rows == 1 ? 0 : 1, runner_up, N, vdot, ord);
    count_launch(1);
    if (vmax) {
        ord_to_float_kernel<<<1, 1, 0, st>>>(ord, vmax);
        count_launch(1);
    }
    FIODE_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int fiode_runner_up_mask(const float* eta, long long N, int label, float margin, uint8_t* mask, void* stream) {
    if (!eta || !mask || N <= 0 || label < 0 || label >= NH) { set_error("fiode_runner_up_mask: bad arguments"); return -1; }
    runner_up_mask_kernel<<<grid_for(N), 256, 0, (cudaStream_t)stream>>>(eta, N, label, margin, mask);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
