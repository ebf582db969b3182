// fiode_b200/csrc/segway_train.cu -- the Segway barrier model at arbitrary points, with its gradients: what the
// adversarial training loop of control/train_segway.py:119-158 needs from the device.
//
//   val(x) = SegwaySingleBarrierModel.forward_adv(x) = sum(P^T (P f(x, u(x))) * x)      control_utils.py:97-111, :229-234
//   f      = Segway.forward                                                              segway_system.py:19-29
//   u      = NNController 3 -> 32 -> 32 -> 1 (ReLU)                                       nn_controller.py:6-34
//
//   fiode_segway_hdot        val at N points (PGD's forward, the loss's forward)
//   fiode_segway_hdot_bwd    d val / d x  (LinfPGDAttack's gradient step, train_segway.py:119-123, :145-146) and the
//                            gradients of  sum_r gout_r val_r  w.r.t. the controller weights and P (loss.backward(), :151)
//   fiode_segway_pgd_step    the sign-gradient step + the two projections of advertorch's perturb_iterative (ord = inf)
//
// One thread per point; the 1 226 parameter gradients are reduced per block in shared memory (rank-1 updates accumulated
// by the threads that own an entry, fixed order) and then across blocks by a second kernel: deterministic.
#include "common.cuh"

namespace fiode {

namespace {
constexpr int SH = 32;
constexpr int kNP = SH * 3 + SH + SH * SH + SH + SH + 1 + 9;        // W0 96 | b0 32 | W1 1024 | b1 32 | W2 32 | b2 1 | P 9 = 1226
constexpr int oW0 = 0, oB0 = 96, oW1 = 128, oB1 = 1152, oW2 = 1184, oB2 = 1216, oP = 1217;
constexpr int TB = 64;                                              // points per block (3 x TB x 33 floats of staging stay under 48 KB)

struct SegWt { float W0[SH * 3], b0[SH], W1[SH * SH], b1[SH], W2[SH], b2, P[9]; };

__device__ __forceinline__ void load_w(const fiode_segway_weights& w, SegWt* S) {
    for (int e = threadIdx.x; e < SH * 3; e += blockDim.x) S->W0[e] = w.W0[e];
    for (int e = threadIdx.x; e < SH * SH; e += blockDim.x) S->W1[e] = w.W1[e];
    for (int e = threadIdx.x; e < SH; e += blockDim.x) { S->b0[e] = w.b0[e]; S->b1[e] = w.b1[e]; S->W2[e] = w.W2[e]; }
    if (threadIdx.x < 9) S->P[threadIdx.x] = w.P[threadIdx.x];
    if (threadIdx.x == 0) S->b2 = w.b2[0];
    __syncthreads();
}

// forward with everything the backward needs kept by the caller
struct Fwd { float a1[SH], z2[SH], u, c, s, f0, f1, f2, n1, n2, d1, d2; };

__device__ __forceinline__ void forward(const SegWt* S, const float (&x)[3], Fwd& F) {
#pragma unroll
    for (int j = 0; j < SH; ++j) {
        float a = S->b0[j];
        a = fmaf(x[0], S->W0[j * 3 + 0], a); a = fmaf(x[1], S->W0[j * 3 + 1], a); a = fmaf(x[2], S->W0[j * 3 + 2], a);
        F.a1[j] = fmaxf(a, 0.f);
    }
    float u = S->b2;
#pragma unroll 4
    for (int j = 0; j < SH; ++j) {
        float a = S->b1[j];
#pragma unroll
        for (int k = 0; k < SH; ++k) a = fmaf(F.a1[k], S->W1[j * SH + k], a);
        F.z2[j] = a;
        u = fmaf(fmaxf(a, 0.f), S->W2[j], u);
    }
    F.u = u;
    const float phi = x[0], v = x[1], pd = x[2];
    F.c = (float)cos((double)phi); F.s = (float)sin((double)phi);
    const float c = F.c, s = F.s;
    F.f0 = pd;
    F.n1 = c * (-1.8f * u + 11.5f * v + 9.8f * s) - 10.9f * u + 68.4f * v - 1.2f * pd * pd * s;
    F.d1 = c - 24.7f;
    F.f1 = F.n1 / F.d1;
    F.n2 = (9.3f * u - 58.8f * v) * c + 38.6f * u - 234.5f * v - s * (208.3f + pd * pd * c);
    F.d2 = c * c - 24.7f;
    F.f2 = F.n2 / F.d2;
}
__device__ __forceinline__ float barrier(const SegWt* S, const float (&x)[3], const Fwd& F, float (&pf)[3], float (&px)[3]) {
    const float f[3] = {F.f0, F.f1, F.f2};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        pf[i] = fmaf(f[2], S->P[i * 3 + 2], fmaf(f[1], S->P[i * 3 + 1], f[0] * S->P[i * 3 + 0]));
        px[i] = fmaf(x[2], S->P[i * 3 + 2], fmaf(x[1], S->P[i * 3 + 1], x[0] * S->P[i * 3 + 0]));
    }
    return pf[0] * px[0] + pf[1] * px[1] + pf[2] * px[2];          // (P f) . (P x) == sum(P^T (P f) * x)
}

__global__ void __launch_bounds__(TB) segway_hdot_kernel(const fiode_segway_weights w, const float* __restrict__ xs, long long N,
                                                         float* __restrict__ val) {
    __shared__ SegWt S;
    load_w(w, &S);
    for (long long r = blockIdx.x * (long long)TB + threadIdx.x; r < N; r += (long long)gridDim.x * TB) {
        const float x[3] = {xs[3 * r], xs[3 * r + 1], xs[3 * r + 2]};
        Fwd F; float pf[3], px[3];
        forward(&S, x, F);
        val[r] = barrier(&S, x, F, pf, px);
    }
}

// One block = TB points (one pass; the grid covers N).  part: [gridDim.x][kNP] per-block parameter-gradient sums (or NULL).
__global__ void __launch_bounds__(TB) segway_hdot_bwd_kernel(const fiode_segway_weights w, const float* __restrict__ xs, long long N,
                                                             const float* __restrict__ gout, float* __restrict__ grad_x,
                                                             float* __restrict__ part) {
    __shared__ SegWt S;
    __shared__ float sA1[TB][SH + 1];      // layer-1 activations of the block's points
    __shared__ float sG2[TB][SH + 1];      // d / d z2
    __shared__ float sG1[TB][SH + 1];      // d / d z1
    __shared__ float sX[TB][4];            // x, g_u
    __shared__ float sPf[TB][8];           // g * (P f), g * (P x), pad
    load_w(w, &S);
    const int t = threadIdx.x;
    const long long r = blockIdx.x * (long long)TB + t;
    const bool on = r < N;
    float x[3] = {0.f, 0.f, 0.f};
    float g = 0.f;
    if (on) { x[0] = xs[3 * r]; x[1] = xs[3 * r + 1]; x[2] = xs[3 * r + 2]; g = gout ? gout[r] : 1.f; }
    Fwd F; float pf[3], px[3];
    forward(&S, x, F);
    barrier(&S, x, F, pf, px);
    // d val / d f = P^T (P x) =: m ;  d val / d x (direct) = P^T (P f) =: nvec
    float m[3], nvec[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        m[j] = fmaf(px[2], S.P[2 * 3 + j], fmaf(px[1], S.P[1 * 3 + j], px[0] * S.P[0 * 3 + j]));
        nvec[j] = fmaf(pf[2], S.P[2 * 3 + j], fmaf(pf[1], S.P[1 * 3 + j], pf[0] * S.P[0 * 3 + j]));
    }
    const float gf0 = g * m[0], gf1 = g * m[1], gf2 = g * m[2];
    const float c = F.c, s = F.s, u = F.u, v = x[1], pd = x[2];
    const float i1 = 1.f / F.d1, i2 = 1.f / F.d2;
    // partial derivatives of f1 = n1 / d1, f2 = n2 / d2
    const float f1_u = (-1.8f * c - 10.9f) * i1, f1_v = (11.5f * c + 68.4f) * i1, f1_w = (-2.4f * pd * s) * i1;
    const float n1p = -s * (-1.8f * u + 11.5f * v + 9.8f * s) + 9.8f * c * c - 1.2f * pd * pd * c;
    const float f1_p = (n1p * F.d1 - F.n1 * (-s)) * i1 * i1;
    const float f2_u = (9.3f * c + 38.6f) * i2, f2_v = (-58.8f * c - 234.5f) * i2, f2_w = (-2.f * pd * s * c) * i2;
    const float n2p = -(9.3f * u - 58.8f * v) * s - c * (208.3f + pd * pd * c) + pd * pd * s * s;
    const float f2_p = (n2p * F.d2 - F.n2 * (-2.f * c * s)) * i2 * i2;
    const float gu = gf1 * f1_u + gf2 * f2_u;
    float gx[3] = {g * nvec[0] + gf1 * f1_p + gf2 * f2_p, g * nvec[1] + gf1 * f1_v + gf2 * f2_v, g * nvec[2] + gf0 + gf1 * f1_w + gf2 * f2_w};
    // controller backward: u = W2 relu(z2) + b2, z2 = W1 a1 + b1, a1 = relu(W0 x + b0)
    float ga1[SH];
#pragma unroll
    for (int k = 0; k < SH; ++k) ga1[k] = 0.f;
#pragma unroll 4
    for (int j = 0; j < SH; ++j) {
        const float gz2 = F.z2[j] > 0.f ? gu * S.W2[j] : 0.f;
        sG2[t][j] = gz2;
#pragma unroll
        for (int k = 0; k < SH; ++k) ga1[k] = fmaf(gz2, S.W1[j * SH + k], ga1[k]);
    }
#pragma unroll
    for (int k = 0; k < SH; ++k) {
        const float gz1 = F.a1[k] > 0.f ? ga1[k] : 0.f;
        sG1[t][k] = gz1;
        sA1[t][k] = F.a1[k];
        gx[0] = fmaf(gz1, S.W0[k * 3 + 0], gx[0]); gx[1] = fmaf(gz1, S.W0[k * 3 + 1], gx[1]); gx[2] = fmaf(gz1, S.W0[k * 3 + 2], gx[2]);
    }
    if (on && grad_x) { grad_x[3 * r] = gx[0]; grad_x[3 * r + 1] = gx[1]; grad_x[3 * r + 2] = gx[2]; }
    if (!part) return;
    sX[t][0] = x[0]; sX[t][1] = x[1]; sX[t][2] = x[2]; sX[t][3] = on ? gu : 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) { sPf[t][i] = on ? g * pf[i] : 0.f; sPf[t][3 + i] = on ? g * px[i] : 0.f; }
    // (points past N carry g = 0, so every staged gradient of theirs is already zero)
    __syncthreads();
    float* dst = part + (size_t)blockIdx.x * kNP;
    // W1 (32 x 32): entry e = j * 32 + k, 8 entries per thread, summed over the block's points in index order
    for (int e = t; e < SH * SH; e += TB) {
        const int j = e >> 5, k = e & 31;
        float acc = 0.f;
        for (int p = 0; p < TB; ++p) acc = fmaf(sG2[p][j], sA1[p][k], acc);
        dst[oW1 + e] = acc;
    }
    for (int e = t; e < SH * 3; e += TB) { // W0 (32 x 3)
        const int j = e / 3, i = e - 3 * j;
        float acc = 0.f;
        for (int p = 0; p < TB; ++p) acc = fmaf(sG1[p][j], sX[p][i], acc);
        dst[oW0 + e] = acc;
    }
    if (t < SH) {                          // b0, b1
        float b0 = 0.f, b1 = 0.f;
        for (int p = 0; p < TB; ++p) { b0 += sG1[p][t]; b1 += sG2[p][t]; }
        dst[oB0 + t] = b0; dst[oB1 + t] = b1;
    }
    __syncthreads();
    // W2 needs relu(z2) and P needs f: staged through sG1 / sPf's spare columns, which are free now
#pragma unroll 4
    for (int j = 0; j < SH; ++j) sG1[t][j] = fmaxf(F.z2[j], 0.f);
    sPf[t][6] = F.f1; sPf[t][7] = F.f2;    // f0 = x2
    __syncthreads();
    if (t < SH) {                          // W2
        float acc = 0.f;
        for (int p = 0; p < TB; ++p) acc = fmaf(sX[p][3], sG1[p][t], acc);
        dst[oW2 + t] = acc;
    } else if (t == SH) {                  // b2
        float acc = 0.f;
        for (int p = 0; p < TB; ++p) acc += sX[p][3];
        dst[oB2] = acc;
    } else if (t >= 40 && t < 40 + 9) {
        // P: the reference's graph reaches P through the inner self.P(f) only -- LyaQuadratic.h_dot refreshes the transposed
        // copy P_t through .data (control_utils.py:230) -- so d / dP = g (P x) f^T; the symmetric half g (P f) x^T is NOT part of
        // what loss.backward() hands the optimizer (train_segway.py:128-131 optimises lya.P only)
        const int i = (t - 40) / 3, j = (t - 40) - 3 * i;
        float acc = 0.f;
        for (int p = 0; p < TB; ++p) {
            const float fj = j == 0 ? sX[p][2] : (j == 1 ? sPf[p][6] : sPf[p][7]);
            acc = fmaf(sPf[p][3 + i], fj, acc);            // g (P x)_i f_j
        }
        dst[oP + (t - 40)] = acc;
    }
}

__global__ void segway_param_reduce_kernel(const float* __restrict__ part, int nblocks, float* __restrict__ out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= kNP) return;
    float acc = 0.f;
    for (int b = 0; b < nblocks; ++b) acc += part[(size_t)b * kNP + e];
    out[e] = acc;
}

// advertorch perturb_iterative, ord = inf:  delta += eps_iter sign(grad); delta = clamp(delta, -eps, eps);
// delta = clamp(x + delta, lo, hi) - x   (in place on x_adv = x + delta)
__global__ void segway_pgd_step_kernel(float* __restrict__ x_adv, const float* __restrict__ x0, const float* __restrict__ grad,
                                       long long n, float eps, float eps_iter, float lo, float hi) {
    for (long long e = blockIdx.x * 256ll + threadIdx.x; e < n; e += (long long)gridDim.x * 256) {
        const float g = grad[e];
        const float sg = g > 0.f ? 1.f : (g < 0.f ? -1.f : 0.f);
        float d = (x_adv[e] - x0[e]) + eps_iter * sg;
        d = fminf(fmaxf(d, -eps), eps);
        x_adv[e] = fminf(fmaxf(x0[e] + d, lo), hi);
    }
}
}  // namespace
}  // namespace fiode

using namespace fiode;

extern "C" int fiode_segway_param_count(void) { return kNP; }
extern "C" size_t fiode_segway_hdot_bwd_workspace_bytes(long long N) {
    return (size_t)((N + TB - 1) / TB) * kNP * sizeof(float) + 256;
}
extern "C" int fiode_segway_hdot(const fiode_segway_weights* w, const float* x, long long N, float* val, void* stream) {
    if (!w || !x || !val) { set_error("fiode_segway_hdot: null pointer"); return -1; }
    if (N <= 0) { set_error("fiode_segway_hdot: N=%lld", N); return -1; }
    const long long want = (N + TB - 1) / TB;
    const int grid = (int)(want < 16ll * sm_count() ? want : 16ll * sm_count());
    segway_hdot_kernel<<<grid, TB, 0, (cudaStream_t)stream>>>(*w, x, N, val);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
extern "C" int fiode_segway_hdot_bwd(const fiode_segway_weights* w, const float* x, long long N, const float* gout, float* grad_x,
                                     float* grad_params, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!w || !x || (!grad_x && !grad_params)) { set_error("fiode_segway_hdot_bwd: null pointer"); return -1; }
    if (N <= 0) { set_error("fiode_segway_hdot_bwd: N=%lld", N); return -1; }
    const long long nb = (N + TB - 1) / TB;
    if (grad_params && (!ws || ws_bytes < fiode_segway_hdot_bwd_workspace_bytes(N))) { set_error("fiode_segway_hdot_bwd: workspace too small"); return -1; }
    segway_hdot_bwd_kernel<<<(unsigned)nb, TB, 0, st>>>(*w, x, N, gout, grad_x, grad_params ? reinterpret_cast<float*>(ws) : nullptr);
    if (grad_params) segway_param_reduce_kernel<<<(kNP + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float*>(ws), (int)nb, grad_params);
    FIODE_CUDA(cudaGetLastError());
    count_launch(grad_params ? 2 : 1);
    return 0;
}
extern "C" int fiode_segway_pgd_step(float* x_adv, const float* x0, const float* grad, long long n, float eps, float eps_iter,
                                     float clip_min, float clip_max, void* stream) {
    if (!x_adv || !x0 || !grad) { set_error("fiode_segway_pgd_step: null pointer"); return -1; }
    if (n <= 0) { set_error("fiode_segway_pgd_step: n=%lld", n); return -1; }
    const long long want = (n + 255) / 256;
    segway_pgd_step_kernel<<<(unsigned)(want < 4096 ? want : 4096), 256, 0, (cudaStream_t)stream>>>(x_adv, x0, grad, n, eps, eps_iter, clip_min, clip_max);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
