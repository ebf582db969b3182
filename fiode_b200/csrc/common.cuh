// fiode_b200/csrc/common.cuh -- shared device helpers for the FI-ODE hot path (sm_100a).
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include "../../include/fiode_b200.h"

namespace fiode {

constexpr int NH = 10;    // ODE state width (classes)            dynamics/classification.py:32
constexpr int NM = 128;   // MLP width                             dynamics/classification.py:35
constexpr int SROW = 12;  // padded row stride (floats) of 10-wide smem rows (16-B aligned)

// thread-local error string for the C-ABI (fiode_last_error)
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
#define FIODE_CUDA(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) return ::fiode::cuda_fail(e__, #x); } while (0)

constexpr int kMaxDevices = 64;
int current_device();
int sm_count();                 // of the current device (cached per device)
bool first_use_on_device(std::atomic<bool>* flags /*[kMaxDevices]*/);
void count_launch(int n);
// A branch beside the caller's stream, for independent kernels of ONE entry point (a layer's weight gradient next to its data
// gradient): fork_side(st) returns an internal stream that has waited for everything queued on st so far (call it again to
// make the branch see later work of st); join_side(st) makes st wait for the branch.  Every fork is joined before the entry
// point returns, so callers see plain stream semantics and a CUDA-graph capture of st stays closed.  Streams and events are
// created once per device on first use (outside capture: run one eager call first); FIODE_FORK=0 returns st itself.
cudaStream_t fork_side(cudaStream_t st);
int join_side(cudaStream_t st);

// ---- Dormand-Prince(5,4) tableau (torchdiffeq 0.2.2 dopri5.py) -------------
static __device__ __constant__ const float kBeta[6][6] = {
    {(float)(1.0 / 5), 0, 0, 0, 0, 0},
    {(float)(3.0 / 40), (float)(9.0 / 40), 0, 0, 0, 0},
    {(float)(44.0 / 45), (float)(-56.0 / 15), (float)(32.0 / 9), 0, 0, 0},
    {(float)(19372.0 / 6561), (float)(-25360.0 / 2187), (float)(64448.0 / 6561), (float)(-212.0 / 729), 0, 0},
    {(float)(9017.0 / 3168), (float)(-355.0 / 33), (float)(46732.0 / 5247), (float)(49.0 / 176),
     (float)(-5103.0 / 18656), 0},
    {(float)(35.0 / 384), 0.f, (float)(500.0 / 1113), (float)(125.0 / 192), (float)(-2187.0 / 6784),
     (float)(11.0 / 84)}};
static __device__ __constant__ const float kCErr[7] = {
    (float)(35.0 / 384 - 1951.0 / 21600), 0.f, (float)(500.0 / 1113 - 22642.0 / 50085),
    (float)(125.0 / 192 - 451.0 / 720), (float)(-2187.0 / 6784 + 12231.0 / 42400),
    (float)(11.0 / 84 - 649.0 / 6300), (float)(-1.0 / 60)};
static __device__ __constant__ const float kCMid[7] = {
    (float)(6025192743.0 / 30085553152.0 / 2), 0.f, (float)(51252292925.0 / 65400821598.0 / 2),
    (float)(-2691868925.0 / 45128329728.0 / 2), (float)(187940372067.0 / 1594534317056.0 / 2),
    (float)(-1776094331.0 / 19743644256.0 / 2), (float)(11237099.0 / 235043384.0 / 2)};

// fp64 step-size controller fed by the fp32 error ratio (torchdiffeq misc._optimal_step_size)
__device__ __forceinline__ double next_dt(double dt, float ratio) {
    if (ratio == 0.f) return dt * 10.0;
    double dfac = (ratio < 1.f) ? 1.0 : 0.2;
    double f = 0.9 / pow((double)ratio, 0.2);
    // torch.min(ifactor, torch.max(f, dfactor)) : NaN propagates like torch.max/min
    if (!(f == f)) return dt * f;
    f = f > dfac ? f : dfac;
    f = f < 10.0 ? f : 10.0;
    return dt * f;
}

// Hairer initial step from the three fp32 norms (torchdiffeq misc._select_initial_step, order arg 4)
__device__ __forceinline__ float init_h0(float d0, float d1) {
    return (d0 < 1e-5f || d1 < 1e-5f) ? 1e-6f : 0.01f * d0 / d1;
}
__device__ __forceinline__ double init_dt(float h0, float d1, float d2) {
    float h1 = (d1 <= 1e-15f && d2 <= 1e-15f) ? fmaxf(1e-6f, h0 * 1e-3f) : powf(0.01f / fmaxf(d1, d2), 0.2f);
    return (double)fminf(100.f * h0, h1);
}

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// order-preserving float <-> uint map so that atomicMax works on floats of either sign
__device__ __forceinline__ unsigned f2ord(float f) {
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// ---- barrier projection, one row per thread (barrier_projection.py:232-255) -------------------
// Runs max_iter bisection midpoints; bit k of the returned word is set when |sum v| >= tol at
// midpoint k (NOT converged), mu_k is written to mus[k*mu_stride].  The caller combines the
// not-converged words of every row of the batch (OR) to find the batch-global stop iteration.
__device__ __forceinline__ unsigned bisect_all(const float (&ft)[NH], const float (&lo)[NH], int max_iter,
                                               float tol, float* mus, int mu_stride) {
    float hi = ft[0] - lo[0], lw = ft[0];
#pragma unroll
    for (int j = 1; j < NH; ++j) { hi = fmaxf(hi, ft[j] - lo[j]); lw = fminf(lw, ft[j]); }
    unsigned notconv = 0;
    for (int k = 0; k < max_iter; ++k) {
        float mu = (hi - lw) / 2.f + lw;
        float e = 0.f;
#pragma unroll
        for (int j = 0; j < NH; ++j) e += fmaxf(ft[j] - mu, lo[j]);
        mus[k * mu_stride] = mu;
        if (!(fabsf(e) < tol)) notconv |= (1u << k);
        if (e > 0.f) lw = mu;
        if (e < 0.f) hi = mu;
    }
    return notconv;
}
// Same bisection for a caller where ONE warp runs it while the rest of the CTA waits (the dopri5 kernel: ncu r01g, 21 % of all
// warp samples sit at the barrier behind it).  The dependent chain per midpoint is cut from ~16 to 8 operations: the ten terms are
// summed as a tree (the reference's torch.sum fixes no order either), and the next midpoint is computed for either sign of e while
// e is still being summed -- (hi - mu) / 2 + mu and (mu - lw) / 2 + lw are exactly what the sequential form evaluates next.
__device__ __forceinline__ unsigned bisect_all_ilp(const float (&ft)[NH], const float (&lo)[NH], int max_iter,
                                                   float tol, float* mus, int mu_stride) {
    static_assert(NH == 10, "tree sum written for 10 terms");
    float hi = ft[0] - lo[0], lw = ft[0];
#pragma unroll
    for (int j = 1; j < NH; ++j) { hi = fmaxf(hi, ft[j] - lo[j]); lw = fminf(lw, ft[j]); }
    unsigned notconv = 0;
    float mu = (hi - lw) / 2.f + lw;
    for (int k = 0; k < max_iter; ++k) {
        float t[NH];
#pragma unroll
        for (int j = 0; j < NH; ++j) t[j] = fmaxf(ft[j] - mu, lo[j]);
        const float mu_pos = (hi - mu) / 2.f + mu, mu_neg = (mu - lw) / 2.f + lw;
        const float e = (((t[0] + t[1]) + (t[2] + t[3])) + ((t[4] + t[5]) + (t[6] + t[7]))) + (t[8] + t[9]);
        mus[k * mu_stride] = mu;
        if (!(fabsf(e) < tol)) notconv |= (1u << k);
        if (e > 0.f) { lw = mu; mu = mu_pos; }
        else if (e < 0.f) { hi = mu; mu = mu_neg; }
    }
    return notconv;
}
// stop iteration from the OR of the not-converged words: first k whose bit is clear, else the last
__device__ __forceinline__ int stop_iter(unsigned notconv_or, int max_iter) {
    unsigned conv = ~notconv_or & (max_iter >= 32 ? 0xffffffffu : ((1u << max_iter) - 1u));
    return conv ? (__ffs(conv) - 1) : (max_iter - 1);
}

// exp rounded correctly to fp32 (via fp64).  `lower` = -alpha_1 (e^{sigma_1 h} - 1) cancels ~9 bits, so a
// 1-ulp difference in exp moves it by ~1e-5: the correctly rounded value is the one every
// libm-quality reference path (torch CPU: SLEEF u10) agrees on.  B200 fp64 rate makes this cheap.
__device__ __forceinline__ float exp_cr(float x) { return (float)exp((double)x); }
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + exp_cr(-x)); }
// dynamics/classification.py:108: lower = -alpha_1 * (exp(sigma_1 * h) - 1)
__device__ __forceinline__ float barrier_lower(float alpha_1, float sigma_1, float h) {
    return -alpha_1 * (exp_cr(sigma_1 * h) - 1.f);
}
// classification.py:109-111: f = (upper - lower) * sigmoid(f) + lower, upper = alpha_2 (1 - h); separate
// roundings like the reference's separate torch ops (no FMA contraction)
__device__ __forceinline__ float scale_nominal_f(float alpha_2, float h, float lo, float f) {
    const float up = alpha_2 * (1.f - h);
    return __fadd_rn(__fmul_rn(up - lo, sigmoidf_(f)), lo);
}
// every rounding spelled out (`up - lo` cannot be contracted into an FMA whatever the calling context): the solver's accept / reject
// sequence on the stress fixtures follows the reference only with the reference's separate roundings
__device__ __forceinline__ float scale_nominal_strict(float alpha_2, float h, float lo, float f) {
    const float up = __fmul_rn(alpha_2, __fsub_rn(1.f, h));
    return __fadd_rn(__fmul_rn(__fsub_rn(up, lo), sigmoidf_(f)), lo);
}


// ---- counter-based dropout masks (Philox4x32-10, Salmon et al. 2011; the generator family torch's CUDA dropout uses) ----
// One call yields the four mask decisions of four consecutive hidden units of one row: element (row, k) of layer l at
// step `step` is lane k & 3 of philox(counter = (row lo, row hi, (k >> 2) | l << 8, step lo), key = seed ^ step hi).
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const unsigned hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}
struct DropCfg {            // in-kernel dropout of the dynamics MLP (classification.py:96-102); thresh == 0 -> off
    unsigned thresh;        // drop iff u32 < thresh  (thresh = p * 2^32)
    float scale;            // 1 / (1 - p)
    unsigned long long seed, step;
};
__device__ __forceinline__ float4 dropout_mask4(const DropCfg& d, long long row, int k4, int layer) {
    const uint4 r = philox4x32_10(make_uint4((unsigned)row, (unsigned)((unsigned long long)row >> 32),
                                             (unsigned)k4 | ((unsigned)layer << 8), (unsigned)d.step),
                                  make_uint2((unsigned)d.seed ^ (unsigned)(d.step >> 32), (unsigned)(d.seed >> 32)));
    return make_float4(r.x >= d.thresh ? d.scale : 0.f, r.y >= d.thresh ? d.scale : 0.f,
                       r.z >= d.thresh ? d.scale : 0.f, r.w >= d.thresh ? d.scale : 0.f);
}
}  // namespace fiode
