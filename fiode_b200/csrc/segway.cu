// fiode_b200/csrc/segway.cu -- Segway NN-controller forward-invariance grid check and closed-loop
// zero-order-hold simulation.
//
// Replaces (control/certify_segway.py:55-59 and :104-110):
//   * control/grid_sampler.py:109-117 grid_uniform_3d (np.meshgrid 'xy' order) and :91-98
//     reject_sampling (level-set shell lb <= V <= ub);
//   * control/control_utils.py:215-234 LyaQuadratic (V = ||P x||^2, h_dot), :21-28 AutoLirpaModel,
//     :97-106 SegwaySingleBarrierModel (h = sum(P^T (P f) * x));
//   * libs/core/core/controllers/nn_controller.py:6-34 NNController 3 -> 32 -> 32 -> 1 (ReLU);
//   * libs/core/core/systems/segway_system.py:19-29 Segway.forward;
//   * libs/core/core/dynamics/system_dynamics.py:30-91 step / simulate: one torchdiffeq dopri5 solve
//     (rtol = atol = 1e-6, batch-wide RMS norm over the B trajectories) per sample interval.
// Nothing is read from HBM but the 1.2 K weights: grid points are generated from their index.
#include "common.cuh"

namespace fiode {

constexpr int SH = 32;   // hidden width of the controller

struct SegW {            // controller + P, staged in shared memory
    float W0[SH * 3], b0[SH], W1[SH * SH], b1[SH], W2[SH], b2, P[9];
};

__device__ __forceinline__ void load_segw(const fiode_segway_weights& w, SegW* S) {
    for (int e = threadIdx.x; e < SH * 3; e += blockDim.x) S->W0[e] = w.W0[e];
    for (int e = threadIdx.x; e < SH * SH; e += blockDim.x) S->W1[e] = w.W1[e];
    for (int e = threadIdx.x; e < SH; e += blockDim.x) { S->b0[e] = w.b0[e]; S->b1[e] = w.b1[e]; S->W2[e] = w.W2[e]; }
    if (threadIdx.x < 9) S->P[threadIdx.x] = w.P[threadIdx.x];
    if (threadIdx.x == 0) S->b2 = w.b2[0];
    __syncthreads();
}

// u = MLP(x)   (nn.Linear: y = x W^T + b)
__device__ __forceinline__ float controller(const SegW* S, float x0, float x1, float x2) {
    float z0[SH];
#pragma unroll
    for (int j = 0; j < SH; ++j) {
        float a = S->b0[j];
        a = fmaf(x0, S->W0[j * 3 + 0], a); a = fmaf(x1, S->W0[j * 3 + 1], a); a = fmaf(x2, S->W0[j * 3 + 2], a);
        z0[j] = fmaxf(a, 0.f);
    }
    float u = S->b2;
#pragma unroll 4
    for (int j = 0; j < SH; ++j) {
        float a = S->b1[j];
#pragma unroll
        for (int k = 0; k < SH; ++k) a = fmaf(z0[k], S->W1[j * SH + k], a);
        u = fmaf(fmaxf(a, 0.f), S->W2[j], u);
    }
    return u;
}

// segway_system.py:19-29 with correctly rounded sin / cos (via fp64), same operation order
__device__ __forceinline__ void segway_f(float phi, float v, float pd, float u, float& f0, float& f1, float& f2) {
    const float c = (float)cos((double)phi), s = (float)sin((double)phi);
    f0 = pd;
    const float t1 = __fmul_rn(c, __fadd_rn(__fadd_rn(__fmul_rn(-1.8f, u), __fmul_rn(11.5f, v)), __fmul_rn(9.8f, s)));
    const float num1 = __fsub_rn(__fadd_rn(__fsub_rn(t1, __fmul_rn(10.9f, u)), __fmul_rn(68.4f, v)),
                                 __fmul_rn(__fmul_rn(__fmul_rn(1.2f, pd), pd), s));
    f1 = num1 / __fsub_rn(c, 24.7f);
    const float a = __fmul_rn(__fsub_rn(__fmul_rn(9.3f, u), __fmul_rn(58.8f, v)), c);
    const float b = __fsub_rn(__fadd_rn(a, __fmul_rn(38.6f, u)), __fmul_rn(234.5f, v));
    const float d = __fmul_rn(s, __fadd_rn(208.3f, __fmul_rn(__fmul_rn(pd, pd), c)));
    f2 = __fsub_rn(b, d) / __fsub_rn(__fmul_rn(c, c), 24.7f);
}

__device__ __forceinline__ float lya_V(const SegW* S, float x0, float x1, float x2) {
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float p = fmaf(x2, S->P[i * 3 + 2], fmaf(x1, S->P[i * 3 + 1], x0 * S->P[i * 3 + 0]));
        acc = __fadd_rn(acc, __fmul_rn(p, p));
    }
    return acc;
}
// h = sum_i (P^T (P f))_i x_i      (control_utils.py:105, :229-234)
__device__ __forceinline__ float barrier_h(const SegW* S, float x0, float x1, float x2, float f0, float f1, float f2) {
    float pf[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) pf[i] = fmaf(f2, S->P[i * 3 + 2], fmaf(f1, S->P[i * 3 + 1], f0 * S->P[i * 3 + 0]));
    const float xs[3] = {x0, x1, x2};
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const float q = fmaf(pf[2], S->P[2 * 3 + j], fmaf(pf[1], S->P[1 * 3 + j], pf[0] * S->P[0 * 3 + j]));
        acc = __fadd_rn(acc, __fmul_rn(q, xs[j]));
    }
    return acc;
}

struct GridArgs {
    fiode_segway_weights w;
    const float* d0; const float* d1; const float* d2; int n0, n1, n2;
    float lb, ub;
    long long first, count;
    unsigned* block_counts;       // pass 0 out / pass 1 in (exclusive offsets)
    float* eta; float* val;       // pass 1 out, compacted (M,3), (M)
    unsigned* max_ord;            // running max of val (ordered uint)
};

// index -> point, np.meshgrid(d0, d1, d2) default 'xy' indexing then reshape(-1): idx = (j * n0 + i) * n2 + k
__device__ __forceinline__ void grid_point(const GridArgs& a, long long idx, float& x0, float& x1, float& x2) {
    const int k = (int)(idx % a.n2);
    const long long r = idx / a.n2;
    const int i = (int)(r % a.n0), j = (int)(r / a.n0);
    x0 = a.d0[i]; x1 = a.d1[j]; x2 = a.d2[k];
}

constexpr int GB = 256;           // points per block per pass-iteration
template <int PASS>
__global__ void __launch_bounds__(GB) segway_grid_kernel(const GridArgs a) {
    __shared__ SegW S;
    __shared__ unsigned wsum[GB / 32];
    __shared__ unsigned base_sh;
    load_segw(a.w, &S);
    const long long chunk = (a.count + gridDim.x - 1) / gridDim.x;      // contiguous slab per block: ordered output
    const long long lo = blockIdx.x * chunk, hi = min(a.count, lo + chunk);
    unsigned running = PASS ? a.block_counts[blockIdx.x] : 0u;
    float vmax = -INFINITY;
    for (long long p0 = lo; p0 < hi; p0 += GB) {
        const long long p = p0 + threadIdx.x;
        bool in = false; float x0 = 0, x1 = 0, x2 = 0;
        if (p < hi) {
            grid_point(a, a.first + p, x0, x1, x2);
            const float V = lya_V(&S, x0, x1, x2);
            in = (V >= a.lb) && (V <= a.ub);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, in);
        const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
        if (lane == 0) wsum[wp] = __popc(bal);
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned acc = running;
            for (int i = 0; i < GB / 32; ++i) { const unsigned c = wsum[i]; wsum[i] = acc; acc += c; }
            base_sh = acc;
        }
        __syncthreads();
        if (PASS && in) {
            const unsigned pos = wsum[wp] + __popc(bal & ((1u << lane) - 1u));
            float f0, f1, f2;
            const float u = controller(&S, x0, x1, x2);
            segway_f(x0, x1, x2, u, f0, f1, f2);
            const float hv = barrier_h(&S, x0, x1, x2, f0, f1, f2);
            a.eta[3ull * pos] = x0; a.eta[3ull * pos + 1] = x1; a.eta[3ull * pos + 2] = x2;
            a.val[pos] = hv;
            vmax = fmaxf(vmax, hv);
        }
        running = base_sh;
        __syncthreads();
    }
    if (!PASS) { if (threadIdx.x == 0) a.block_counts[blockIdx.x] = running; }
    else {
        vmax = warp_max(vmax);
        if ((threadIdx.x & 31) == 0 && vmax > -INFINITY) atomicMax(a.max_ord, f2ord(vmax));
    }
}
// exclusive scan of the per-block counts (single thread: a few hundred entries); total -> counts[nb]
__global__ void segway_scan_kernel(unsigned* counts, int nb) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        unsigned acc = 0;
        for (int i = 0; i < nb; ++i) { const unsigned c = counts[i]; counts[i] = acc; acc += c; }
        counts[nb] = acc;
    }
}
__global__ void segway_finish_kernel(const unsigned* counts, int nb, const unsigned* max_ord, float* stats) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        stats[0] = (float)counts[nb];
        stats[1] = counts[nb] ? ord2f(*max_ord) : -INFINITY;
    }
}

// ---- closed-loop ZOH simulation: one warp, lane b = trajectory b ------------------------------------
struct SimArgs {
    fiode_segway_weights w;
    const float* x0; int B; const float* ts; int N; float rtol, atol;
    float* xs; float* us; int* stats;     // (B,N,3), (B,N-1,1), {status, total accepted, total rejected, nfe}
};

__device__ __forceinline__ float rms3(float a, float b, float c, bool on, int n) {
    double s = on ? ((double)a * a + (double)b * b + (double)c * c) : 0.0;
    s = warp_sum(s);
    return sqrtf((float)(s / (double)n));
}

__global__ void __launch_bounds__(32) segway_simulate_kernel(const SimArgs a) {
    __shared__ SegW S;
    load_segw(a.w, &S);
    const int b = threadIdx.x;
    const bool on = b < a.B;
    const int n = a.B * 3;
    float y[3] = {0, 0, 0};
    if (on) { y[0] = a.x0[b * 3]; y[1] = a.x0[b * 3 + 1]; y[2] = a.x0[b * 3 + 2]; }
    if (on) { float* o = a.xs + ((size_t)b * a.N) * 3; o[0] = y[0]; o[1] = y[1]; o[2] = y[2]; }
    int status = 0, nacc = 0, nrej = 0, nfe = 0;
    for (int j = 0; j + 1 < a.N && status == 0; ++j) {
        const double t0 = (double)a.ts[j], tf = (double)a.ts[j + 1];
        const float u = on ? controller(&S, y[0], y[1], y[2]) : 0.f;          // held over the interval
        if (on) a.us[(size_t)b * (a.N - 1) + j] = u;
        float k[7][3];
        segway_f(y[0], y[1], y[2], u, k[0][0], k[0][1], k[0][2]); ++nfe;
        // initial step (misc._select_initial_step)
        float sc[3], q0[3], q1[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) { sc[i] = a.atol + fabsf(y[i]) * a.rtol; q0[i] = y[i] / sc[i]; q1[i] = k[0][i] / sc[i]; }
        const float d0 = rms3(q0[0], q0[1], q0[2], on, n), d1 = rms3(q1[0], q1[1], q1[2], on, n);
        const float h0 = init_h0(d0, d1);
        float g[3];
        segway_f(y[0] + h0 * k[0][0], y[1] + h0 * k[0][1], y[2] + h0 * k[0][2], u, g[0], g[1], g[2]); ++nfe;
        const float d2 = rms3((g[0] - k[0][0]) / sc[0], (g[1] - k[0][1]) / sc[1], (g[2] - k[0][2]) / sc[2], on, n) / h0;
        double dt = init_dt(h0, d1, d2);
        double tcur = t0, tleft = t0;
        float ce[3] = {y[0], y[1], y[2]}, cd[3] = {y[0], y[1], y[2]}, cc[3] = {y[0], y[1], y[2]},
              cb[3] = {y[0], y[1], y[2]}, ca[3] = {y[0], y[1], y[2]};
        long long nsteps = 0;
        while (tf > tcur) {
            if (!(tcur + dt > tcur)) { status = 1; break; }
            const unsigned bad = __ballot_sync(0xffffffffu, on && !(isfinite(y[0]) && isfinite(y[1]) && isfinite(y[2])));
            if (bad) { status = 2; break; }
            const float dtf = (float)dt;
            float ys[3];
#pragma unroll
            for (int s = 0; s < 6; ++s) {
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    float acc = k[0][i] * (kBeta[s][0] * dtf);
#pragma unroll
                    for (int q = 1; q <= s; ++q) acc = fmaf(k[q][i], kBeta[s][q] * dtf, acc);
                    ys[i] = y[i] + acc;
                }
                segway_f(ys[0], ys[1], ys[2], u, k[s + 1][0], k[s + 1][1], k[s + 1][2]); ++nfe;
            }
            float qe[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                float err = k[0][i] * (dtf * kCErr[0]);
#pragma unroll
                for (int q = 1; q < 7; ++q) err = fmaf(k[q][i], dtf * kCErr[q], err);
                qe[i] = err / (a.atol + a.rtol * fmaxf(fabsf(y[i]), fabsf(ys[i])));
            }
            const float ratio = fabsf(rms3(qe[0], qe[1], qe[2], on, n));
            if (ratio <= 1.f) {
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    float mid = k[0][i] * (dtf * kCMid[0]);
#pragma unroll
                    for (int q = 1; q < 7; ++q) mid = fmaf(k[q][i], dtf * kCMid[q], mid);
                    const float y0 = y[i], y1 = ys[i], f0 = k[0][i], f1 = k[6][i], ym = y0 + mid;
                    ca[i] = 2.f * dtf * (f1 - f0) - 8.f * (y1 + y0) + 16.f * ym;
                    cb[i] = dtf * (5.f * f0 - 3.f * f1) + 18.f * y0 + 14.f * y1 - 32.f * ym;
                    cc[i] = dtf * (f1 - 4.f * f0) - 11.f * y0 - 5.f * y1 + 16.f * ym;
                    cd[i] = dtf * f0;
                    ce[i] = y0;
                    y[i] = y1;
                    k[0][i] = f1;
                }
                tleft = tcur; tcur = tcur + dt; ++nacc;
            } else ++nrej;
            dt = next_dt(dt, ratio);
            if (++nsteps > 1000000) { status = 3; break; }
        }
        if (status) break;
        const float x = (float)((tf - tleft) / (tcur - tleft));
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            float tot = ce[i] + x * cd[i];
            float xp = x * x; tot = tot + xp * cc[i];
            xp = xp * x; tot = tot + xp * cb[i];
            xp = xp * x; tot = tot + xp * ca[i];
            y[i] = tot;                                   // step() returns the interpolated state; next interval restarts
        }
        if (on) { float* o = a.xs + ((size_t)b * a.N + j + 1) * 3; o[0] = y[0]; o[1] = y[1]; o[2] = y[2]; }
    }
    if (b == 0) { a.stats[0] = status; a.stats[1] = nacc; a.stats[2] = nrej; a.stats[3] = nfe; }
}

}  // namespace fiode

using namespace fiode;

extern "C" size_t fiode_segway_grid_workspace_bytes(void) { return (size_t)(4 * 148 * 8 + 8) * sizeof(unsigned) + 256; }

extern "C" int fiode_segway_grid(const fiode_segway_weights* w, const float* d0, int n0, const float* d1, int n1,
                                 const float* d2, int n2, float level_lb, float level_ub, long long first,
                                 long long count, long long capacity, float* eta, float* val, float* stats2, void* ws,
                                 size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!w || !d0 || !d1 || !d2 || !eta || !val || !stats2 || !ws) { set_error("fiode_segway_grid: null pointer"); return -1; }
    const long long total = (long long)n0 * n1 * n2;
    if (n0 <= 0 || n1 <= 0 || n2 <= 0 || first < 0 || count <= 0 || first + count > total) { set_error("fiode_segway_grid: bad range"); return -1; }
    if (ws_bytes < fiode_segway_grid_workspace_bytes()) { set_error("fiode_segway_grid: workspace too small"); return -1; }
    (void)capacity;   // the caller sizes eta/val for `count` points (worst case) or for a bound it has proven
    const long long want = (count + GB - 1) / GB;
    int nb = (int)(want < 4ll * sm_count() ? want : 4ll * sm_count());
    if (nb > 4 * 148 * 8) nb = 4 * 148 * 8;
    unsigned* counts = reinterpret_cast<unsigned*>(ws);
    unsigned* max_ord = counts + nb + 1;
    FIODE_CUDA(cudaMemsetAsync(ws, 0, (size_t)(nb + 2) * sizeof(unsigned), st));
    GridArgs a{};
    a.w = *w; a.d0 = d0; a.d1 = d1; a.d2 = d2; a.n0 = n0; a.n1 = n1; a.n2 = n2; a.lb = level_lb; a.ub = level_ub;
    a.first = first; a.count = count; a.block_counts = counts; a.eta = eta; a.val = val; a.max_ord = max_ord;
    segway_grid_kernel<0><<<nb, GB, 0, st>>>(a);
    segway_scan_kernel<<<1, 32, 0, st>>>(counts, nb);
    segway_grid_kernel<1><<<nb, GB, 0, st>>>(a);
    segway_finish_kernel<<<1, 32, 0, st>>>(counts, nb, max_ord, stats2);
    FIODE_CUDA(cudaGetLastError());
    count_launch(4);
    return 0;
}

extern "C" int fiode_segway_simulate(const fiode_segway_weights* w, const float* x0, int B, const float* ts, int N,
                                     double rtol, double atol, float* xs, float* us, int* stats4, void* stream) {
    if (!w || !x0 || !ts || !xs || !us || !stats4) { set_error("fiode_segway_simulate: null pointer"); return -1; }
    if (B < 1 || B > 32 || N < 2) { set_error("fiode_segway_simulate: need 1 <= B <= 32 trajectories and N >= 2 (got B=%d N=%d)", B, N); return -1; }
    SimArgs a{};
    a.w = *w; a.x0 = x0; a.B = B; a.ts = ts; a.N = N; a.rtol = (float)rtol; a.atol = (float)atol; a.xs = xs; a.us = us;
    a.stats = stats4;
    segway_simulate_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
