// fiode_b200/csrc/dopri5.cu -- device-resident adaptive Dormand-Prince solver for the FI-ODE
// simplex dynamics.  One cooperative persistent kernel per solve; no host sync per step.
//
// Replaces torchdiffeq==0.2.2 odeint(..., method='dopri5') as called at the reference's
// models.py:236-241 (func = dyn_fun.ode_forward, dynamics/classification.py:104-132).
// Semantics follow SURVEY.md Appendix A: fp32 state/stages, fp64 time + controller, batch-global
// RMS error norm, accept iff ratio <= 1, dense output inside the last accepted step; the barrier
// projection keeps the reference's batch-global stop iteration (barrier_projection.py:247).
//
// Work split: the batch is cut into contiguous slabs of R rows, one slab per CTA for the whole
// solve.  All row state (y, 7 stages, interpolant, MLP activations) lives in that CTA's shared
// memory; the 128x128 mixing matrix lives in registers (64 per thread).  The only global traffic
// per RHS evaluation is one grid barrier carrying a 32-bit OR word (projection stop iteration);
// per step one more barrier carries the error-norm partial sums (ordered, deterministic).
#include "common.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace fiode {

constexpr int kThreads = 256;
constexpr int kMaxGrid = 1024;
constexpr int kMaxIter = 32;

struct SolveCtl {                 // lives at the start of the caller's workspace (zeroed per solve)
    unsigned count;               // monotonic arrival counter
    unsigned orw[3];              // rotating OR words (projection not-converged bits)
    double slots[2][2 * kMaxGrid];  // per-CTA partial sums, double-buffered by barrier parity
};

struct SolveArgs {
    fiode_dyn_weights w;
    const float* h0; const float* s; int s_rows; int B; int R;
    const double* t; int nt; float rtol, atol; long long max_steps;
    float* out; fiode_solve_stats* stats; SolveCtl* ctl;
    int G, C;                     // cluster mode: G independent solves of B rows each, C CTAs (one cluster) per solve
    double* xlog_dt; float* xlog_ratio; long long xlog_cap;   // optional caller-owned attempt log (per solve: xlog_cap entries)
};

struct ClusterSlot { double s0, s1; unsigned orw; unsigned pad; };

// who am I inside my solve, and how do its CTAs talk
struct Comm {
    int rank, size;               // CTA index / CTA count of this solve
    unsigned epoch;
    SolveCtl* ctl;                // cooperative-grid mode: global-memory barrier
    ClusterSlot* slots;           // cluster mode: [2] in this CTA's shared memory, read by peers through DSMEM
};

struct Red { double s0, s1; unsigned orw; };

// One solve-wide barrier that also (a) ORs a 32-bit word and (b) sums two doubles in CTA order.
// Caller passes the CTA-level values in thread 0.
//   cooperative-grid mode: arrival counter + rotating OR words + per-CTA slots in global memory;
//   cluster mode: each CTA publishes into its own shared memory, one hardware cluster barrier, every
//   CTA then reads its peers' slots over DSMEM in rank order (deterministic).  Slots are double-buffered
//   by barrier parity, so one barrier per reduction suffices.
template <bool CLUSTER>
__device__ __forceinline__ Red comm_reduce(Comm& c, double s0, double s1, unsigned orw, bool want_sums, Red* sh) {
    const unsigned e = c.epoch;
    if (CLUSTER) {
        cg::cluster_group cl = cg::this_cluster();
        ClusterSlot* mine = c.slots + (e & 1);
        if (threadIdx.x == 0) { mine->s0 = s0; mine->s1 = s1; mine->orw = orw; }
        cl.sync();
        if (threadIdx.x == 0) {
            double a0 = 0, a1 = 0; unsigned o = 0;
            for (int r = 0; r < c.size; ++r) {
                const ClusterSlot* peer = cl.map_shared_rank(c.slots, r) + (e & 1);
                a0 += peer->s0; a1 += peer->s1; o |= peer->orw;
            }
            sh->s0 = a0; sh->s1 = a1; sh->orw = o;
        }
        __syncthreads();
        c.epoch = e + 1;
        return *sh;
    }
    SolveCtl* ctl = c.ctl;
    const unsigned G = c.size;
    if (threadIdx.x == 0) {
        if (want_sums) {
            double* sl = ctl->slots[e & 1];
            __stcg(sl + 2 * c.rank, s0);
            __stcg(sl + 2 * c.rank + 1, s1);
        }
        if (orw) atomicOr(&ctl->orw[e % 3], orw);
        __threadfence();
        atomicAdd(&ctl->count, 1u);
        const unsigned target = (e + 1u) * G;
        while (ld_acquire_u32(&ctl->count) < target) {}
        if (c.rank == 0) __stcg(&ctl->orw[(e + 2) % 3], 0u);
        sh->orw = __ldcg(&ctl->orw[e % 3]);
    }
    __syncthreads();
    if (want_sums && threadIdx.x < 32) {
        const double* sl = ctl->slots[e & 1];
        double a0 = 0, a1 = 0;
        for (unsigned i = threadIdx.x; i < G; i += 32) { a0 += __ldcg(sl + 2 * i); a1 += __ldcg(sl + 2 * i + 1); }
        a0 = warp_sum(a0); a1 = warp_sum(a1);
        if (threadIdx.x == 0) { sh->s0 = a0; sh->s1 = a1; }
    }
    if (want_sums) __syncthreads();
    c.epoch = e + 1;
    return *sh;
}

// CTA-level sum of a per-thread double pair -> valid in thread 0
__device__ __forceinline__ void cta_sum2(double& a, double& b, double* scratch /*[16]*/) {
    a = warp_sum(a); b = warp_sum(b);
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { scratch[w] = a; scratch[8 + w] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double x = 0, y = 0;
#pragma unroll
        for (int i = 0; i < kThreads / 32; ++i) { x += scratch[i]; y += scratch[8 + i]; }
        a = x; b = y;
    }
    __syncthreads();
}

struct Smem {
    float *c, *z1, *z2;          // [Rp][128]
    float *y, *ys, *ft, *lo;     // [Rp][SROW]
    float *k;                    // [7][Rp][SROW]
    float *ce, *cd, *cc, *cb, *ca;  // interpolant coefficients [Rp][SROW]
    float *mus;                  // [kMaxIter][Rp]
    double* red;                 // [16]
    Red* rr;
    unsigned* orw;
};

// One evaluation of f(h) for the CTA's slab: ys -> kout (both [Rp][SROW] in smem).
// Follows classification.py:96-115 + barrier_projection.py:232-255.
template <bool CLUSTER>
__device__ __forceinline__ void eval_rhs(const SolveArgs& a, const Smem& S, int R, int Rp, const float (&q1)[NH],
                                         const float4 (&q2)[16], const float (&q3)[16], float b2j, float b2j1, float b3o,
                                         float* kout, Comm& comm) {
    const int t = threadIdx.x;
    // layer 1: z1[r][j] = relu(c[r][j] + Q1[j,:] . h[r,:])
    {
        const int j = t & 127;
        // four rows per pass: the ten dependent FMAs of a row are the latency, independent rows hide it (same order per row)
        for (int r0 = t >> 7; r0 < R; r0 += 8) {
            float z[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int r = min(r0 + 2 * u, Rp - 1);
                const float4 h0 = *reinterpret_cast<const float4*>(S.ys + r * SROW);
                const float4 h1 = *reinterpret_cast<const float4*>(S.ys + r * SROW + 4);
                const float2 h2 = *reinterpret_cast<const float2*>(S.ys + r * SROW + 8);
                float zz = S.c[r * NM + j];
                zz = fmaf(q1[0], h0.x, zz); zz = fmaf(q1[1], h0.y, zz); zz = fmaf(q1[2], h0.z, zz); zz = fmaf(q1[3], h0.w, zz);
                zz = fmaf(q1[4], h1.x, zz); zz = fmaf(q1[5], h1.y, zz); zz = fmaf(q1[6], h1.z, zz); zz = fmaf(q1[7], h1.w, zz);
                zz = fmaf(q1[8], h2.x, zz); zz = fmaf(q1[9], h2.y, zz);
                z[u] = zz;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (r0 + 2 * u < R) S.z1[(r0 + 2 * u) * NM + j] = fmaxf(z[u], 0.f);
        }
    }
    __syncthreads();
    // layer 2: thread (jg, ks) owns Q2[4 jg .. 4 jg + 3][k], k = 32 m + 4 ks + c (m, c < 4), in registers: every LDS.128 of
    // z1 feeds 16 FFMAs (4 outputs x 4 k) and the 8 ks-lanes of a row read 128 contiguous bytes (no bank conflicts);
    // 4 rows per pass -> 16 independent accumulators, reduced over the 8 ks-lanes by a 3-step reduce-scatter
    {
        const int ks = t & 7, jg = t >> 3;
        for (int r = 0; r < Rp; r += 4) {
            float acc[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[i] = 0.f;
#pragma unroll
            for (int m = 0; m < 4; ++m) {
#pragma unroll
                for (int rr = 0; rr < 4; ++rr) {
                    const float4 v = *reinterpret_cast<const float4*>(S.z1 + (r + rr) * NM + m * 32 + ks * 4);
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        const float4 q = q2[jj * 4 + m];
                        float a = acc[rr * 4 + jj];
                        a = fmaf(q.x, v.x, a); a = fmaf(q.y, v.y, a); a = fmaf(q.z, v.z, a); a = fmaf(q.w, v.w, a);
                        acc[rr * 4 + jj] = a;
                    }
                }
            }
            float w8[8], w4[4], w2[2];
            const bool h4 = ks & 4, h2 = ks & 2, h1 = ks & 1;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float keep = h4 ? acc[i + 8] : acc[i], send = h4 ? acc[i] : acc[i + 8];
                w8[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float keep = h2 ? w8[i + 4] : w8[i], send = h2 ? w8[i] : w8[i + 4];
                w4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
            }
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const float keep = h1 ? w4[i + 2] : w4[i], send = h1 ? w4[i] : w4[i + 2];
                w2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
            }
            // this lane now holds accumulator indices idx0, idx0 + 1 (idx = 4 row + output)
            const int idx0 = (h4 ? 8 : 0) + (h2 ? 4 : 0) + (h1 ? 2 : 0);
            const int rr = idx0 >> 2, jj = idx0 & 3;
            *reinterpret_cast<float2*>(S.z2 + (r + rr) * NM + jg * 4 + jj) =
                make_float2(fmaxf(w2[0] + b2j, 0.f), fmaxf(w2[1] + b2j1, 0.f));
        }
    }
    __syncthreads();
    // layer 3: 8 threads per output (k = 32 kk + 4 ks + c, conflict-free like layer 2), 3 rows x 10 outputs per pass
    // (threads 240..255 shadow group 0)
    {
        const int g = (t >> 3) % 30, ks = t & 7, o = g % NH, rr = g / NH;
        // four passes of 3 rows in flight per thread: independent 16-FMA + 3-shuffle chains, same order per output as before
        for (int r0 = 0; r0 < R; r0 += 12) {
            float acc[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int r = min(r0 + 3 * u + rr, Rp - 1);
                const float* z = S.z2 + r * NM + ks * 4;
                float a_ = 0.f;
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const float4 v = *reinterpret_cast<const float4*>(z + kk * 32);
                    a_ = fmaf(q3[4 * kk + 0], v.x, a_); a_ = fmaf(q3[4 * kk + 1], v.y, a_);
                    a_ = fmaf(q3[4 * kk + 2], v.z, a_); a_ = fmaf(q3[4 * kk + 3], v.w, a_);
                }
                acc[u] = a_;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], 1);
                acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], 2);
                acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], 4);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ks == 0 && t < 240 && r0 + 3 * u + rr < R) S.ft[(r0 + 3 * u + rr) * SROW + o] = acc[u] + b3o;
        }
    }
    __syncthreads();
    // barrier projection.  The element-wise part first, spread over the whole CTA (one correctly rounded exp through fp64 per
    // element: ten per row on the one or two row warps was a quarter of pass A, with every other warp waiting at the barrier)
    for (int e = t; e < R * NH; e += kThreads) {
        const int r = e / NH, i = e - r * NH;
        const float h = S.ys[r * SROW + i];
        const float lo = barrier_lower(a.w.alpha_1, a.w.sigma_1, h);
        if (a.w.scale_nominal) S.ft[r * SROW + i] = scale_nominal_strict(a.w.alpha_2, h, lo, S.ft[r * SROW + i]);
        S.lo[r * SROW + i] = lo;
    }
    __syncthreads();
    // pass A: one row per thread, all max_iter midpoints, record mu_k + mask
    unsigned nc = 0;
    if (t < R) {
        float ft[NH], lo[NH];
#pragma unroll
        for (int i = 0; i < NH; ++i) { ft[i] = S.ft[t * SROW + i]; lo[i] = S.lo[t * SROW + i]; }
        nc = bisect_all_ilp(ft, lo, a.w.max_iter, a.w.tol, S.mus + t, Rp);
    }
    if (t < 64) {  // R <= 64 -> rows live in warps 0 and 1
        nc = __reduce_or_sync(0xffffffffu, nc);
        if ((t & 31) == 0 && nc) atomicOr(S.orw, nc);
    }
    __syncthreads();
    unsigned cta_or = 0;
    if (t == 0) { cta_or = *S.orw; *S.orw = 0; }
    const Red red = comm_reduce<CLUSTER>(comm, 0.0, 0.0, cta_or, false, S.rr);
    const int K = stop_iter(red.orw, a.w.max_iter);
    // pass B: v = max(ftilde - mu_K, lower)
    for (int e = t; e < R * NH; e += kThreads) {
        const int r = e / NH, i = e - r * NH;
        kout[r * SROW + i] = fmaxf(S.ft[r * SROW + i] - S.mus[K * Rp + r], S.lo[r * SROW + i]);
    }
    __syncthreads();
}

template <bool CLUSTER>
__global__ void __launch_bounds__(kThreads, 1) dopri5_simplex_kernel(const SolveArgs a) {
    extern __shared__ float4 smem4[];
    const int t = threadIdx.x;
    Comm comm;
    comm.rank = CLUSTER ? (int)(blockIdx.x % a.C) : (int)blockIdx.x;
    comm.size = CLUSTER ? a.C : (int)gridDim.x;
    comm.epoch = 0; comm.ctl = a.ctl;
    const int solve = CLUSTER ? (int)(blockIdx.x / a.C) : 0;
    const int lrow0 = comm.rank * a.R;                       // first row inside the solve
    const int R = max(0, min(a.R, a.B - lrow0));
    const int row0 = solve * a.B + lrow0;                    // first row in the global (G*B) tensors
    const int Rp = (a.R + 3) & ~3;
    fiode_solve_stats* stats = a.stats + solve;

    Smem S;
    float* p = reinterpret_cast<float*>(smem4);
    S.c = p; p += Rp * NM; S.z1 = p; p += Rp * NM; S.z2 = p; p += Rp * NM;
    S.y = p; p += Rp * SROW; S.ys = p; p += Rp * SROW; S.ft = p; p += Rp * SROW; S.lo = p; p += Rp * SROW;
    S.k = p; p += 7 * Rp * SROW;
    S.ce = p; p += Rp * SROW; S.cd = p; p += Rp * SROW; S.cc = p; p += Rp * SROW; S.cb = p; p += Rp * SROW;
    S.ca = p; p += Rp * SROW;
    S.mus = p; p += kMaxIter * Rp;
    S.red = reinterpret_cast<double*>(p); p += 32;
    S.rr = reinterpret_cast<Red*>(p); p += 8;
    S.orw = reinterpret_cast<unsigned*>(p); p += 4;
    comm.slots = reinterpret_cast<ClusterSlot*>(p); p += 2 * sizeof(ClusterSlot) / sizeof(float);

    // ---- one-time setup: weights to registers, c = b1 + bx + Qx s, y = h0 -----------------------
    float q1[NH]; float4 q2[16]; float q3[16]; float b2j, b2j1, b3o;
    {
        const int j = t & 127;
#pragma unroll
        for (int i = 0; i < NH; ++i) q1[i] = a.w.Q1[j * NH + i];
        {   // layer-2 slice: q2[jj * 4 + m] = Q2[4 jg + jj][32 m + 4 ks .. + 4]
            const int ks2 = t & 7, jg = t >> 3;
#pragma unroll
            for (int jj = 0; jj < 4; ++jj)
#pragma unroll
                for (int m = 0; m < 4; ++m)
                    q2[jj * 4 + m] = *reinterpret_cast<const float4*>(a.w.Q2 + (jg * 4 + jj) * NM + m * 32 + ks2 * 4);
            const int j0 = jg * 4 + 2 * (ks2 & 1);          // the two outputs this lane stores after the reduce-scatter
            b2j = a.w.b2[j0]; b2j1 = a.w.b2[j0 + 1];
        }
        const int g = (t >> 3) % 30, ks = t & 7, o = g % NH;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
#pragma unroll
            for (int c = 0; c < 4; ++c) q3[4 * kk + c] = a.w.Q3[o * NM + kk * 32 + ks * 4 + c];
        b3o = a.w.b3[o];
        for (int e = t; e < Rp * NM; e += kThreads) { S.z1[e] = 0.f; S.z2[e] = 0.f; }
        float qx[NH];
#pragma unroll
        for (int i = 0; i < NH; ++i) qx[i] = a.w.Qx[j * NH + i];
        const float bj = a.w.b1[j] + a.w.bx[j];
        for (int r = t >> 7; r < Rp; r += 2) {
            float z = 0.f;
            if (r < R) {
                const float* sr = a.s + (size_t)(a.s_rows == 1 ? 0 : row0 + r) * NH;
                // (Q1 h + b1) + (Qx s + bx) is regrouped as Q1 h + (Qx s + (b1 + bx)): <= 1 ulp apart
#pragma unroll
                for (int i = 0; i < NH; ++i) z = fmaf(qx[i], sr[i], z);
                z += bj;
            }
            S.c[r * NM + j] = z;
        }
        for (int e = t; e < Rp * SROW; e += kThreads) {
            const int r = e / SROW, i = e - r * SROW;
            const float v = (r < R && i < NH) ? a.h0[(size_t)(row0 + r) * NH + i] : 0.f;
            S.y[e] = v; S.ys[e] = v;
            S.ce[e] = v; S.cd[e] = v; S.cc[e] = v; S.cb[e] = v; S.ca[e] = v;  // interp_coeff = [y0]*5
        }
        for (int e = t; e < R * NH; e += kThreads) {
            const int r = e / NH, i = e - r * NH;
            a.out[(size_t)(row0 + r) * NH + i] = a.h0[(size_t)(row0 + r) * NH + i];  // solution[0] = y0
        }
        if (t == 0) *S.orw = 0;
    }
    __syncthreads();

    const double inv_n = 1.0 / ((double)a.B * NH);
    const size_t out_stride = (size_t)a.B * NH * (CLUSTER ? a.G : 1);
    float* k0 = S.k;
    auto kst = [&](int s) { return S.k + s * Rp * SROW; };

    int n_accept = 0, n_reject = 0, nfe = 0, status = 0, n_logged = 0;
    double tcur = a.t[0], t_left = tcur, dt;

    // ---- f0 and the initial step (misc._select_initial_step) ------------------------------------
    eval_rhs<CLUSTER>(a, S, R, Rp, q1, q2, q3, b2j, b2j1, b3o, k0, comm); ++nfe;
    float h0s;
    {
        double s0 = 0, s1 = 0;
        for (int e = t; e < R * NH; e += kThreads) {
            const int r = e / NH, i = e - r * NH;
            const float y = S.y[r * SROW + i], f = k0[r * SROW + i];
            const float sc = a.atol + fabsf(y) * a.rtol;
            const float u = y / sc, v = f / sc;
            s0 += (double)u * u; s1 += (double)v * v;
        }
        cta_sum2(s0, s1, S.red);
        const Red rd = comm_reduce<CLUSTER>(comm, s0, s1, 0u, true, S.rr);
        const float d0 = sqrtf((float)(rd.s0 * inv_n)), d1 = sqrtf((float)(rd.s1 * inv_n));
        h0s = init_h0(d0, d1);
        for (int e = t; e < R * NH; e += kThreads) {
            const int r = e / NH, i = e - r * NH;
            S.ys[r * SROW + i] = S.y[r * SROW + i] + h0s * k0[r * SROW + i];
        }
        __syncthreads();
        eval_rhs<CLUSTER>(a, S, R, Rp, q1, q2, q3, b2j, b2j1, b3o, kst(1), comm); ++nfe;
        double s2 = 0, dummy = 0;
        for (int e = t; e < R * NH; e += kThreads) {
            const int r = e / NH, i = e - r * NH;
            const float sc = a.atol + fabsf(S.y[r * SROW + i]) * a.rtol;
            const float u = (kst(1)[r * SROW + i] - k0[r * SROW + i]) / sc;
            s2 += (double)u * u;
        }
        cta_sum2(s2, dummy, S.red);
        const Red rd2 = comm_reduce<CLUSTER>(comm, s2, 0.0, 0u, true, S.rr);
        const float d2 = sqrtf((float)(rd2.s0 * inv_n)) / h0s;
        dt = init_dt(h0s, d1, d2);
    }

    // ---- main loop: integrate through each requested output time ---------------------------------
    long long nsteps_total = 0;
    for (int it = 1; it < a.nt && status == 0; ++it) {
        const double tnext = a.t[it];
        long long nsteps = 0;
        while (tnext > tcur) {
            if (nsteps >= a.max_steps) { status = 3; break; }
            if (!(tcur + dt > tcur)) { status = 1; break; }
            const float dtf = (float)dt;
            // stages 1..6
            for (int s = 0; s < 6; ++s) {
                for (int e = t; e < R * NH; e += kThreads) {
                    const int r = e / NH, i = e - r * NH, o = r * SROW + i;
                    float acc = kst(0)[o] * (kBeta[s][0] * dtf);
                    for (int j = 1; j <= s; ++j) acc = fmaf(kst(j)[o], kBeta[s][j] * dtf, acc);
                    S.ys[o] = S.y[o] + acc;
                }
                __syncthreads();
                eval_rhs<CLUSTER>(a, S, R, Rp, q1, q2, q3, b2j, b2j1, b3o, kst(s + 1), comm); ++nfe;
            }
            // error estimate + norm (+ count of non-finite entries of the step's start state)
            double se = 0, nonfin = 0;
            for (int e = t; e < R * NH; e += kThreads) {
                const int r = e / NH, i = e - r * NH, o = r * SROW + i;
                float err = kst(0)[o] * (dtf * kCErr[0]);
#pragma unroll
                for (int j = 1; j < 7; ++j) err = fmaf(kst(j)[o], dtf * kCErr[j], err);
                const float y0 = S.y[o], y1 = S.ys[o];
                const float tol = a.atol + a.rtol * fmaxf(fabsf(y0), fabsf(y1));
                const float q = err / tol;
                se += (double)q * q;
                if (!isfinite(y0)) nonfin += 1.0;
            }
            cta_sum2(se, nonfin, S.red);
            const Red rd = comm_reduce<CLUSTER>(comm, se, nonfin, 0u, true, S.rr);
            if (rd.s1 > 0.0) { status = 2; break; }
            const float ratio = fabsf(sqrtf((float)(rd.s0 * inv_n)));
            if (comm.rank == 0 && t == 0 && n_logged < FIODE_LOG_STEPS) {
                stats->dts[n_logged] = dt; stats->ratios[n_logged] = ratio;
            }
            if (n_logged < FIODE_LOG_STEPS) ++n_logged;
            if (a.xlog_dt && comm.rank == 0 && t == 0 && nsteps_total < a.xlog_cap) {
                a.xlog_dt[(size_t)solve * a.xlog_cap + nsteps_total] = dt;
                a.xlog_ratio[(size_t)solve * a.xlog_cap + nsteps_total] = ratio;
            }
            if (ratio <= 1.f) {
                // accept: fit the quartic interpolant (interp._interp_fit), advance, FSAL
                for (int e = t; e < R * NH; e += kThreads) {
                    const int r = e / NH, i = e - r * NH, o = r * SROW + i;
                    float mid = kst(0)[o] * (dtf * kCMid[0]);
#pragma unroll
                    for (int j = 1; j < 7; ++j) mid = fmaf(kst(j)[o], dtf * kCMid[j], mid);
                    const float y0 = S.y[o], y1 = S.ys[o], f0 = kst(0)[o], f1 = kst(6)[o];
                    const float ym = y0 + mid;
                    S.ca[o] = 2.f * dtf * (f1 - f0) - 8.f * (y1 + y0) + 16.f * ym;
                    S.cb[o] = dtf * (5.f * f0 - 3.f * f1) + 18.f * y0 + 14.f * y1 - 32.f * ym;
                    S.cc[o] = dtf * (f1 - 4.f * f0) - 11.f * y0 - 5.f * y1 + 16.f * ym;
                    S.cd[o] = dtf * f0;
                    S.ce[o] = y0;
                    S.y[o] = y1;
                    kst(0)[o] = f1;
                }
                __syncthreads();
                t_left = tcur; tcur = tcur + dt; ++n_accept;
            } else {
                ++n_reject;
            }
            dt = next_dt(dt, ratio);
            ++nsteps; ++nsteps_total;
        }
        if (status) break;
        // dense output at tnext inside [t_left, tcur] (interp._interp_evaluate)
        const float x = (float)((tnext - t_left) / (tcur - t_left));
        for (int e = t; e < R * NH; e += kThreads) {
            const int r = e / NH, i = e - r * NH, o = r * SROW + i;
            float total = S.ce[o] + x * S.cd[o];
            float xp = x * x; total = total + xp * S.cc[o];
            xp = xp * x; total = total + xp * S.cb[o];
            xp = xp * x; total = total + xp * S.ca[o];
            a.out[it * out_stride + (size_t)(row0 + r) * NH + i] = total;
        }
    }
    if (comm.rank == 0 && t == 0) {
        stats->n_accept = n_accept; stats->n_reject = n_reject; stats->nfe = nfe;
        stats->status = status; stats->t_final = tcur; stats->last_dt = dt;
        stats->n_logged = n_logged;
    }
    if (CLUSTER) cg::this_cluster().sync();     // peers may still be reading this CTA's slots over DSMEM
}

static size_t solver_smem_bytes(int R) {
    const int Rp = (R + 3) & ~3;
    size_t f = (size_t)3 * Rp * NM + (size_t)(4 + 7 + 5) * Rp * SROW + (size_t)kMaxIter * Rp + 32 + 8 + 4 + 16;
    return f * sizeof(float);
}

}  // namespace fiode

using namespace fiode;

static int ensure_smem_attr() {
    static std::atomic<bool> done[kMaxDevices];
    if (first_use_on_device(done)) {
        const int mx = (int)solver_smem_bytes(64);
        FIODE_CUDA(cudaFuncSetAttribute(dopri5_simplex_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
        FIODE_CUDA(cudaFuncSetAttribute(dopri5_simplex_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
        // clusters of 16 CTAs (B up to 1024 rows per solve in the batched launch) are beyond the portable size of 8
        if (cudaFuncSetAttribute(dopri5_simplex_kernel<true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess)
            (void)cudaGetLastError();       // not granted: a launch with 16 then reports its own error
    }
    return 0;
}

extern "C" size_t fiode_dopri5_workspace_bytes(int B) { (void)B; return sizeof(SolveCtl); }

extern "C" int fiode_dopri5_simplex(const fiode_dyn_weights* w, const float* h0, const float* s, int s_rows,
                                    int B, const double* t, int nt, double rtol, double atol,
                                    long long max_num_steps, float* out, fiode_solve_stats* stats, void* ws,
                                    size_t ws_bytes, int rows_per_cta, void* stream) {
    return fiode_dopri5_simplex_logged(w, h0, s, s_rows, B, t, nt, rtol, atol, max_num_steps, out, stats, ws, ws_bytes,
                                       rows_per_cta, nullptr, nullptr, 0, stream);
}

// Same solve with a caller-owned log of EVERY step attempt (dt as the controller held it in fp64, error ratio in fp32):
// the differentiable solve replays the accepted ones.  Attempts beyond log_capacity are not recorded; the attempt count is
// stats->n_accept + stats->n_reject, so a caller that sees more attempts than capacity re-runs with a larger log.
extern "C" int fiode_dopri5_simplex_logged(const fiode_dyn_weights* w, const float* h0, const float* s, int s_rows,
                                           int B, const double* t, int nt, double rtol, double atol,
                                           long long max_num_steps, float* out, fiode_solve_stats* stats, void* ws,
                                           size_t ws_bytes, int rows_per_cta, double* log_dts, float* log_ratios,
                                           long long log_capacity, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if ((log_dts == nullptr) != (log_ratios == nullptr) || log_capacity < 0 || (log_dts && log_capacity == 0)) {
        set_error("fiode_dopri5_simplex_logged: log_dts / log_ratios / log_capacity must be given together"); return -1;
    }
    if (!w || !h0 || !s || !t || !out || !stats || !ws) { set_error("fiode_dopri5_simplex: null pointer"); return -1; }
    if (B <= 0 || nt < 1) { set_error("fiode_dopri5_simplex: B=%d nt=%d", B, nt); return -1; }
    if (s_rows != 1 && s_rows != B) { set_error("fiode_dopri5_simplex: s_rows must be 1 or B"); return -1; }
    if (ws_bytes < sizeof(SolveCtl)) { set_error("fiode_dopri5_simplex: workspace too small"); return -1; }
    if (w->max_iter < 1 || w->max_iter > kMaxIter) { set_error("fiode_dopri5_simplex: max_iter must be in [1,32]"); return -1; }
    const int nsm = sm_count();
    int R = rows_per_cta;
    if (R <= 0) {  // auto: spread over the SMs, at least 2 rows per CTA
        R = (B + nsm - 1) / nsm;
        if (R < 2) R = 2;
    }
    if (R > 64) { set_error("fiode_dopri5_simplex: batch %d needs %d rows per CTA (> 64); split the batch", B, R); return -2; }
    const int G = (B + R - 1) / R;
    if (G > nsm || G > kMaxGrid) { set_error("fiode_dopri5_simplex: grid %d exceeds co-resident CTAs %d", G, nsm); return -2; }
    const size_t smem = solver_smem_bytes(R);
    int rc = ensure_smem_attr();
    if (rc) return rc;
    FIODE_CUDA(cudaMemsetAsync(ws, 0, sizeof(SolveCtl), st));
    SolveArgs a;
    a.w = *w; a.h0 = h0; a.s = s; a.s_rows = s_rows; a.B = B; a.R = R; a.t = t; a.nt = nt;
    a.rtol = (float)rtol; a.atol = (float)atol; a.max_steps = max_num_steps; a.out = out; a.stats = stats;
    a.ctl = reinterpret_cast<SolveCtl*>(ws); a.G = 1; a.C = G;
    a.xlog_dt = log_dts; a.xlog_ratio = log_ratios; a.xlog_cap = log_capacity;
    void* args[] = {&a};
    FIODE_CUDA(cudaLaunchCooperativeKernel((void*)dopri5_simplex_kernel<false>, dim3(G), dim3(kThreads), args, smem, st));
    count_launch(1);
    return 0;
}

// G independent solves of B rows each in ONE launch: a thread-block cluster per solve (hardware cluster
// barrier + DSMEM reductions), so each solve keeps its own batch-wide error norm / projection stop rule.
extern "C" int fiode_dopri5_simplex_batched(const fiode_dyn_weights* w, const float* h0, const float* s, int s_rows,
                                            int G, int B, const double* t, int nt, double rtol, double atol,
                                            long long max_num_steps, float* out, fiode_solve_stats* stats,
                                            int cluster_size, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!w || !h0 || !s || !t || !out || !stats) { set_error("fiode_dopri5_simplex_batched: null pointer"); return -1; }
    if (G <= 0 || B <= 0 || nt < 1) { set_error("fiode_dopri5_simplex_batched: G=%d B=%d nt=%d", G, B, nt); return -1; }
    if (s_rows != 1 && s_rows != G * B) { set_error("fiode_dopri5_simplex_batched: s_rows must be 1 or G*B"); return -1; }
    if (w->max_iter < 1 || w->max_iter > kMaxIter) { set_error("fiode_dopri5_simplex_batched: max_iter must be in [1,32]"); return -1; }
    int C = cluster_size;
    if (C <= 0) { C = B > 512 ? 16 : 8; while (C > 1 && (B + C - 1) / C < 8) C >>= 1; }     // at least 8 rows per CTA
    if (C != 1 && C != 2 && C != 4 && C != 8 && C != 16) { set_error("fiode_dopri5_simplex_batched: cluster size must be 1, 2, 4, 8 or 16"); return -1; }
    const int R = (B + C - 1) / C;
    if (R > 64) { set_error("fiode_dopri5_simplex_batched: B=%d exceeds %d rows (cluster of %d x 64); use fiode_dopri5_simplex", B, C * 64, C); return -2; }
    int rc = ensure_smem_attr();
    if (rc) return rc;
    SolveArgs a;
    a.w = *w; a.h0 = h0; a.s = s; a.s_rows = s_rows; a.B = B; a.R = R; a.t = t; a.nt = nt;
    a.rtol = (float)rtol; a.atol = (float)atol; a.max_steps = max_num_steps; a.out = out; a.stats = stats;
    a.ctl = nullptr; a.G = G; a.C = C;
    a.xlog_dt = nullptr; a.xlog_ratio = nullptr; a.xlog_cap = 0;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(G * C)); cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = solver_smem_bytes(R); cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    FIODE_CUDA(cudaLaunchKernelEx(&cfg, dopri5_simplex_kernel<true>, a));
    count_launch(1);
    return 0;
}

// How many solves of B rows one launch of fiode_dopri5_simplex_batched keeps co-resident (one cluster each).  A launch with
// more solves than this runs in waves, each wave costing a full solve latency, so callers size G to a multiple of it.
extern "C" int fiode_dopri5_max_concurrent_solves(int B, int cluster_size) {
    int C = cluster_size;
    if (C <= 0) { C = B > 512 ? 16 : 8; while (C > 1 && (B + C - 1) / C < 8) C >>= 1; }
    if (C != 1 && C != 2 && C != 4 && C != 8 && C != 16) { set_error("fiode_dopri5_max_concurrent_solves: cluster size must be 1, 2, 4, 8 or 16"); return -1; }
    const int R = (B + C - 1) / C;
    if (B <= 0 || R > 64) { set_error("fiode_dopri5_max_concurrent_solves: B=%d does not fit a cluster of %d x 64 rows", B, C); return -2; }
    int rc = ensure_smem_attr();
    if (rc) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(C * 1024)); cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = solver_smem_bytes(R);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    int n = 0;
    FIODE_CUDA(cudaOccupancyMaxActiveClusters(&n, dopri5_simplex_kernel<true>, &cfg));
    return n;
}
