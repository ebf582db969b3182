// fiode_b200/csrc/certify_tc.cu -- Lipschitz grid certification on the sm_100a tensor cores.
//
// Replaces the per-image body of robustness/certify_lipschitz.py:104-143 (h_vdot :37-42 ->
// dynamics/classification.py:96-126 + barrier_projection/barrier_projection.py:232-255) like the FFMA
// kernel in rhs_rows.cu, with the three dense layers of the dynamics MLP on tcgen05:
//
//   CTA = (image m, part p), one per SM.  640 threads in five warpgroups (setmaxnreg re-splits the registers):
//     2 x row warpgroups ("slots"; thread = one grid point = one TMEM lane): S0, S1, S2 and the D3 read-out
//     2 x projection warpgroups: S3 on the rows their partner row warp hands over through a shared-memory ring
//     2 MMA-issuing warps (one thread each, one per slot).
//   TMEM (512 columns) = 2 slots x {D: 128 fp32 columns, A: 64 + 64 columns of packed fp16 hi / lo}.  Operand B
//   (the Cayley matrices, fp16x2 split, x64) lives in shared memory for the whole kernel; the activations never
//   leave TMEM / registers:
//
//     S0  grid numerators -> [eta, 1] fp16 hi/lo (LUT)  --tcgen05.st-->  A1           (K = 16)
//     L1  D1 = A1 . [Q1 | c]^T                          3 MMAs  128x128x16    c = b1 + bx + Qx s (per image)
//     S1  relu(D1) -> fp16 hi/lo                        --ld / st-->     A2
//     L2  D2 = A2 . Q2^T                                24 MMAs 128x128x16
//     S2  relu(D2 + b2) -> fp16 hi/lo                   --ld / st-->     A3
//     L3  D3 = A3_hi . [Q3_hi; Q3_lo]^T (N = 32) + A3_lo . Q3_hi^T (N = 16)      16 MMAs
//     S3  ftilde = D3 + b3 -> barrier lower bound (LUT), bisection projection with every candidate stop
//         iteration tracked (the stop rule is global per reference chunk), running maxima of the violation
//
//   The two slots run the same sequence on alternating tiles; while one waits for its MMAs the other owns the
//   FP32 pipes, and S0 of a slot's next tile is issued before its D3 hand-over.  Measured bound (ncu, round 1):
//   instruction issue (60 % busy; ALU pipe 45 %, tensor pipe 34 %), not the tensor pipe.
//   Numerics: fp16x2 split (see tc.cuh) ~ fp32; everything after D3 is the FFMA kernel's arithmetic.
#include "common.cuh"
#include "rows.cuh"
#include "tc.cuh"
#include <cstdlib>
#include <cstring>

namespace fiode {
using namespace tc;

constexpr int TM = 128;                    // rows per tile = TMEM lanes
constexpr int kTcThreads = 32 * 20;        // warpgroups: 2 x row, 2 x projection, 1 x {MMA warp + 3 idle}
constexpr int kMmaWarp = 16;
constexpr int kRegsRow = 80, kRegsProj = 136, kRegsMma = 40;   // setmaxnreg budgets; launch = 96 x 640
static_assert(256 * kRegsRow + 256 * kRegsProj + 128 * kRegsMma <= 96 * kTcThreads, "setmaxnreg.inc would block: the CTA pool is what the launch allocated");
constexpr float kWS = 64.f, kWSInv = 1.f / 64.f;

// shared-memory map (bytes).  B operands: [k chunk of 8 halves][n][8 halves]  (K-major core matrices)
constexpr int OFF_B2H = 0;                 // Q2 hi   [16][128][8]
constexpr int OFF_B2L = 32768;             // Q2 lo
constexpr int OFF_B1H = 65536;             // [Q1 | c | 0] hi   [2][128][8]
constexpr int OFF_B1L = 69632;
constexpr int OFF_B3 = 73728;              // rows 0..15 Q3 hi (10 used), rows 16..31 Q3 lo   [16][32][8]
constexpr int OFF_B2S = 81920;             // b2 / 2 fp32 [128]
constexpr int OFF_B3S = OFF_B2S + 512;     // b3 fp32 [16]
constexpr int OFF_LUT_HL = OFF_B3S + 64;   // u32 [256]: fp16 hi | lo << 16 of eta(z)
constexpr int OFF_LUT_LO = OFF_LUT_HL + 1024;   // fp32 [256]: barrier lower bound at eta(z)
constexpr int OFF_LUT_PT = OFF_LUT_LO + 1024;   // fp32 [256]: Lipschitz perturbation for eta_max = eta(z)
constexpr int OFF_BARS = OFF_LUT_PT + 1024;     // mbarriers: A ready [2], D ready [2], ring full [16], ring empty [16]
constexpr int OFF_TMEM = OFF_BARS + 36 * 8;
constexpr int OFF_RING = OFF_TMEM + 32;         // [slot 2][quarter 4][entry 2] x { ft [10][32] fp32, z [3][32] u32 }
constexpr int kRingEntry = 13 * 32 * 4;
constexpr int kTcSmemBytes = 120 * 1024;   // > half an SM: one CTA (and one 512-column TMEM allocation) per SM
static_assert(OFF_RING + 16 * kRingEntry <= kTcSmemBytes, "smem map");
static_assert(OFF_RING % 16 == 0, "ring alignment");

// optional in-kernel stage timers (build with -DFIODE_TC_PROF; read back with fiode_debug_tc_prof)
#ifdef FIODE_TC_PROF
__device__ long long g_tc_prof[32];
#define PROF_DECL long long pacc_[16] = {0}; long long pt_ = clock64(), pn_
#define PROF(i) do { pn_ = clock64(); pacc_[(i) & 15] += pn_ - pt_; pt_ = pn_; } while (0)
#define PROF_FLUSH(cond) do { if ((cond) && blockIdx.x == 0 && lane == 0) for (int i_ = 0; i_ < 14; ++i_) if (pacc_[i_]) g_tc_prof[i_] = pacc_[i_]; } while (0)
#else
#define PROF_DECL
#define PROF(i)
#define PROF_FLUSH(cond)
#endif

__device__ __forceinline__ void put_b(uint8_t* base, int N, int n, int k, unsigned short v) {
    *reinterpret_cast<unsigned short*>(base + ((size_t)(k >> 3) * N + n) * 16 + (k & 7) * 2) = v;
}

// S1 / S2: D (128 fp32 columns of this thread's lane) -> relu(D / 64 [+ bias]) -> fp16 hi / lo -> A
// relu costs no ALU-pipe slot here: with y = x / 2,  y + |y| == relu(x) exactly (|.| is a free FADD source modifier;
// FMNMX / FSEL / ISETP share the half-rate ALU pipe, which is what bounds this kernel), so D is scaled by 1/128
// instead of 1/64 and the bias is pre-halved.
template <bool kBias>
__device__ __forceinline__ void act16(const uint32_t (&v)[16], uint32_t tA, int c, const float* half_bias) {
    const float2 sc = make_float2(0.5f * kWSInv, 0.5f * kWSInv), neg1 = make_float2(-1.f, -1.f);
    uint32_t hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float2 x = make_float2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
        if (kBias) x = __ffma2_rn(x, sc, *reinterpret_cast<const float2*>(half_bias + 16 * c + 2 * i));
        else x = __fmul2_rn(x, sc);
        x.x = x.x + fabsf(x.x); x.y = x.y + fabsf(x.y);
        const __half2 h = __float22half2_rn(x);
        const float2 r = __ffma2_rn(__half22float2(h), neg1, x);      // x - float(h), exact
        const __half2 l = __float22half2_rn(r);
        hi[i] = *reinterpret_cast<const uint32_t*>(&h);
        lo[i] = *reinterpret_cast<const uint32_t*>(&l);
    }
    tmem_st8(tA + 8 * c, hi);
    tmem_st8(tA + 64 + 8 * c, lo);
}
// 16-column chunks, the TMEM load of chunk c+1 in flight while chunk c is converted.  (The operand cannot be
// published in halves to start the MMAs earlier: they overwrite the very D columns still being read here.)
template <bool kBias>
__device__ __forceinline__ void act_to_a(uint32_t tD, uint32_t tA, const float* bias) {
    uint32_t v0[16], v1[16];
    tmem_ld16(tD, v0);
#pragma unroll
    for (int c = 0; c < 8; c += 2) {
        tmem_wait_ld();
        tmem_ld16(tD + 16 * (c + 1), v1);
        act16<kBias>(v0, tA, c, bias);
        tmem_wait_ld();
        if (c + 2 < 8) tmem_ld16(tD + 16 * (c + 2), v0);
        act16<kBias>(v1, tA, c + 1, bias);
    }
}

__global__ void __launch_bounds__(kTcThreads, 1) certify_tc_kernel(const RowsArgs a, unsigned* __restrict__ status) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int m = blockIdx.x / a.P, p = blockIdx.x - m * a.P;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BARS);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + OFF_TMEM);
    float* b2s = reinterpret_cast<float*>(smem + OFF_B2S);
    float* b3s = reinterpret_cast<float*>(smem + OFF_B3S);
    uint32_t* lut_hl = reinterpret_cast<uint32_t*>(smem + OFF_LUT_HL);
    float* lut_lo = reinterpret_cast<float*>(smem + OFF_LUT_LO);
    float* lut_pt = reinterpret_cast<float*>(smem + OFF_LUT_PT);
    const int label = a.labels[m];

    // ---- one-time setup ---------------------------------------------------------------------------
    if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
    if (t == 0) {
        mbar_init(&bars[0], TM); mbar_init(&bars[1], TM); mbar_init(&bars[2], 1); mbar_init(&bars[3], 1);
        for (int i = 4; i < 36; ++i) mbar_init(&bars[i], 32);
        mbar_init_fence();
    }
    for (int e = t; e < NM * NM; e += kTcThreads) {           // Q2 (n, k) row-major
        const int n = e >> 7, k = e & 127;
        unsigned short h, l;
        split1(a.w.Q2[e] * kWS, h, l);
        put_b(smem + OFF_B2H, NM, n, k, h); put_b(smem + OFF_B2L, NM, n, k, l);
    }
    for (int e = t; e < NM * 16; e += kTcThreads) {           // [Q1 | c | 0]: c = b1 + bx + Qx s (rhs_rows.cu set_static_feature)
        const int n = e >> 4, k = e & 15;
        float v = 0.f;
        if (k < NH) v = a.w.Q1[n * NH + k];
        else if (k == NH) {
            float z = 0.f;
#pragma unroll
            for (int i = 0; i < NH; ++i) z = fmaf(__ldg(a.w.Qx + n * NH + i), __ldg(a.s + (size_t)m * NH + i), z);
            v = z + (a.w.b1[n] + a.w.bx[n]);
        }
        unsigned short h, l;
        split1(v * kWS, h, l);
        put_b(smem + OFF_B1H, NM, n, k, h); put_b(smem + OFF_B1L, NM, n, k, l);
    }
    for (int e = t; e < 32 * NM; e += kTcThreads) {           // rows 0..15: Q3 hi, rows 16..31: Q3 lo
        const int n = e >> 7, k = e & 127, o = n & 15;
        unsigned short h = 0, l = 0;
        if (o < NH) split1(a.w.Q3[o * NM + k] * kWS, h, l);
        put_b(smem + OFF_B3, 32, n, k, n < 16 ? h : l);
    }
    for (int e = t; e < NM; e += kTcThreads) b2s[e] = 0.5f * a.w.b2[e];   // pre-halved, see act16
    if (t < 16) b3s[t] = t < NH ? a.w.b3[t] : 0.f;
    for (int z = t; z < 256; z += kTcThreads) {
        const float eta = z <= a.T ? a.lut[z] : 0.f;
        unsigned short h, l;
        split1(eta, h, l);
        lut_hl[z] = (uint32_t)h | ((uint32_t)l << 16);
        lut_lo[z] = barrier_lower(a.w.alpha_1, a.w.sigma_1, eta);
        // certify_lipschitz.py:122-124 (same roundings as the FFMA kernel)
        const float Lf = __fadd_rn(__fmul_rn(a.sqrtn, a.sa * exp_cr(a.w.sigma_1 * (eta + a.inv_T))), 1.f);
        lut_pt[z] = (a.sqrt2 * Lf) * a.dist;
    }
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    const long long stride = 2LL * a.P;
    auto tiles_in_chunk = [&](int ch) -> long long {
        const long long c0 = (long long)ch * a.chunk;
        const long long c1 = (ch == a.n_chunks - 1) ? a.Ng : min(a.Ng, c0 + a.chunk);
        return (c1 - c0 + TM - 1) / TM;
    };

    if (warp >= kMmaWarp) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsMma));
        // ================= MMA issuers: warp 16 -> slot 0, warp 17 -> slot 1 (one thread each) ========
        // (one issuer per slot on different SM sub-partitions: an issuer shares its scheduler with four
        //  ALU-bound warps, so its ~10 dependent instructions per MMA take longer than the MMA itself)
        if (warp < kMmaWarp + 2 && lane == 0) {
            const int s = warp - kMmaWarp;
            long long remaining = 0;
            for (int ch = 0; ch < a.n_chunks; ++ch) {
                const long long nt = tiles_in_chunk(ch);
                const long long first = (long long)p + (long long)s * a.P;
                if (nt > first) remaining += (nt - first + stride - 1) / stride;
            }
            const uint32_t sb = smem_u32(smem);
            constexpr uint32_t id128 = idesc_f16(128, 128), id32 = idesc_f16(128, 32), id16 = idesc_f16(128, 16);
            const uint32_t tD = tmem + s * 256, tA = tD + 128;
            const uint64_t b1h = smem_desc(sb + OFF_B1H, NM * 16, 128), b1l = smem_desc(sb + OFF_B1L, NM * 16, 128);
            const uint64_t b2h = smem_desc(sb + OFF_B2H, NM * 16, 128), b2l = smem_desc(sb + OFF_B2L, NM * 16, 128);
            const uint64_t b3 = smem_desc(sb + OFF_B3, 32 * 16, 128);
            constexpr uint64_t k2 = (2 * NM * 16) >> 4, k3 = (2 * 32 * 16) >> 4;   // descriptor step per 16-wide k slab
            uint64_t* a_rdy = &bars[s];
            uint64_t* d_rdy = &bars[2 + s];
            uint32_t par = 0;
            PROF_DECL;
            for (; remaining > 0; --remaining) {
                mbar_wait(a_rdy, par); par ^= 1; fence_after_sync();
                PROF(10);
                mma_ts_f16(tD, tA, b1h, id128, 0);
                mma_ts_f16(tD, tA + 8, b1h, id128, 1);
                mma_ts_f16(tD, tA, b1l, id128, 1);
                mma_commit(d_rdy);
                PROF(11);
                mbar_wait(a_rdy, par); par ^= 1; fence_after_sync();
                PROF(10);
                mma_ts_f16_x8(tD, tA, b2h, k2, id128, 0);          // hi . hi
                mma_ts_f16_x8(tD, tA + 64, b2h, k2, id128, 1);     // lo . hi
                mma_ts_f16_x8(tD, tA, b2l, k2, id128, 1);          // hi . lo
                mma_commit(d_rdy);
                PROF(12);
                mbar_wait(a_rdy, par); par ^= 1; fence_after_sync();
                PROF(10);
                mma_ts_f16_x8(tD, tA, b3, k3, id32, 0);            // hi . [hi; lo]  (N = 32)
                mma_ts_f16_x8(tD, tA + 64, b3, k3, id16, 1);       // lo . hi        (N = 16)
                mma_commit(d_rdy);
                PROF(13);
            }
            PROF_FLUSH(s == 0);
        }
    } else {
        // row warp w (0..7) and projection warp 8 + w work on the same 32 rows: slot = w / 4, quarter = w % 4
        const bool is_proj = warp >= 8;
        const int w = warp & 7;
        const int slot = w >> 2, q = w & 3;
        const long long first = (long long)p + (long long)slot * a.P;
        uint8_t* ring = smem + OFF_RING + (size_t)(w * 2) * kRingEntry;
        uint64_t* full = &bars[4 + w * 2];
        uint64_t* empty = &bars[20 + w * 2];
        int ch = 0;
        long long tile = first;
        auto settle = [&]() -> bool {
            while (ch < a.n_chunks && tile >= tiles_in_chunk(ch)) { ++ch; tile = first; }
            return ch < a.n_chunks;
        };
        bool have = settle();
        int e = 0;

        if (!is_proj) {
            asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsRow));
            // ================= row warps: TMEM lane = 32 * quarter + lane =============================
            const uint32_t tD = tmem + ((uint32_t)(q * 32) << 16) + slot * 256, tA = tD + 128;
            uint64_t* a_rdy = &bars[slot];
            uint64_t* d_rdy = &bars[2 + slot];
            uint32_t par_d = 0, par_e = 3;     // ring entries start empty
            const int r = q * 32 + lane;
            // S0: numerators of this thread's grid point (label swap of eval_utils.py:64-67 applied) -> A1
            // the 10 numerator bytes of this thread's grid point, fetched one tile ahead (global-load latency
            // would otherwise sit on the slot's critical path); raw[5] bit 0 = row exists
            auto load_z = [&](int chx, long long tilex, unsigned (&raw)[6]) {
                const long long c0 = (long long)chx * a.chunk;
                const long long c1 = (chx == a.n_chunks - 1) ? a.Ng : min(a.Ng, c0 + a.chunk);
                const long long g = c0 + tilex * TM + r;
                raw[5] = g < c1;
#pragma unroll
                for (int i = 0; i < NH / 2; ++i) raw[i] = 0;
                if (raw[5]) {
                    const unsigned short* src = reinterpret_cast<const unsigned short*>(a.gridz + g * NH);
#pragma unroll
                    for (int i = 0; i < NH / 2; ++i) raw[i] = __ldg(src + i);
                }
            };
            auto stage0 = [&](const unsigned (&raw)[6], uint32_t (&zw)[3]) {
                const bool valid = raw[5] != 0;
                int z[NH];
#pragma unroll
                for (int i = 0; i < NH / 2; ++i) { z[2 * i] = raw[i] & 0xff; z[2 * i + 1] = raw[i] >> 8; }
                int zl = z[0];
#pragma unroll
                for (int i = 1; i < NH; ++i) if (i == label) zl = z[i];
                const int z0 = z[0];
#pragma unroll
                for (int i = 1; i < NH; ++i) if (i == label) z[i] = z0;
                z[0] = zl;
                zw[0] = z[0] | (z[1] << 8) | (z[2] << 16) | (z[3] << 24);
                zw[1] = z[4] | (z[5] << 8) | (z[6] << 16) | (z[7] << 24);
                zw[2] = z[8] | (z[9] << 8) | (valid ? 0x80000000u : 0u);
                uint32_t hi[8], lo[8];
#pragma unroll
                for (int i = 0; i < NH / 2; ++i) {
                    const uint32_t u0 = lut_hl[z[2 * i]], u1 = lut_hl[z[2 * i + 1]];
                    hi[i] = __byte_perm(u0, u1, 0x5410);
                    lo[i] = __byte_perm(u0, u1, 0x7632);
                }
                hi[5] = 0x3C00u; lo[5] = 0u;          // k = 10: constant 1 -> picks up c
                hi[6] = hi[7] = lo[6] = lo[7] = 0u;
                tmem_st8(tA, hi);
                tmem_st8(tA + 8, lo);
                tmem_wait_st();
                fence_before_sync();
                mbar_arrive(a_rdy);
            };
            uint32_t zc[3] = {0, 0, 0};
            unsigned raw[6] = {0, 0, 0, 0, 0, 0};
            if (have) { load_z(ch, tile, raw); stage0(raw, zc); }
            PROF_DECL;
#ifdef FIODE_TC_PROF
#define RPROF(i) PROF(i)
#else
#define RPROF(i)
#endif
            while (have) {
                tile += stride;                      // the slot's next tile: numerators requested now, used after D3
                have = settle();
                if (have) load_z(ch, tile, raw);
                mbar_wait(d_rdy, par_d); par_d ^= 1; fence_after_sync();              // D1
                RPROF(0);
                act_to_a<false>(tD, tA, nullptr);
                tmem_wait_st(); fence_before_sync(); mbar_arrive(a_rdy);
                RPROF(1);
                mbar_wait(d_rdy, par_d); par_d ^= 1; fence_after_sync();              // D2
                RPROF(2);
                act_to_a<true>(tD, tA, b2s);
                tmem_wait_st(); fence_before_sync(); mbar_arrive(a_rdy);
                RPROF(3);
                mbar_wait(d_rdy, par_d); par_d ^= 1; fence_after_sync();              // D3
                RPROF(4);
                uint32_t v[32];
                tmem_ld32(tD, v);
                tmem_wait_ld();
                // next tile of this slot first: its layer-1 MMAs overlap the hand-over below
                uint32_t zn[3] = {0, 0, 0};
                if (have) stage0(raw, zn);
                // hand ftilde + numerators to the projection warp
                RPROF(5);
                mbar_wait(&empty[e], (par_e >> e) & 1); par_e ^= 1u << e;
                RPROF(6);
                float* rf = reinterpret_cast<float*>(ring + e * kRingEntry);
                uint32_t* rz = reinterpret_cast<uint32_t*>(rf + NH * 32);
#pragma unroll
                for (int i = 0; i < NH; ++i)
                    rf[i * 32 + lane] = fmaf(__uint_as_float(v[i]) + __uint_as_float(v[16 + i]), kWSInv, b3s[i]);
                rz[lane] = zc[0]; rz[32 + lane] = zc[1]; rz[64 + lane] = zc[2];
                mbar_arrive(&full[e]);
                RPROF(7);
                e ^= 1;
                zc[0] = zn[0]; zc[1] = zn[1]; zc[2] = zn[2];
            }
            PROF_FLUSH(warp == 0);
        } else {
            // ================= projection warps: S3 (rhs_rows.cu kCertify epilogue) ====================
            asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsProj));
            float m1[kIt], m2[kIt];
#pragma unroll
            for (int k = 0; k < kIt; ++k) { m1[k] = -INFINITY; m2[k] = -INFINITY; }
            unsigned notconv = 0;
            bool bad = false;
            uint32_t par_f = 0;
            PROF_DECL;
#ifdef FIODE_TC_PROF
#define PPROF(i) PROF(i)
#else
#define PPROF(i)
#endif
            while (have) {
                const int cur_ch = ch;
                mbar_wait(&full[e], (par_f >> e) & 1); par_f ^= 1u << e;
                PPROF(8);
                const float* rf = reinterpret_cast<const float*>(ring + e * kRingEntry);
                const uint32_t* rz = reinterpret_cast<const uint32_t*>(rf + NH * 32);
                float ft[NH];
#pragma unroll
                for (int i = 0; i < NH; ++i) ft[i] = rf[i * 32 + lane];
                const uint32_t z0 = rz[lane], z1 = rz[32 + lane], z2 = rz[64 + lane];
                mbar_arrive(&empty[e]);
                e ^= 1;
                tile += stride;
                have = settle();
                if (z2 & 0x80000000u) {
                    int zc[NH];
                    zc[0] = z0 & 0xff; zc[1] = (z0 >> 8) & 0xff; zc[2] = (z0 >> 16) & 0xff; zc[3] = z0 >> 24;
                    zc[4] = z1 & 0xff; zc[5] = (z1 >> 8) & 0xff; zc[6] = (z1 >> 16) & 0xff; zc[7] = z1 >> 24;
                    zc[8] = z2 & 0xff; zc[9] = (z2 >> 8) & 0xff;
                    float lo[NH];
                    int zmax = zc[0];
#pragma unroll
                    for (int i = 0; i < NH; ++i) {
                        lo[i] = lut_lo[zc[i]];
                        zmax = max(zmax, zc[i]);
                        if (a.w.scale_nominal) ft[i] = scale_nominal_f(a.w.alpha_2, a.lut[zc[i]], lo[i], ft[i]);
                    }
                    float chk = 0.f;
#pragma unroll
                    for (int i = 0; i < NH; ++i) chk += ft[i];
                    if (!(fabsf(chk) <= 3.0e38f)) bad = true;
                    // max over the tie set W = {i != label : eta_i == eta_max} of max(ft_i - mu, lo_i)
                    //   = max(max_W ft_i - mu, lower(eta_max))   (rounding is monotone; lo_i is equal on W)
                    float FW = -INFINITY, fty = ft[0], loy = lo[0];
#pragma unroll
                    for (int i = 0; i < NH; ++i) {
                        if (zc[i] == zmax && i != label) FW = fmaxf(FW, ft[i]);
                        if (i == label) { fty = ft[i]; loy = lo[i]; }
                    }
                    float LW = (FW == -INFINITY) ? -INFINITY : lut_lo[zmax];
                    asm volatile("" : "+f"(LW));
                    float perturb = lut_pt[zmax];
                    // keep these five in registers: ptxas otherwise re-derives the label selects in every iteration
                    asm volatile("" : "+f"(fty), "+f"(loy), "+f"(FW), "+f"(perturb));
                    float hi = ft[0] - lo[0], lw = ft[0];
#pragma unroll
                    for (int j = 1; j < NH; ++j) { hi = fmaxf(hi, ft[j] - lo[j]); lw = fminf(lw, ft[j]); }
                    // the bisection runs on the NEGATED bounds: -((hi - lw) / 2 + lw) == ((-hi) - (-lw)) / 2 + (-lw)
                    // bit for bit (rounding is symmetric), and every use of mu below is "x - mu"
                    float nhi = -hi, nlw = -lw;
#pragma unroll
                    for (int k = 0; k < kIt; ++k) {
                        if (k < a.w.max_iter) {
                            const float nm = fmaf(nhi - nlw, 0.5f, nlw);
                            const float2 nmu = make_float2(nm, nm);
                            float e_ = 0.f;
#pragma unroll
                            for (int j = 0; j < NH; j += 2) {
                                const float2 d = __fadd2_rn(make_float2(ft[j], ft[j + 1]), nmu);   // ft - mu
                                e_ += fmaxf(d.x, lo[j]);
                                e_ += fmaxf(d.y, lo[j + 1]);
                            }
                            const float fy = fmaxf(fty + nm, loy), fw = fmaxf(FW + nm, LW);
                            const float hx = -fy + fw;                     // certify_lipschitz.py:41
                            m2[k] = fmaxf(m2[k], hx);
                            m1[k] = fmaxf(m1[k], hx + perturb);
                            if (!(fabsf(e_) < a.w.tol)) notconv |= (1u << k);
                            if (e_ > 0.f) nlw = nm;
                            if (e_ < 0.f) nhi = nm;
                        }
                    }
                }
                PPROF(9);
                if (!have || ch != cur_ch) {
                    // chunk finished for these rows: warp reduce and publish
                    notconv = __reduce_or_sync(0xffffffffu, notconv);
                    unsigned* dst = a.cmax + ((size_t)m * a.n_chunks + cur_ch) * 2 * kIt;
#pragma unroll
                    for (int k = 0; k < kIt; ++k) {
                        const float x1 = warp_max(m1[k]), x2 = warp_max(m2[k]);
                        if (lane == 0 && k < a.w.max_iter) {
                            atomicMax(dst + k, f2ord(x1));
                            atomicMax(dst + kIt + k, f2ord(x2));
                        }
                        m1[k] = -INFINITY; m2[k] = -INFINITY;
                    }
                    if (lane == 0 && notconv) atomicOr(a.cnot + (size_t)m * a.n_chunks + cur_ch, notconv);
                    notconv = 0;
                }
            }
            if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(status, 1u);
            PROF_FLUSH(warp == 8);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == kMmaWarp) tmem_dealloc(tmem, 512);
}

static int g_mlp_path = -1;
int mlp_path() {
    if (g_mlp_path < 0) {
        const char* e = getenv("FIODE_MLP_PATH");
        g_mlp_path = (e && !strcmp(e, "ffma")) ? 1 : 0;
    }
    return g_mlp_path;
}

int launch_certify_tc(const RowsArgs& a0, cudaStream_t st) {
    RowsArgs a = a0;
    FIODE_CUDA(cudaFuncSetAttribute(certify_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
    const int slots = sm_count();            // one CTA per SM
    int P = slots / a.M; if (P < 1) P = 1;
    const long long max_tiles = (a.chunk + TM - 1) / TM + 1;
    if (P > max_tiles) P = (int)max_tiles;
    a.P = P;
    unsigned* status = a.cnot + (size_t)a.M * a.n_chunks;
    certify_tc_kernel<<<a.M * P, kTcThreads, kTcSmemBytes, st>>>(a, status);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}

}  // namespace fiode

using namespace fiode;

#ifdef FIODE_TC_PROF
extern "C" __attribute__((visibility("default"))) int fiode_debug_tc_prof(long long* out32, int reset) {
    if (out32) cudaMemcpyFromSymbol(out32, fiode::g_tc_prof, sizeof(long long) * 32);
    if (reset) { long long z[32] = {0}; cudaMemcpyToSymbol(fiode::g_tc_prof, z, sizeof(z)); }
    return 0;
}
#endif

extern "C" void fiode_set_mlp_path(int path) { fiode::g_mlp_path = path ? 1 : 0; }
extern "C" int fiode_get_mlp_path(void) { return fiode::mlp_path(); }

extern "C" size_t fiode_certify_workspace_bytes(int M, int n_chunks) {
    return ((size_t)M * n_chunks * (2 * kIt + 1) + 1) * sizeof(unsigned);
}

extern "C" int fiode_certify_grid(const fiode_dyn_weights* w, const uint8_t* gridz, long long Ng, int T,
                                  const float* lut, const float* s, const int* labels, int M, int batches,
                                  float kappa, float* out, int* iters, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!w || !gridz || !lut || !s || !labels || !out || !ws) { set_error("fiode_certify_grid: null pointer"); return -1; }
    if (M <= 0 || Ng <= 0 || batches <= 0 || T <= 0 || T > 255) { set_error("fiode_certify_grid: bad sizes"); return -1; }
    if (w->max_iter < 1 || w->max_iter > kIt) { set_error("fiode_certify_grid: max_iter must be in [1,32]"); return -1; }
    // chunking of certify_lipschitz.py:100-102, 116-119
    const long long chunk = Ng / batches;
    if (chunk == 0) { set_error("fiode_certify_grid: grid has fewer rows (%lld) than batches (%d)", Ng, batches); return -1; }
    const int n_chunks = batches + ((Ng % batches) != 0 ? 1 : 0);
    if (ws_bytes < fiode_certify_workspace_bytes(M, n_chunks)) { set_error("fiode_certify_grid: workspace too small"); return -1; }
    unsigned* cmax = reinterpret_cast<unsigned*>(ws);
    unsigned* cnot = cmax + (size_t)M * n_chunks * 2 * kIt;
    unsigned* status = cnot + (size_t)M * n_chunks;
    FIODE_CUDA(cudaMemsetAsync(ws, 0, fiode_certify_workspace_bytes(M, n_chunks), st));  // ord(0) < ord(any float)
    RowsArgs a{};
    a.w = *w; a.s = s; a.gridz = gridz; a.lut = lut; a.T = T; a.Ng = Ng; a.chunk = chunk; a.n_chunks = n_chunks;
    a.labels = labels; a.M = M; a.cmax = cmax; a.cnot = cnot;
    // certify_lipschitz.py:122-124: Lf = sqrt(N) * (sigma_1*alpha_1*exp(sigma_1*eta_ub)) + 1 ; perturb = (sqrt(2)*Lf)*dist
    a.sa = (float)((double)w->sigma_1 * (double)w->alpha_1);
    a.sqrtn = (float)sqrt((double)NH);
    a.sqrt2 = (float)sqrt(2.0);
    a.dist = (float)(sqrt((double)NH) / (double)T);
    a.inv_T = (float)(1.0 / (double)T);
    const int rc = mlp_path() == 0 ? launch_certify_tc(a, st) : launch_certify_ffma(a, st);
    if (rc) return rc;
    return launch_certify_finish(cmax, cnot, status, M, n_chunks, w->max_iter, kappa, out, iters, st);
}
