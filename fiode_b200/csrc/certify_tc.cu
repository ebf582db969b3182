// fiode_b200/csrc/certify_tc.cu -- Lipschitz grid certification on the sm_100a tensor cores.
//
// Replaces the per-image body of robustness/certify_lipschitz.py:104-143 (h_vdot :37-42 ->
// dynamics/classification.py:96-126 + barrier_projection/barrier_projection.py:232-255) like the FFMA
// kernel in rhs_rows.cu, with the three dense layers of the dynamics MLP on tcgen05:
//
//   CTA = (image m, part p).  288 threads: two row warpgroups ("slots", 4 warps each; thread = one grid
//   point = one TMEM lane) + one MMA-issuing warp.  TMEM (512 columns) = 2 slots x {D: 128 fp32 columns,
//   A: 64 + 64 columns of packed fp16 hi / lo}.  Operand B (the Cayley matrices, fp16x2 split, x64) lives in
//   shared memory for the whole kernel; the activations never leave TMEM / registers:
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
//   The two slots run the same sequence on alternating tiles; while one waits for its MMAs the other
//   owns the FP32 pipes, and S0 of a slot's next tile is issued before S3 of the current one.
//   Numerics: fp16x2 split (see tc.cuh) ~ fp32; everything after D3 is the FFMA kernel's arithmetic.
#include "common.cuh"
#include "rows.cuh"
#include "tc.cuh"
#include <cstdlib>
#include <cstring>

namespace fiode {
using namespace tc;

constexpr int TM = 128;                    // rows per tile = TMEM lanes
constexpr int kTcThreads = 288;            // 8 row warps + 1 MMA warp
constexpr float kWS = 64.f, kWSInv = 1.f / 64.f;

// shared-memory map (bytes).  B operands: [k chunk of 8 halves][n][8 halves]  (K-major core matrices)
constexpr int OFF_B2H = 0;                 // Q2 hi   [16][128][8]
constexpr int OFF_B2L = 32768;             // Q2 lo
constexpr int OFF_B1H = 65536;             // [Q1 | c | 0] hi   [2][128][8]
constexpr int OFF_B1L = 69632;
constexpr int OFF_B3 = 73728;              // rows 0..15 Q3 hi (10 used), rows 16..31 Q3 lo   [16][32][8]
constexpr int OFF_B2S = 81920;             // b2 fp32 [128]
constexpr int OFF_B3S = OFF_B2S + 512;     // b3 fp32 [16]
constexpr int OFF_LUT_HL = OFF_B3S + 64;   // u32 [256]: fp16 hi | lo << 16 of eta(z)
constexpr int OFF_LUT_LO = OFF_LUT_HL + 1024;   // fp32 [256]: barrier lower bound at eta(z)
constexpr int OFF_LUT_PT = OFF_LUT_LO + 1024;   // fp32 [256]: Lipschitz perturbation for eta_max = eta(z)
constexpr int OFF_BARS = OFF_LUT_PT + 1024;     // 4 mbarriers
constexpr int OFF_TMEM = OFF_BARS + 32;
constexpr int kTcSmemBytes = 120 * 1024;   // > half an SM: one CTA (and one 512-column TMEM allocation) per SM
static_assert(OFF_TMEM + 16 <= kTcSmemBytes, "smem map");

__device__ __forceinline__ void put_b(uint8_t* base, int N, int n, int k, unsigned short v) {
    *reinterpret_cast<unsigned short*>(base + ((size_t)(k >> 3) * N + n) * 16 + (k & 7) * 2) = v;
}

// S1 / S2: D (128 fp32 columns of this thread's lane) -> relu(D / 64 [+ bias]) -> fp16 hi / lo -> A
template <bool kBias>
__device__ __forceinline__ void act_to_a(uint32_t tD, uint32_t tA, const float* bias) {
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
        uint32_t v[32], hi[16], lo[16];
        tmem_ld32(tD + 32 * c, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            float x0 = __uint_as_float(v[2 * i]), x1 = __uint_as_float(v[2 * i + 1]);
            if (kBias) {
                const float2 b = *reinterpret_cast<const float2*>(bias + 32 * c + 2 * i);
                x0 = fmaf(x0, kWSInv, b.x); x1 = fmaf(x1, kWSInv, b.y);
            } else {
                x0 *= kWSInv; x1 *= kWSInv;
            }
            split2(fmaxf(x0, 0.f), fmaxf(x1, 0.f), hi[i], lo[i]);
        }
        tmem_st16(tA + 16 * c, hi);
        tmem_st16(tA + 64 + 16 * c, lo);
    }
}

__global__ void __launch_bounds__(kTcThreads, 1) certify_tc_kernel(const RowsArgs a, unsigned* __restrict__ status) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int m = blockIdx.x / a.P, p = blockIdx.x - m * a.P;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BARS);   // [0,1] A ready (slot), [2,3] D ready (slot)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + OFF_TMEM);
    float* b2s = reinterpret_cast<float*>(smem + OFF_B2S);
    float* b3s = reinterpret_cast<float*>(smem + OFF_B3S);
    uint32_t* lut_hl = reinterpret_cast<uint32_t*>(smem + OFF_LUT_HL);
    float* lut_lo = reinterpret_cast<float*>(smem + OFF_LUT_LO);
    float* lut_pt = reinterpret_cast<float*>(smem + OFF_LUT_PT);
    const int label = a.labels[m];

    // ---- one-time setup ---------------------------------------------------------------------------
    if (warp == 8) tmem_alloc(tmem_slot, 512);
    if (t == 0) {
        mbar_init(&bars[0], TM); mbar_init(&bars[1], TM); mbar_init(&bars[2], 1); mbar_init(&bars[3], 1);
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
    for (int e = t; e < NM; e += kTcThreads) b2s[e] = a.w.b2[e];
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
    const long long first_of_slot[2] = {(long long)p, (long long)p + a.P};
    const long long stride = 2LL * a.P;
    auto tiles_in_chunk = [&](int ch) -> long long {
        const long long c0 = (long long)ch * a.chunk;
        const long long c1 = (ch == a.n_chunks - 1) ? a.Ng : min(a.Ng, c0 + a.chunk);
        return (c1 - c0 + TM - 1) / TM;
    };

    if (warp == 8) {
        // ================= MMA issuer (one thread) ===================================================
        if (lane == 0) {
            long long remaining[2] = {0, 0};
            for (int ch = 0; ch < a.n_chunks; ++ch) {
                const long long nt = tiles_in_chunk(ch);
                for (int s = 0; s < 2; ++s)
                    if (nt > first_of_slot[s]) remaining[s] += 3 * ((nt - first_of_slot[s] + stride - 1) / stride);
            }
            const uint32_t sb = smem_u32(smem);
            constexpr uint32_t id128 = idesc_f16(128, 128), id32 = idesc_f16(128, 32), id16 = idesc_f16(128, 16);
            int stage[2] = {0, 0};
            uint32_t par[2] = {0, 0};
            while (remaining[0] > 0 || remaining[1] > 0) {
#pragma unroll
                for (int s = 0; s < 2; ++s) {
                    if (remaining[s] <= 0 || !mbar_test_wait(&bars[s], par[s])) continue;
                    fence_after_sync();
                    const uint32_t tD = tmem + s * 256, tA = tD + 128;
                    if (stage[s] == 0) {
                        const uint64_t bh = smem_desc(sb + OFF_B1H, NM * 16, 128), bl = smem_desc(sb + OFF_B1L, NM * 16, 128);
                        mma_ts_f16(tD, tA, bh, id128, 0);
                        mma_ts_f16(tD, tA + 8, bh, id128, 1);
                        mma_ts_f16(tD, tA, bl, id128, 1);
                    } else if (stage[s] == 1) {
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks)
                            mma_ts_f16(tD, tA + ks * 8, smem_desc(sb + OFF_B2H + ks * 2 * NM * 16, NM * 16, 128), id128, ks > 0);
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks)
                            mma_ts_f16(tD, tA + 64 + ks * 8, smem_desc(sb + OFF_B2H + ks * 2 * NM * 16, NM * 16, 128), id128, 1);
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks)
                            mma_ts_f16(tD, tA + ks * 8, smem_desc(sb + OFF_B2L + ks * 2 * NM * 16, NM * 16, 128), id128, 1);
                    } else {
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks)
                            mma_ts_f16(tD, tA + ks * 8, smem_desc(sb + OFF_B3 + ks * 2 * 32 * 16, 32 * 16, 128), id32, ks > 0);
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks)
                            mma_ts_f16(tD, tA + 64 + ks * 8, smem_desc(sb + OFF_B3 + ks * 2 * 32 * 16, 32 * 16, 128), id16, 1);
                    }
                    mma_commit(&bars[2 + s]);
                    par[s] ^= 1;
                    stage[s] = stage[s] == 2 ? 0 : stage[s] + 1;
                    --remaining[s];
                }
            }
        }
    } else {
        // ================= row warps: slot = warp / 4, TMEM lane = 32 * (warp % 4) + lane ============
        const int slot = warp >> 2, q = warp & 3;
        const uint32_t tD = tmem + ((uint32_t)(q * 32) << 16) + slot * 256, tA = tD + 128;
        uint64_t* a_rdy = &bars[slot];
        uint64_t* d_rdy = &bars[2 + slot];
        uint32_t par_d = 0;
        const int r = q * 32 + lane;

        int ch = 0;
        long long tile = first_of_slot[slot];
        auto settle = [&]() -> bool {
            while (ch < a.n_chunks && tile >= tiles_in_chunk(ch)) { ++ch; tile = first_of_slot[slot]; }
            return ch < a.n_chunks;
        };
        // S0: numerators of this thread's grid point (label swap of eval_utils.py:64-67 applied) -> A1
        auto stage0 = [&](int chx, long long tilex, int (&z)[NH], bool& valid) {
            const long long c0 = (long long)chx * a.chunk;
            const long long c1 = (chx == a.n_chunks - 1) ? a.Ng : min(a.Ng, c0 + a.chunk);
            const long long g = c0 + tilex * TM + r;
            valid = g < c1;
#pragma unroll
            for (int i = 0; i < NH; ++i) z[i] = 0;
            if (valid) {
                const unsigned short* src = reinterpret_cast<const unsigned short*>(a.gridz + g * NH);
#pragma unroll
                for (int i = 0; i < NH / 2; ++i) {
                    const unsigned v = __ldg(src + i);
                    z[2 * i] = v & 0xff; z[2 * i + 1] = v >> 8;
                }
            }
            int zl = z[0];
#pragma unroll
            for (int i = 1; i < NH; ++i) if (i == label) zl = z[i];
            const int z0 = z[0];
#pragma unroll
            for (int i = 1; i < NH; ++i) if (i == label) z[i] = z0;
            z[0] = zl;
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

        float m1[kIt], m2[kIt];
#pragma unroll
        for (int k = 0; k < kIt; ++k) { m1[k] = -INFINITY; m2[k] = -INFINITY; }
        unsigned notconv = 0;
        bool bad = false;
        int zc[NH];
        bool vc = false;
        bool have = settle();
        if (have) stage0(ch, tile, zc, vc);
        while (have) {
            const int cur_ch = ch;
            mbar_wait(d_rdy, par_d); par_d ^= 1; fence_after_sync();              // D1
            act_to_a<false>(tD, tA, nullptr);
            tmem_wait_st(); fence_before_sync(); mbar_arrive(a_rdy);
            mbar_wait(d_rdy, par_d); par_d ^= 1; fence_after_sync();              // D2
            act_to_a<true>(tD, tA, b2s);
            tmem_wait_st(); fence_before_sync(); mbar_arrive(a_rdy);
            mbar_wait(d_rdy, par_d); par_d ^= 1; fence_after_sync();              // D3
            float ft[NH];
            {
                uint32_t v[32];
                tmem_ld32(tD, v);
                tmem_wait_ld();
#pragma unroll
                for (int i = 0; i < NH; ++i)
                    ft[i] = fmaf(__uint_as_float(v[i]) + __uint_as_float(v[16 + i]), kWSInv, b3s[i]);
            }
            // next tile of this slot: its layer-1 MMAs overlap the projection below
            tile += stride;
            have = settle();
            int zn[NH];
            bool vn = false;
            if (have) stage0(ch, tile, zn, vn);

            // ---- S3: projection with all candidate stop iterations (rhs_rows.cu kCertify epilogue) ----
            if (vc) {
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
                const float LW = (FW == -INFINITY) ? -INFINITY : lut_lo[zmax];
                const float perturb = lut_pt[zmax];
                float hi = ft[0] - lo[0], lw = ft[0];
#pragma unroll
                for (int j = 1; j < NH; ++j) { hi = fmaxf(hi, ft[j] - lo[j]); lw = fminf(lw, ft[j]); }
#pragma unroll
                for (int k = 0; k < kIt; ++k) {
                    if (k < a.w.max_iter) {
                        const float mu = (hi - lw) / 2.f + lw;
                        float e = 0.f;
#pragma unroll
                        for (int j = 0; j < NH; ++j) e += fmaxf(ft[j] - mu, lo[j]);
                        const float fy = fmaxf(fty - mu, loy), fw = fmaxf(FW - mu, LW);
                        const float hx = -fy + fw;                     // certify_lipschitz.py:41
                        m2[k] = fmaxf(m2[k], hx);
                        m1[k] = fmaxf(m1[k], hx + perturb);
                        if (!(fabsf(e) < a.w.tol)) notconv |= (1u << k);
                        if (e > 0.f) lw = mu;
                        if (e < 0.f) hi = mu;
                    }
                }
            }
            if (!have || ch != cur_ch) {
                // chunk finished for this slot: warp reduce and publish
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
#pragma unroll
            for (int i = 0; i < NH; ++i) zc[i] = zn[i];
            vc = vn;
        }
        if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(status, 1u);
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 8) tmem_dealloc(tmem, 512);
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
