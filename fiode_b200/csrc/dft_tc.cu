// fiode_b200/csrc/dft_tc.cu -- model.0 of the backbone on the sm_100a tensor cores:
//   per-frequency channel mixing (3 -> 32) + inverse real 2-D DFT 32x32 + bias + GroupSort
// (libs/ortho_conv/layers.py:72-86 `yfft = Q @ xfft; y = irfft2(yfft) + bias`, :127-131 GroupSort), the stage that was
// `irfft32_mix3_wide_kernel` (CUDA-core FFTs, 0.59 ms per 8192 images, issue-bound).  The inverse transform is
// evaluated as two dense real GEMMs on tcgen05 (fp16x2 split, see tc.cuh); only the K = 3 mixing, the hi/lo splits
// and one transposition stay on the CUDA cores.
//
//   column pass   T[co][a][v] = sum_u Y[co][u][v] w^(ua)            D1[(co,v) 128][(a, re/im) 64] = A1[128][(u, re/im) 64] . M1
//   row pass      y[co][a][b] = c2r_v(T[co][a][.]) / 1024           D2[(co,a) 128][b 32]          = A2[128][k 32]          . M2
//
// Columns v = 0 and v = 16 of a real signal's spectrum transform to REAL columns of T (the c2r step ignores their imaginary
// parts, as torch.fft.irfft2 and the FFT kernel do), so they travel as ONE complex column P = Y[.][0] + i Y[.][16]:
// 16 complex columns per channel, 8 channels x 16 columns = one 128-row MMA tile, and the row pass has exactly K = 32 real
// inputs per output row (17 real parts, 15 imaginary parts).
//
// A CTA owns 8 output channels (4 GroupSort pairs): its slice of the transposed spectrum Qt (104 KB) stays in shared
// memory for the whole launch, the DFT matrices (20 KB, fp16 hi/lo, K-major core matrices) beside it.  Three warpgroups
// work on three images at a time, each a self-contained pipeline (no cross-warpgroup synchronisation):
//   TMA bulk copy of the image's 3-channel spectrum (13 KB) -> mixing, thread = tile row (channel, column) ->
//   hi/lo split -> tcgen05.st (A operand in TMEM) -> 12 MMAs -> tcgen05.ld -> transposition through a warp-private 4 KB
//   smem tile -> split -> tcgen05.st -> 2 x 6 MMAs -> tcgen05.ld -> + bias, GroupSort (partner = lane ^ 16) ->
//   smem -> coalesced 512-byte stores.
// Every warp keeps one GroupSort pair (c, c + 16) of its image from the spectrum to the store.
#include "common.cuh"
#include "tc.cuh"
#include <cstdlib>

namespace fiode {
using namespace tc;

namespace {
constexpr int D_N = 32, D_H = 17, D_F = D_N * D_H;          // 544 frequencies, torch rfft2 layout phi = u * 17 + v
constexpr int D_WG = 3;                                      // warpgroups (images in flight) per CTA
constexpr int D_THREADS = D_WG * 128;
constexpr int D_CH = 8;                                      // channels per CTA
constexpr uint32_t kQsBytes = D_CH * 3 * D_F * 8;            // 104 448
constexpr uint32_t kM1Halves = 64 * 64, kM2Halves = 32 * 32;
constexpr uint32_t kConstBytes = (2 * kM1Halves + 2 * kM2Halves) * 2;      // 20 480: M1 hi, M1 lo, M2 hi, M2 lo
constexpr uint32_t kXsBytes = 3 * D_F * 8;                   // 13 056
constexpr uint32_t kSWords = 2 * 16 * 32 + 16;               // warp-private tile: [j 2][a' 16][k 32] floats, j = 1 block shifted by 16 words
constexpr uint32_t kSBytes = kSWords * 4;                    // 4 160
constexpr uint32_t kWgBytes = kXsBytes + 4 * kSBytes;        // 29 696
constexpr uint32_t kSmemBytes = kQsBytes + kConstBytes + D_WG * kWgBytes + 128;
constexpr float kWS = 64.f, kWSInv = 1.f / 64.f;             // DFT matrices are stored x 64 (their lo halves stay normal fp16)
constexpr float kS1 = 1.f / 2048.f;                          // D1 -> T / 32: undoes the x 64 and keeps |T| inside fp16 range

__device__ __forceinline__ void mbar_expect_tx_(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s_(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void named_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void tmem_ld32f(uint32_t addr, float* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]),
                   "=f"(v[8]), "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15]),
                   "=f"(v[16]), "=f"(v[17]), "=f"(v[18]), "=f"(v[19]), "=f"(v[20]), "=f"(v[21]), "=f"(v[22]), "=f"(v[23]),
                   "=f"(v[24]), "=f"(v[25]), "=f"(v[26]), "=f"(v[27]), "=f"(v[28]), "=f"(v[29]), "=f"(v[30]), "=f"(v[31])
                 : "r"(addr) : "memory");
}
__device__ __forceinline__ void tmem_st32_(uint32_t addr, const uint32_t* v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
                 "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                 ::"r"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                 "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
                 "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
                 "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
}
// hi / lo fp16 pair words of two fp32 values (element 0 in the low half)
__device__ __forceinline__ void split_pair(float x0, float x1, uint32_t& hi, uint32_t& lo) {
    const float2 x = make_float2(x0, x1);
    const __half2 h = __float22half2_rn(x);
    const __half2 l = __float22half2_rn(__ffma2_rn(__half22float2(h), make_float2(-1.f, -1.f), x));   // x - float(h), exact
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}
}  // namespace

// DFT matrices as tcgen05 B operands: K-major 8 x 16-byte core matrices, element (k, n) of an N-wide matrix at
// half index ((k >> 3) * N + n) * 8 + (k & 7); block order M1 hi | M1 lo | M2 hi | M2 lo, all scaled by 64.
//   M1[k = 2u + r][n = 32 h + 2 a' + r2], a = 16 h + a':  (pr, pi) of P[u] -> (re, im) of sum_u P[u] e^{+2 pi i u a / 32}
//   M2[k][b]: k < 16 real part of column v' = k, k >= 16 imaginary part of column v' = k - 16 (v' = 0: the real parts of
//             T[.][0] and T[.][16]);  c2r weights 1, (-1)^b, 2 cos, -2 sin, times 1/1024 * 32 (A2 holds T / 32)
__global__ void dft32_consts_kernel(__half* __restrict__ out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    double val;
    size_t idx_hi, idx_lo;
    if (e < 64 * 64) {
        const int k = e / 64, n = e % 64;
        const int u = k >> 1, r = k & 1, h = n >> 5, a = 16 * h + ((n & 31) >> 1), r2 = n & 1;
        double s, c;
        sincospi((double)((u * a) % 32) / 16.0, &s, &c);
        val = r == 0 ? (r2 == 0 ? c : s) : (r2 == 0 ? -s : c);
        const size_t off = ((size_t)(k >> 3) * 64 + n) * 8 + (k & 7);
        idx_hi = off; idx_lo = kM1Halves + off;
    } else if (e < 64 * 64 + 32 * 32) {
        const int f = e - 64 * 64, k = f / 32, b = f % 32;
        const int v = k & 15;
        double s, c;
        sincospi((double)((v * b) % 32) / 16.0, &s, &c);
        if (k < 16) val = v == 0 ? 1.0 : 2.0 * c;
        else val = v == 0 ? ((b & 1) ? -1.0 : 1.0) : -2.0 * s;
        val *= 1.0 / 32.0;
        const size_t off = ((size_t)(k >> 3) * 32 + b) * 8 + (k & 7);
        idx_hi = 2 * kM1Halves + off; idx_lo = 2 * kM1Halves + kM2Halves + off;
    } else return;
    const float x = (float)(val * (double)kWS);
    unsigned short h16, l16;
    split1(x, h16, l16);
    out[idx_hi] = __ushort_as_half(h16);
    out[idx_lo] = __ushort_as_half(l16);
}

__global__ void __launch_bounds__(D_THREADS, 1)
irfft32_mix3_tc_kernel(const float2* __restrict__ Xf, const float2* __restrict__ Qt, const __half* __restrict__ consts, int B,
                       const float* __restrict__ bias, int group_sort, float* __restrict__ y) {
    extern __shared__ __align__(1024) uint8_t smem[];
    float2* Qs = reinterpret_cast<float2*>(smem);                          // [slot 8][c 3][phi 544]
    uint8_t* cst = smem + kQsBytes;                                        // M1 hi | M1 lo | M2 hi | M2 lo
    uint8_t* wgs = cst + kConstBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(wgs + D_WG * kWgBytes);   // per warpgroup: x_full, d1_full, d2_full
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * D_WG);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int wg = warp >> 2, q = warp & 3, j = lane >> 4, i = lane & 15;
    const int cg = blockIdx.x & 3, slice = blockIdx.x >> 2, nslices = gridDim.x >> 2;
    const int co = 4 * cg + q + 16 * j;                                    // this lane's channel: GroupSort pair (c, c + 16) per warp

    if (warp == 0) tmem_alloc(tmem_slot, 512);
    if (t == 0) {
        for (int k = 0; k < 3 * D_WG; ++k) mbar_init(&bars[k], 1);
        mbar_init_fence();
    }
    // resident operands: the 8 channel rows of Qt (slot = 2 q + j) and the DFT matrices
    for (int e = t; e < (int)(kQsBytes / 16); e += D_THREADS) {
        const int slot = e / (3 * D_F / 2), rem = e - slot * (3 * D_F / 2);
        const int ch = 4 * cg + (slot >> 1) + 16 * (slot & 1);
        reinterpret_cast<float4*>(Qs)[e] = __ldg(reinterpret_cast<const float4*>(Qt + (size_t)ch * 3 * D_F) + rem);
    }
    for (int e = t; e < (int)(kConstBytes / 16); e += D_THREADS)
        reinterpret_cast<float4*>(cst)[e] = __ldg(reinterpret_cast<const float4*>(consts) + e);
    fence_proxy_async();                                                   // the tensor core reads the matrices through the async proxy
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    uint8_t* my = wgs + wg * kWgBytes;
    float2* Xs = reinterpret_cast<float2*>(my);                            // [c 3][phi 544] of the current image
    float* S = reinterpret_cast<float*>(my + kXsBytes + q * kSBytes);      // warp-private transposition / output tile
    uint64_t* x_full = &bars[3 * wg], *d1_full = &bars[3 * wg + 1], *d2_full = &bars[3 * wg + 2];
    const bool leader = (q == 0 && lane == 0);
    const int stride = nslices * D_WG;
    int img = slice + nslices * wg;
    if (leader && img < B) {
        mbar_expect_tx_(x_full, kXsBytes);
        bulk_g2s_(Xs, Xf + (size_t)img * 3 * D_F, kXsBytes, x_full);
    }
    // TMEM columns of this warpgroup: A [0, 64): A1 hi [0,32) lo [32,64); A2 tile h: hi [32h, 32h+16) lo [32h+16, 32h+32)
    //                                  D [64, 128): D1 64 columns; D2 tile h: [64 + 32h, 64 + 32h + 32)
    const uint32_t colA = wg * 128, colD = wg * 128 + 64;
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
    const uint32_t m1 = smem_u32(cst), m2 = m1 + 4 * kM1Halves;
    const float2* Qrow = Qs + (size_t)(2 * q + j) * 3 * D_F;
    const float bs = bias ? __ldg(bias + co) : 0.f;
    const uint32_t sj = j * (512 + 16);                                    // word offset of this lane's j block in S

    for (uint32_t it = 0; img < B; img += stride, ++it) {
        const uint32_t ph = it & 1;
        mbar_wait(x_full, ph);
        // ---- mixing: acc[u] = sum_c Q[co][c][u][v'] X[c][u][v'], v' = i; the Nyquist column v = 16 is split over the lanes ----
        float2 acc[D_N];
#pragma unroll
        for (int u = 0; u < D_N; ++u) acc[u] = make_float2(0.f, 0.f);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
#pragma unroll
            for (int u = 0; u < D_N; ++u) {
                const float2 qv = Qrow[c * D_F + u * D_H + i], xv = Xs[c * D_F + u * D_H + i];
                acc[u].x = fmaf(qv.x, xv.x, acc[u].x); acc[u].x = fmaf(-qv.y, xv.y, acc[u].x);
                acc[u].y = fmaf(qv.x, xv.y, acc[u].y); acc[u].y = fmaf(qv.y, xv.x, acc[u].y);
            }
        }
        {
            float2 n0 = make_float2(0.f, 0.f), n1 = make_float2(0.f, 0.f);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const int p0 = c * D_F + (2 * i) * D_H + 16, p1 = p0 + D_H;
                const float2 qa = Qrow[p0], xa = Xs[p0], qb = Qrow[p1], xb = Xs[p1];
                n0.x = fmaf(qa.x, xa.x, n0.x); n0.x = fmaf(-qa.y, xa.y, n0.x); n0.y = fmaf(qa.x, xa.y, n0.y); n0.y = fmaf(qa.y, xa.x, n0.y);
                n1.x = fmaf(qb.x, xb.x, n1.x); n1.x = fmaf(-qb.y, xb.y, n1.x); n1.y = fmaf(qb.x, xb.y, n1.y); n1.y = fmaf(qb.y, xb.x, n1.y);
            }
            // S doubles as the exchange buffer: [j][u 32] complex
            *reinterpret_cast<float4*>(S + j * 64 + 4 * i) = make_float4(n0.x, n0.y, n1.x, n1.y);
            __syncwarp();
            if (i == 0) {                       // packed column: P[u] = Y[u][0] + i Y[u][16]
#pragma unroll
                for (int u = 0; u < D_N; u += 2) {
                    const float4 w = *reinterpret_cast<const float4*>(S + j * 64 + 2 * u);
                    acc[u].x -= w.y; acc[u].y += w.x; acc[u + 1].x -= w.w; acc[u + 1].y += w.z;
                }
            }
            __syncwarp();
        }
        // ---- A1: hi / lo fp16 of (re, im) pairs, k = 2u + (re | im) ----
        {
            uint32_t hi[32], lo[32];
#pragma unroll
            for (int u = 0; u < D_N; ++u) split_pair(acc[u].x, acc[u].y, hi[u], lo[u]);
            tmem_st32_(lane_base + colA, hi);
            tmem_st32_(lane_base + colA + 32, lo);
        }
        tmem_wait_st();
        fence_before_sync();
        named_sync(1 + wg, 128);
        if (leader) {
            fence_after_sync();
            const uint32_t idesc = idesc_f16(128, 64);
            const uint64_t bh = smem_desc(m1, 64 * 16, 128), bl = smem_desc(m1 + 2 * kM1Halves, 64 * 16, 128);
            constexpr uint64_t step = (uint64_t)(2 * 64 * 16) >> 4;
            mma_ts_f16_x4(tmem + colD, tmem + colA, bh, step, idesc, 0);            // hi . hi
            mma_ts_f16_x4(tmem + colD, tmem + colA + 32, bh, step, idesc, 1);       // lo . hi
            mma_ts_f16_x4(tmem + colD, tmem + colA, bl, step, idesc, 1);            // hi . lo
            mma_commit(d1_full);
            const int nxt = img + stride;                                          // every lane has finished reading Xs
            if (nxt < B) {
                mbar_expect_tx_(x_full, kXsBytes);
                bulk_g2s_(Xs, Xf + (size_t)nxt * 3 * D_F, kXsBytes, x_full);
            }
        }
        mbar_wait(d1_full, ph);
        fence_after_sync();
        // ---- transposition (channel, column) x row  ->  (channel, row) x column, half of the rows at a time ----
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float d[32];
            tmem_ld32f(lane_base + colD + h * 32, d);
            tmem_wait_ld();
            // d[2a'] = Re T[16h + a'][v'], d[2a' + 1] = Im; S[j][a'][k ^ swizzle(a')], k = v' (re), 16 + v' (im)
#pragma unroll
            for (int a = 0; a < 16; ++a) {
                float* row = S + sj + a * 32;
                row[i ^ ((a & 7) << 2)] = d[2 * a] * kS1;
                row[(16 + i) ^ ((a & 7) << 2)] = d[2 * a + 1] * kS1;
            }
            __syncwarp();
            {   // this lane's row (channel j, a' = i): 32 inputs of the row pass
                const float* row = S + sj + i * 32;
                const int xi = (i & 7) << 2;
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const float4 w = *reinterpret_cast<const float4*>(row + ((c << 2) ^ xi));
                    split_pair(w.x, w.y, hi[2 * c], lo[2 * c]);
                    split_pair(w.z, w.w, hi[2 * c + 1], lo[2 * c + 1]);
                }
                tmem_st16(lane_base + colA + h * 32, hi);
                tmem_st16(lane_base + colA + h * 32 + 16, lo);
            }
            __syncwarp();
        }
        tmem_wait_st();
        fence_before_sync();
        named_sync(1 + wg, 128);
        if (leader) {
            fence_after_sync();
            const uint32_t idesc = idesc_f16(128, 32);
            const uint64_t bh = smem_desc(m2, 32 * 16, 128), bl = smem_desc(m2 + 2 * kM2Halves, 32 * 16, 128);
            constexpr uint64_t step = (uint64_t)(2 * 32 * 16) >> 4;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t tD = tmem + colD + h * 32, tA = tmem + colA + h * 32;
                mma_ts_f16(tD, tA, bh, idesc, 0);             mma_ts_f16(tD, tA + 8, bh + step, idesc, 1);         // hi . hi
                mma_ts_f16(tD, tA + 16, bh, idesc, 1);        mma_ts_f16(tD, tA + 24, bh + step, idesc, 1);        // lo . hi
                mma_ts_f16(tD, tA, bl, idesc, 1);             mma_ts_f16(tD, tA + 8, bl + step, idesc, 1);         // hi . lo
            }
            mma_commit(d2_full);
        }
        mbar_wait(d2_full, ph);
        fence_after_sync();
        // ---- epilogue: / 64, + bias, GroupSort with the partner lane, smem, coalesced stores ----
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float d[32];
            tmem_ld32f(lane_base + colD + h * 32, d);
            tmem_wait_ld();
            float* row = S + sj + i * 32;
            const int xi = (i & 7) << 2;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                float v[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float u = fmaf(d[4 * c + e], kWSInv, bs);
                    const float w = __shfl_xor_sync(0xffffffffu, u, 16);
                    v[e] = group_sort ? (j ? fminf(u, w) : fmaxf(u, w)) : u;
                }
                *reinterpret_cast<float4*>(row + ((c << 2) ^ xi)) = make_float4(v[0], v[1], v[2], v[3]);
            }
            __syncwarp();
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                const int R = 4 * s + (lane >> 3), jj = R >> 4, a = R & 15, c = lane & 7;
                const float4 w = *reinterpret_cast<const float4*>(S + jj * (512 + 16) + a * 32 + ((c ^ (a & 7)) << 2));
                const int ch = 4 * cg + q + 16 * jj;
                *reinterpret_cast<float4*>(y + (((size_t)img * 32 + ch) * 32 + 16 * h + a) * 32 + 4 * c) = w;
            }
            __syncwarp();
        }
        fence_before_sync();                 // this iteration's tcgen05.ld are ordered before the next iteration's st / MMAs
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}


// =====================================================================================================================
// Version 2: warp-specialised.  The first kernel above delivers every mixing operand through shared memory (2 x 96 complex
// values per lane and unit: 1 536 of its 2 900 shared-memory wavefronts per unit, profiles/r02_dft_tc_ncu.md) and runs one
// serial pipeline per warpgroup.  Here
//   8 mixer warps   thread = two (u, v') frequency pairs over ALL 8 channels of the CTA: its 2 x 24 complex entries of Q
//                   never move (pair A in registers, pair B in 48 TMEM columns of its own lane); per image it loads 6
//                   complex X values, forms the 16 mixed values, splits them and writes the A operand of the column pass
//                   into a shared-memory ring in the tcgen05 K-major core-matrix layout (conflict-free 4-byte stores)
//   2 Nyquist warps the v = 16 column (32 frequencies x 8 channels), handed to the lanes that own v' = 0
//   8 tail warps    two warpgroups alternating images: D1 (TMEM) -> transposition -> A2 (TMEM) -> D2 -> GroupSort -> stores
//   1 MMA thread    both GEMMs of both tail warpgroups, issued as their operands become ready (non-blocking polling)
//   1 copy thread   TMA bulk copies of the 3-channel input spectra into a 3-deep ring
// connected by mbarrier rings only (no CTA-wide barrier in the loop); setmaxnreg moves registers to the mixers.
// =====================================================================================================================
namespace {
constexpr int V_THREADS = 640;
constexpr int V_NX = 2, V_NA = 2;
constexpr int V_XR = 20;                                        // row stride of the input spectrum [c][u][V_XR]: 4 u x 4 v' lanes of a half-warp hit 16 different 8-byte banks
constexpr uint32_t kXs2Bytes = 3 * 32 * V_XR * 8;               // 15 360
constexpr int kRegsTail = 88, kRegsMix = 128, kRegsAux = 48;
static_assert(256 * kRegsTail + 256 * kRegsMix + 128 * kRegsAux <= 96 * V_THREADS, "setmaxnreg.inc would block");
constexpr uint32_t kA1Bytes = 2 * 128 * 64 * 2;                 // hi + lo blocks of one A1 tile: 32 768
constexpr uint32_t kS2Words = 2 * 32 * 32 + 16;                 // [j 2][a 32][k 32] floats, j = 1 block shifted by 16 words
constexpr uint32_t kS2Bytes = 8320;
static_assert(kS2Words * 4 <= kS2Bytes, "S tile");
constexpr uint32_t V_OFF_CONST = 0;
constexpr uint32_t V_OFF_A1 = V_OFF_CONST + kConstBytes;                       //  20 480
constexpr uint32_t V_OFF_XS = V_OFF_A1 + V_NA * kA1Bytes;                      //  86 016
constexpr uint32_t V_OFF_Y16 = V_OFF_XS + V_NX * kXs2Bytes;                     // 125 184
constexpr uint32_t V_OFF_Q16 = V_OFF_Y16 + 2 * 2048;                           // 129 280
constexpr uint32_t V_OFF_S = V_OFF_Q16 + 8 * 3 * 32 * 8;                       // 135 424
constexpr uint32_t V_OFF_O = V_OFF_S + 8 * kS2Bytes;                           // output staging tiles, 4 KB per tail warp
constexpr uint32_t V_OFF_BARS = V_OFF_O + 8 * 4096;
constexpr uint32_t V_SMEM = V_OFF_BARS + 32 * 8 + 16;
static_assert(V_NX <= 3, "barrier map below reserves three X slots");
enum { BX_FULL = 0, BX_EMPTY = 3, BA_FULL = 6, BA_EMPTY = 8, BY_FULL = 10, BY_EMPTY = 12, BD1_FULL = 14, BD1_FREE = 16,
       BA2_FULL = 18, BD2_FULL = 20, B_COUNT = 22 };

__device__ __forceinline__ void tmem_ld16f(uint32_t addr, float* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]),
                   "=f"(v[8]), "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15])
                 : "r"(addr) : "memory");
}
// mbarrier.arrive that releases a buffer this thread has READ: the arrive must not be issued before the loads have returned.
// Issue is in order, so it is enough that instructions consuming the loaded values precede it; the (unused) operands tie the
// arrive to their results, which keeps the compiler from hoisting it above them.
__device__ __forceinline__ void mbar_arrive_after(uint64_t* bar, float d0, float d1, float d2, float d3) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)), "f"(d0), "f"(d1), "f"(d2), "f"(d3) : "memory");
}
__device__ __forceinline__ void cmac(float2& acc, float2 q, float2 x) {
    acc.x = fmaf(q.x, x.x, acc.x); acc.x = fmaf(-q.y, x.y, acc.x);
    acc.y = fmaf(q.x, x.y, acc.y); acc.y = fmaf(q.y, x.x, acc.y);
}
}  // namespace

__global__ void __launch_bounds__(V_THREADS, 1)
irfft32_mix3_tc2_kernel(const float2* __restrict__ Xf, const float2* __restrict__ Qt, const __half* __restrict__ consts, int B,
                        const float* __restrict__ bias, int group_sort, float* __restrict__ y) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + V_OFF_BARS);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 32);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int cg = blockIdx.x & 3, slice = blockIdx.x >> 2, nslices = gridDim.x >> 2;
    const int N = slice < B ? (B - slice + nslices - 1) / nslices : 0;       // units (images) of this CTA: img = slice + nslices * n
    auto chan = [&](int slot) { return 4 * cg + (slot >> 1) + 16 * (slot & 1); };   // slot = 2 q + j: GroupSort pairs (c, c + 16)

    if (warp == 16) tmem_alloc(tmem_slot, 512);
    if (t == 0) {
        for (int k = 0; k < V_NX; ++k) { mbar_init(&bars[BX_FULL + k], 1); mbar_init(&bars[BX_EMPTY + k], 256 + 64); }
        for (int k = 0; k < 2; ++k) {
            mbar_init(&bars[BA_FULL + k], 256); mbar_init(&bars[BA_EMPTY + k], 1);
            mbar_init(&bars[BY_FULL + k], 64); mbar_init(&bars[BY_EMPTY + k], 32);
            mbar_init(&bars[BD1_FULL + k], 1); mbar_init(&bars[BD1_FREE + k], 128);
            mbar_init(&bars[BA2_FULL + k], 128); mbar_init(&bars[BD2_FULL + k], 1);
        }
        mbar_init_fence();
    }
    for (int e = t; e < (int)(kConstBytes / 16); e += V_THREADS)
        reinterpret_cast<float4*>(smem + V_OFF_CONST)[e] = __ldg(reinterpret_cast<const float4*>(consts) + e);
    for (int e = t; e < 8 * 3 * 32; e += V_THREADS) {           // Nyquist column of the CTA's Q slice: [slot][c][u]
        const int slot = e / 96, c = (e / 32) % 3, u = e & 31;
        reinterpret_cast<float2*>(smem + V_OFF_Q16)[e] = __ldg(Qt + ((size_t)chan(slot) * 3 + c) * D_F + u * D_H + 16);
    }
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp < 8) {
        // ============================ tails: two warpgroups alternating units ======================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsTail));
        const int tw = warp >> 2, q = warp & 3, j = lane >> 4, i = lane & 15;
        const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16) + tw * 192;    // A2 [0,64) D1 [64,128) D2 [128,192)
        float* S = reinterpret_cast<float*>(smem + V_OFF_S + warp * kS2Bytes);
        const uint32_t sj = j * (1024 + 16);
        const float bs0 = bias ? __ldg(bias + chan(2 * q)) : 0.f, bs1 = bias ? __ldg(bias + chan(2 * q + 1)) : 0.f;
        uint64_t* d1_full = &bars[BD1_FULL + tw], *d1_free = &bars[BD1_FREE + tw], *a2_full = &bars[BA2_FULL + tw], *d2_full = &bars[BD2_FULL + tw];
        float* O = reinterpret_cast<float*>(smem + V_OFF_O + warp * 4096);       // output staging: [j 2][a 32][16 floats], 16-byte chunks swizzled
        // part A of a unit: D1 row (slot 2q + j, v' = i), columns 32 h + 2 a' + (re | im), a = 16 h + a'  ->  S[j][a][k], k = v' | 16 + v'
        auto part_a = [&](uint32_t ph) {
            mbar_wait(d1_full, ph);
            fence_after_sync();
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float d[32];
                tmem_ld32f(lane_base + 64 + h * 32, d);
                tmem_wait_ld();
                if (h == 1) { fence_before_sync(); mbar_arrive(d1_free); }      // D1 is in registers: the next column pass may overwrite it
#pragma unroll
                for (int a = 0; a < 16; ++a) {
                    float* row = S + sj + (16 * h + a) * 32;
                    row[i ^ ((a & 7) << 2)] = d[2 * a] * kS1;
                    row[(16 + i) ^ ((a & 7) << 2)] = d[2 * a + 1] * kS1;
                }
            }
            __syncwarp();
        };
        if (tw < N) part_a(0);
        for (int n = tw, k = 0; n < N; n += 2, ++k) {
            const uint32_t ph = k & 1;
            const int img = slice + nslices * n;
            // part B: rows (channel j', a = lane) of the row pass: tile A = channel slot 2q, tile B = its GroupSort partner 2q + 1
#pragma unroll
            for (int jj = 0; jj < 2; ++jj) {
                const float* row = S + jj * (1024 + 16) + lane * 32;
                const int xi = (lane & 7) << 2;
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const float4 w = *reinterpret_cast<const float4*>(row + ((c << 2) ^ xi));
                    split_pair(w.x, w.y, hi[2 * c], lo[2 * c]);
                    split_pair(w.z, w.w, hi[2 * c + 1], lo[2 * c + 1]);
                }
                tmem_st16(lane_base + jj * 32, hi);
                tmem_st16(lane_base + jj * 32 + 16, lo);
            }
            tmem_wait_st();
            fence_before_sync();
            mbar_arrive(a2_full);
            __syncwarp();                         // every lane has read its S rows
            if (n + 2 < N) part_a((uint32_t)((k + 1) & 1));       // the next unit's column pass goes to S while this unit's row pass runs
            mbar_wait(d2_full, ph);
            fence_after_sync();
            // epilogue: / 64, + bias, GroupSort of (tile A, tile B) in registers; 16 columns at a time through O, 64-byte row pieces out
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                float da[16], db[16];
                tmem_ld16f(lane_base + 128 + cc * 16, da);
                tmem_ld16f(lane_base + 160 + cc * 16, db);
                tmem_wait_ld();
                const int sw = (lane >> 1) & 3;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    float hi4[4], lo4[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float u = fmaf(da[4 * c + e], kWSInv, bs0), w = fmaf(db[4 * c + e], kWSInv, bs1);
                        hi4[e] = group_sort ? fmaxf(u, w) : u;
                        lo4[e] = group_sort ? fminf(u, w) : w;
                    }
                    *reinterpret_cast<float4*>(O + lane * 16 + ((c ^ sw) << 2)) = make_float4(hi4[0], hi4[1], hi4[2], hi4[3]);
                    *reinterpret_cast<float4*>(O + 512 + lane * 16 + ((c ^ sw) << 2)) = make_float4(lo4[0], lo4[1], lo4[2], lo4[3]);
                }
                __syncwarp();
#pragma unroll
                for (int s_ = 0; s_ < 8; ++s_) {
                    const int R = 8 * s_ + (lane >> 2), jj = R >> 5, a = R & 31, c = lane & 3;
                    const float4 w = *reinterpret_cast<const float4*>(O + jj * 512 + a * 16 + ((c ^ ((a >> 1) & 3)) << 2));
                    *reinterpret_cast<float4*>(y + (((size_t)img * 32 + chan(2 * q + jj)) * 32 + a) * 32 + 16 * cc + 4 * c) = w;
                }
                __syncwarp();
            }
            fence_before_sync();
        }
    } else if (warp < 16) {
        // ============================ mixers ==================================================================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsMix));
        const int mw = warp - 8, mg = mw >> 2, ul = lane & 3, vl = lane >> 2;
        const int u = 4 * mw + ul;
        const uint32_t tq = tmem + ((uint32_t)((warp & 3) * 32) << 16) + 384 + mg * 48;     // pair B's Q: column c * 16 + slot * 2 + (re | im)
        float2 qa[3][8];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            uint32_t qb[16];
#pragma unroll
            for (int slot = 0; slot < 8; ++slot) {
                const float2* src = Qt + ((size_t)chan(slot) * 3 + c) * D_F + u * D_H;
                qa[c][slot] = __ldg(src + vl);
                const float2 b = __ldg(src + 8 + vl);
                qb[2 * slot] = __float_as_uint(b.x); qb[2 * slot + 1] = __float_as_uint(b.y);
            }
            tmem_st16(tq + c * 16, qb);
        }
        tmem_wait_st();
        // word offset of this lane's (re, im) pair inside an A1 block: core matrix (u >> 2, row >> 3), row & 7, k & 7 = 2 (u & 3)
        const uint32_t wofs = (uint32_t)(u >> 2) * 512 + (uint32_t)vl * 4 + (uint32_t)(u & 3);      // + row0 * 4, row0 = 16 slot (+ 8 for pair B)
        for (int n = 0; n < N; ++n) {
            const int xb = n % V_NX, ab = n & 1;
            mbar_wait(&bars[BX_FULL + xb], (uint32_t)((n / V_NX) & 1));
            const float2* Xs = reinterpret_cast<const float2*>(smem + V_OFF_XS + xb * kXs2Bytes);
            float2 xa[3], xbv[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) { xa[c] = Xs[(c * 32 + u) * V_XR + vl]; xbv[c] = Xs[(c * 32 + u) * V_XR + 8 + vl]; }
            float2 acc[8];
#pragma unroll
            for (int slot = 0; slot < 8; ++slot) {
                acc[slot] = make_float2(0.f, 0.f);
#pragma unroll
                for (int c = 0; c < 3; ++c) cmac(acc[slot], qa[c][slot], xa[c]);
            }
            if (vl == 0) {                         // packed column: P[u] = Y[u][0] + i Y[u][16]
                mbar_wait(&bars[BY_FULL + ab], (uint32_t)((n >> 1) & 1));
                const float2* Y16 = reinterpret_cast<const float2*>(smem + V_OFF_Y16 + ab * 2048);
#pragma unroll
                for (int slot = 0; slot < 8; ++slot) { const float2 w = Y16[slot * 32 + u]; acc[slot].x -= w.y; acc[slot].y += w.x; }
                mbar_arrive_after(&bars[BY_EMPTY + ab], acc[0].x + acc[1].x, acc[2].x + acc[3].x, acc[4].x + acc[5].x, acc[6].x + acc[7].x);
            }
            mbar_wait(&bars[BA_EMPTY + ab], (uint32_t)(((n >> 1) & 1) ^ 1));
            uint32_t* A1 = reinterpret_cast<uint32_t*>(smem + V_OFF_A1 + ab * kA1Bytes);
#pragma unroll
            for (int slot = 0; slot < 8; ++slot) {
                uint32_t hi, lo;
                split_pair(acc[slot].x, acc[slot].y, hi, lo);
                A1[wofs + slot * 64] = hi;
                A1[4096 + wofs + slot * 64] = lo;
            }
#pragma unroll
            for (int slot = 0; slot < 8; ++slot) acc[slot] = make_float2(0.f, 0.f);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                float qb[16];
                tmem_ld16f(tq + c * 16, qb);
                tmem_wait_ld();
#pragma unroll
                for (int slot = 0; slot < 8; ++slot) cmac(acc[slot], make_float2(qb[2 * slot], qb[2 * slot + 1]), xbv[c]);
            }
            mbar_arrive_after(&bars[BX_EMPTY + xb], acc[0].x, acc[0].y, acc[7].x, acc[7].y);     // xa / xbv of all three channels went into these
#pragma unroll
            for (int slot = 0; slot < 8; ++slot) {
                uint32_t hi, lo;
                split_pair(acc[slot].x, acc[slot].y, hi, lo);
                A1[wofs + slot * 64 + 32] = hi;
                A1[4096 + wofs + slot * 64 + 32] = lo;
            }
            fence_proxy_async();                   // generic stores -> the tensor core's (async proxy) operand reads
            mbar_arrive(&bars[BA_FULL + ab]);
        }
    } else {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsAux));
        if (warp == 16) {
            // ============================ MMA issuer ==========================================================
            if (lane == 0) {
                const uint32_t sb = smem_u32(smem);
                const uint32_t m1 = sb + V_OFF_CONST, m2 = m1 + 4 * kM1Halves;
                constexpr uint32_t id64 = idesc_f16(128, 64), id32 = idesc_f16(128, 32);
                int n1 = 0, n2 = 0;
                while (n2 < N) {
                    bool progressed = false;
                    if (n1 < N) {
                        const int tw = n1 & 1, ab = n1 & 1, k = n1 >> 1;
                        if (mbar_test_wait(&bars[BA_FULL + ab], (uint32_t)(k & 1)) && mbar_test_wait(&bars[BD1_FREE + tw], (uint32_t)((k & 1) ^ 1))) {
                            fence_after_sync();
                            const uint32_t tD = tmem + tw * 192 + 64;
                            const uint32_t a1 = sb + V_OFF_A1 + ab * kA1Bytes;
#pragma unroll
                            for (int pass = 0; pass < 3; ++pass) {          // hi . hi, lo . hi, hi . lo
                                const uint32_t ablk = a1 + (pass == 1 ? 16384 : 0), bblk = m1 + (pass == 2 ? 2 * kM1Halves : 0);
#pragma unroll
                                for (int sl = 0; sl < 4; ++sl)
                                    mma_ss_f16(tD, smem_desc(ablk + sl * 4096, 128 * 16, 128), smem_desc(bblk + sl * 2048, 64 * 16, 128), id64,
                                               (pass | sl) != 0);
                            }
                            mma_commit(&bars[BA_EMPTY + ab]);
                            mma_commit(&bars[BD1_FULL + tw]);
                            ++n1;
                            progressed = true;
                        }
                    }
                    if (n2 < n1) {
                        const int tw = n2 & 1, k = n2 >> 1;
                        if (mbar_test_wait(&bars[BA2_FULL + tw], (uint32_t)(k & 1))) {
                            fence_after_sync();
                            const uint64_t bh = smem_desc(m2, 32 * 16, 128), bl = smem_desc(m2 + 2 * kM2Halves, 32 * 16, 128);
                            constexpr uint64_t step = (uint64_t)(2 * 32 * 16) >> 4;
#pragma unroll
                            for (int jj = 0; jj < 2; ++jj) {
                                const uint32_t tD = tmem + tw * 192 + 128 + jj * 32, tA = tmem + tw * 192 + jj * 32;
                                mma_ts_f16(tD, tA, bh, id32, 0);             mma_ts_f16(tD, tA + 8, bh + step, id32, 1);
                                mma_ts_f16(tD, tA + 16, bh, id32, 1);        mma_ts_f16(tD, tA + 24, bh + step, id32, 1);
                                mma_ts_f16(tD, tA, bl, id32, 1);             mma_ts_f16(tD, tA + 8, bl + step, id32, 1);
                            }
                            mma_commit(&bars[BD2_FULL + tw]);
                            ++n2;
                            progressed = true;
                        }
                    }
                    if (!progressed) __nanosleep(40);      // the polling loop shares a scheduler with two mixer and two tail warps
                }
            }
        } else if (warp == 17) {
            // ============================ input spectra: TMA bulk copies ======================================
            if (lane == 0) {
                for (int n = 0; n < N; ++n) {
                    const int xb = n % V_NX;
                    mbar_wait(&bars[BX_EMPTY + xb], (uint32_t)(((n / V_NX) & 1) ^ 1));
                    mbar_expect_tx_(&bars[BX_FULL + xb], kXs2Bytes);
                    bulk_g2s_(smem + V_OFF_XS + xb * kXs2Bytes, Xf + (size_t)(slice + nslices * n) * 3 * 32 * V_XR, kXs2Bytes, &bars[BX_FULL + xb]);
                }
            }
        } else {
            // ============================ Nyquist column v = 16: thread = (u, 4 channels) =========================
            const int nt = (warp - 18) * 32 + lane, u = nt & 31, half = nt >> 5;
            const float2* Q16 = reinterpret_cast<const float2*>(smem + V_OFF_Q16);
            for (int n = 0; n < N; ++n) {
                const int xb = n % V_NX, yb = n & 1;
                mbar_wait(&bars[BX_FULL + xb], (uint32_t)((n / V_NX) & 1));
                const float2* Xs = reinterpret_cast<const float2*>(smem + V_OFF_XS + xb * kXs2Bytes);
                float2 x16[3];
#pragma unroll
                for (int c = 0; c < 3; ++c) x16[c] = Xs[(c * 32 + u) * V_XR + 16];
                float2 acc[4];
#pragma unroll
                for (int s4 = 0; s4 < 4; ++s4) {
                    acc[s4] = make_float2(0.f, 0.f);
#pragma unroll
                    for (int c = 0; c < 3; ++c) cmac(acc[s4], Q16[((4 * half + s4) * 3 + c) * 32 + u], x16[c]);
                }
                mbar_arrive_after(&bars[BX_EMPTY + xb], acc[0].x, acc[0].y, acc[3].x, acc[3].y);
                mbar_wait(&bars[BY_EMPTY + yb], (uint32_t)(((n >> 1) & 1) ^ 1));
                float2* Y16 = reinterpret_cast<float2*>(smem + V_OFF_Y16 + yb * 2048);
#pragma unroll
                for (int s4 = 0; s4 < 4; ++s4) Y16[(4 * half + s4) * 32 + u] = acc[s4];
                mbar_arrive(&bars[BY_FULL + yb]);
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 16) tmem_dealloc(tmem, 512);
}

size_t dft32_consts_bytes() { return kConstBytes; }
int dft32_tc2_xrow() { return V_XR; }
int launch_dft32_consts(void* out, cudaStream_t st) {
    dft32_consts_kernel<<<(64 * 64 + 32 * 32 + 255) / 256, 256, 0, st>>>(reinterpret_cast<__half*>(out));
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
// Xf [b][3][544] (torch rfft2 layout; the warp-specialised kernel takes rows padded to dft32_tc2_xrow() = 20), Qt [32][3][544] (transposed spectrum), consts from launch_dft32_consts, y [b][32][32][32]
int launch_irfft32_mix3_tc(const float2* Xf, const float2* Qt, const void* consts, int B, const float* bias, int gs, float* y,
                           cudaStream_t st) {
    static std::atomic<bool> done[kMaxDevices];
    if (first_use_on_device(done))
        FIODE_CUDA(cudaFuncSetAttribute(irfft32_mix3_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    int slices = sm_count() / 4;
    if (slices > B) slices = B;
    if (slices < 1) slices = 1;
    irfft32_mix3_tc_kernel<<<4 * slices, D_THREADS, kSmemBytes, st>>>(Xf, Qt, reinterpret_cast<const __half*>(consts), B, bias, gs, y);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
int launch_irfft32_mix3_tc2(const float2* Xf, const float2* Qt, const void* consts, int B, const float* bias, int gs, float* y,
                            cudaStream_t st) {
    static std::atomic<bool> done[kMaxDevices];
    if (first_use_on_device(done))
        FIODE_CUDA(cudaFuncSetAttribute(irfft32_mix3_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)V_SMEM));
    int slices = sm_count() / 4;
    if (slices > B) slices = B;
    if (slices < 1) slices = 1;
    irfft32_mix3_tc2_kernel<<<4 * slices, V_THREADS, V_SMEM, st>>>(Xf, Qt, reinterpret_cast<const __half*>(consts), B, bias, gs, y);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
}  // namespace fiode
