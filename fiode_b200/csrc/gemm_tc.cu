// fiode_b200/csrc/gemm_tc.cu -- the backbone's dense contractions on the sm_100a tensor cores.
//
//   D[m][n] = sum_k A[m][k] * W[n][k] + bias[n]   (+ GroupSort over the channel halves, layers.py:127-131)
//
// used for  CayleyLinear.forward  F.linear(X, Q, b)          (libs/ortho_conv/layers.py:105; 4096->512, 512->512)
// and for the two stride-2 CayleyConv layers, which are real orthogonal 1x1 convolutions on the space-to-depth
// tensor (layers.py:24-48 + SURVEY App. C-1; 128->32 @16x16 and 256->64 @8x8).
//
// Arithmetic: fp16x2 split (tc.cuh) -- an FP32 emulation with three kind::f16 MMAs per product and fp32
// accumulation in TMEM.  Structure (one persistent CTA per SM, or two when the tile is narrow):
//   8 loader warps   two warpgroups; warpgroup p owns the k chunks (64 wide) of parity p: cp.async of the raw fp32
//                    chunk (row-major, or gathered through the space-to-depth map) one own chunk ahead, then
//                    smem -> hi/lo fp16 -> tcgen05.st into A buffer p (thread = output row = TMEM lane)
//   1 copy thread    cp.async.bulk of the pre-split, pre-tiled weight chunk (hi + lo) into smem stage p
//   1 MMA thread     12 MMAs per chunk (4 k slabs x {hi.hi, lo.hi, hi.lo}); tcgen05.commit frees stage p
//   8 sum warps      add each 4-chunk accumulator segment to a register-resident sum (see "Promotion"), then the
//                    epilogue: x 1/64, + bias, GroupSort pairs, store
#include "common.cuh"
#include "tc.cuh"
#include <cstdlib>

namespace fiode {
using namespace tc;

constexpr int GT_THREADS = 576;   // warps 0-7 loaders, 8 MMA, 9 weight copies, 10-17 accumulate / epilogue
constexpr int GT_KC = 64;         // k chunk
constexpr float kGtWS = 64.f, kGtWSInv = 1.f / 64.f;

enum { kLoadRow = 0, kLoadS2d = 1 };

struct GemmTcArgs {
    const float* A; long long lda;   // kLoadRow: A (M, K) row-major
    int cin, n;                       // kLoadS2d: x (B, cin, n, n); M = B * (n/2)^2, K = 4 * cin
    long long M; int N, K, NT;        // NT = columns per CTA tile (multiple of 32, <= 256)
    const __half* Wp;                 // packed weights, see pack_w_kernel
    const float* bias; int group_sort;
    float* out; long long ldo;        // kLoadRow: out (M, N) row-major; kLoadS2d: out (B, N, n/2, n/2)
    int nbatch; long long sA, sW, sO; // kLoadRow only: nbatch independent problems (A + f sA, Wp + f sW, out + f sO)
};

// original output column of local column l of n-tile j.  With GroupSort a tile holds NT/2 columns of the first
// channel half followed by their partners of the second half, so that a pair never straddles two CTAs.
__host__ __device__ __forceinline__ int gt_orig_col(int j, int l, int N, int NT, int gs) {
    if (!gs) return j * NT + l;
    const int h = NT / 2;
    return l < h ? j * h + l : N / 2 + j * h + (l - h);
}

// W (N, K) fp32 row-major -> per (n-tile j, k-chunk c) one contiguous block {hi[8][NT][8 halves], lo[8][NT][8 halves]}
// (K-major 8x16-byte core matrices, the layout tcgen05.mma reads from shared memory), scaled by 64.
__global__ void pack_w_kernel(const float* __restrict__ W0, int N, int K, int NT, int gs, __half* __restrict__ Wp0, int nbatch) {
    const long long per = (long long)N * K, total = per * nbatch;
    const int kcn = K / GT_KC;
    for (long long e0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; e0 < total; e0 += (long long)gridDim.x * blockDim.x) {
        const long long f = e0 / per, e = e0 - f * per;
        const float* W = W0 + f * per;
        __half* Wp = Wp0 + f * per * 2;
        const int k = (int)(e % K);
        const long long jl = e / K;           // tile-major local column index
        const int j = (int)(jl / NT), l = (int)(jl % NT);
        const int n = gt_orig_col(j, l, N, NT, gs);
        unsigned short h, lo;
        split1(W[(size_t)n * K + k] * kGtWS, h, lo);
        const int c = k / GT_KC, kk = k % GT_KC;
        __half* blk = Wp + ((size_t)j * kcn + c) * (size_t)(2 * NT * GT_KC);
        const size_t off = ((size_t)(kk >> 3) * NT + l) * 8 + (kk & 7);
        blk[off] = __ushort_as_half(h);
        blk[(size_t)NT * GT_KC + off] = __ushort_as_half(lo);
    }
}

__device__ __forceinline__ void tmem_ld8(uint32_t addr, uint32_t* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(addr) : "memory");
}
__device__ __forceinline__ void tmem_ld8f(uint32_t addr, float* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "r"(addr) : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t addr, const uint32_t* v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
                 "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                 ::"r"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                 "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
                 "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
                 "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {   // src_bytes 0 -> zero fill
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async8(uint32_t dst, const void* src, int src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

constexpr int GT_PC = 4;             // chunks per accumulation segment (see "promotion" below)
constexpr int GT_RAW = 128 * 256;    // one raw fp32 A chunk: 128 rows x 64 floats
// flattened (tile, chunk) sequence of a persistent CTA; every role walks the same sequence
struct GtIter {
    int tile; unsigned g; int c;       // 32-bit on purpose: the tile -> (batch, m tile, n tile) divisions sit on the epilogue's critical path
    __device__ __forceinline__ void next(int kcn, int stride) { ++g; if (++c == kcn) { c = 0; tile += stride; } }
};

// Promotion: the tensor core adds into its fp32 accumulator with truncation, a bias that grows linearly with the
// number of MMAs chained on one accumulator (measured: 5e-5 relative after 768 MMAs, K = 4096).  MMAs are therefore
// chained over GT_PC chunks (48 MMAs) only; each such segment lands in one of two TMEM accumulators and four
// dedicated warps add it -- in round-to-nearest FFMA-pipe arithmetic -- to the running sum they hold in registers.
// TR (row-major output only): transpose the epilogue's 32 x NT/4 register blocks through warp-private smem tiles so that
// every store instruction writes whole row segments instead of 16 bytes per lane at a row stride.
template <int LOAD, int NT, bool TR = false>
__global__ void __launch_bounds__(GT_THREADS, 1) gemm_tc_kernel(const GemmTcArgs a) {
    constexpr int kTmemCols = NT > 64 ? 512 : 256, kABase = 2 * NT;       // D0 [0,NT) D1 [NT,2NT) A0 A1 (64 columns each)
    constexpr uint32_t kStage = NT * 2 * GT_KC * 2;                       // hi + lo weight blocks of one k chunk
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* raw = smem;                          // [warpgroup 2][stage 2] raw fp32 A chunks (XOR-swizzled 16-byte pieces)
    uint8_t* wst = smem + 4 * GT_RAW;             // [2] weight stages
    uint64_t* bars = reinterpret_cast<uint64_t*>(wst + 2 * kStage);
    uint64_t* a_full = bars;          // [2] count 128: A buffer written (tcgen05.st)
    uint64_t* b_full = bars + 2;      // [2] count 1 + tx bytes: weight stage landed
    uint64_t* empty = bars + 4;       // [2] count 1 (tcgen05.commit): A buffer + weight stage consumed
    uint64_t* seg_full = bars + 6;    // [2] count 1 (tcgen05.commit): segment accumulator complete
    uint64_t* seg_empty = bars + 8;   // [2] count 256: segment accumulator drained
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    if (warp == 8) tmem_alloc(tmem_slot, kTmemCols);
    if (t == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&a_full[i], 128); mbar_init(&b_full[i], 1); mbar_init(&empty[i], 1);
            mbar_init(&seg_full[i], 1); mbar_init(&seg_empty[i], 256);
        }
        mbar_init_fence();
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    const int n_nt = a.N / NT, kcn = a.K / GT_KC;
    const int per_batch = (int)((a.M + 127) / 128) * n_nt, tiles = per_batch * (a.nbatch > 0 ? a.nbatch : 1);
    const int half_n = a.n / 2, P = half_n * half_n;                      // kLoadS2d geometry

    if (warp < 8) {
        // ================= loaders: global fp32 -> smem (cp.async, one own chunk ahead) -> hi/lo fp16 -> TMEM ======
        const int q = warp & 3, par = warp >> 2;
        const uint32_t tA = tmem + ((uint32_t)(q * 32) << 16) + kABase + par * 64;
        const uint32_t raw0 = smem_u32(raw) + par * 2 * GT_RAW;
        const int r = q * 32 + lane;
        // issue the copies of chunk `it` (if it exists) into own stage `st`
        auto issue = [&](const GtIter& it, int st) {
            if (it.tile < tiles) {
                const int fb = it.tile / per_batch;
                const long long mt = (it.tile - fb * per_batch) / n_nt;
                const uint32_t dst = raw0 + st * GT_RAW;
                if (LOAD == kLoadRow) {       // 16 lanes cover the 256 contiguous bytes of one row
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const int rr = q * 32 + 2 * i + (lane >> 4), pc = lane & 15;
                        const long long m = mt * 128 + rr;
                        const bool ok = m < a.M;
                        cp_async16(dst + rr * 256 + ((pc ^ (rr & 7)) << 4), a.A + fb * a.sA + (ok ? m : 0) * a.lda + it.c * GT_KC + pc * 4,
                                   ok ? 16 : 0);
                    }
                } else {                      // lane = own pixel; k = 4 ch + 2 dy + dx (einops 'b c (w 2) (h 2) -> b (c 4) w h')
                    const long long m = mt * 128 + r;
                    const bool ok = m < a.M;
                    const long long b = (ok ? m : 0) / P;
                    const int p = (int)((ok ? m : 0) - b * P), yy = p / half_n, xx = p - yy * half_n;
                    const float* src = a.A + (((size_t)b * a.cin + it.c * 16) * a.n + 2 * yy) * a.n + 2 * xx;
#pragma unroll
                    for (int ch = 0; ch < 16; ++ch) {
                        const uint32_t d = dst + r * 256 + ((ch ^ (r & 7)) << 4);
                        cp_async8(d, src + (size_t)ch * a.n * a.n, ok ? 8 : 0);
                        cp_async8(d + 8, src + (size_t)ch * a.n * a.n + a.n, ok ? 8 : 0);
                    }
                }
            }
            cp_async_commit();
        };
        GtIter cur{(int)blockIdx.x, 0u, 0};
        if (par == 1) cur.next(kcn, gridDim.x);            // warpgroup p owns the chunks with g = p (mod 2)
        GtIter nxt = cur;
        nxt.next(kcn, gridDim.x); nxt.next(kcn, gridDim.x);
        int st = 0;
        issue(cur, 0);
        while (cur.tile < tiles) {
            __syncwarp();                                  // every lane is done reading stage st^1
            issue(nxt, st ^ 1);
            cp_async_wait<1>();
            __syncwarp();
            const uint32_t src = raw0 + st * GT_RAW + r * 256;
            mbar_wait(&empty[par], (uint32_t)((cur.g >> 1) & 1) ^ 1u);
            fence_after_sync();
            const float2 neg1 = make_float2(-1.f, -1.f);
#pragma unroll
            for (int hq = 0; hq < 2; ++hq) {               // 32 k values at a time
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int pc = 0; pc < 8; ++pc) {
                    float4 u;
                    const uint32_t ad = src + (((hq * 8 + pc) ^ (r & 7)) << 4);
                    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(u.x), "=f"(u.y), "=f"(u.z), "=f"(u.w) : "r"(ad));
                    const float2 x0 = make_float2(u.x, u.y), x1 = make_float2(u.z, u.w);
                    const __half2 h0 = __float22half2_rn(x0), h1 = __float22half2_rn(x1);
                    const __half2 l0 = __float22half2_rn(__ffma2_rn(__half22float2(h0), neg1, x0));   // x - float(h), exact
                    const __half2 l1 = __float22half2_rn(__ffma2_rn(__half22float2(h1), neg1, x1));
                    hi[2 * pc] = *reinterpret_cast<const uint32_t*>(&h0); hi[2 * pc + 1] = *reinterpret_cast<const uint32_t*>(&h1);
                    lo[2 * pc] = *reinterpret_cast<const uint32_t*>(&l0); lo[2 * pc + 1] = *reinterpret_cast<const uint32_t*>(&l1);
                }
                tmem_st16(tA + hq * 16, hi);
                tmem_st16(tA + 32 + hq * 16, lo);
            }
            tmem_wait_st();
            fence_before_sync();
            mbar_arrive(&a_full[par]);
            cur = nxt;
            nxt.next(kcn, gridDim.x); nxt.next(kcn, gridDim.x);
            st ^= 1;
        }
        cp_async_wait<0>();
    }
    else if (warp == 8) {
        // ================= MMA issuer =====================================================================
        if (lane == 0) {
            const uint32_t sw = smem_u32(wst);
            const uint32_t idesc = idesc_f16(128, NT);
            constexpr uint64_t step = (uint64_t)(2 * NT * 16) >> 4;        // descriptor units per 16-wide k slab
            unsigned sg = 0;                                               // segment counter
            for (GtIter it{(int)blockIdx.x, 0u, 0}; it.tile < tiles; it.next(kcn, gridDim.x)) {
                const int b = (int)(it.g & 1), cs = it.c % GT_PC, db = (int)(sg & 1);
                const uint32_t ph = (uint32_t)((it.g >> 1) & 1);
                mbar_wait(&a_full[b], ph);
                mbar_wait(&b_full[b], ph);
                if (cs == 0) mbar_wait(&seg_empty[db], (uint32_t)((sg >> 1) & 1) ^ 1u);   // accumulator drained
                fence_after_sync();
                const uint32_t tD = tmem + db * NT, tA = tmem + kABase + b * 64;
                const uint64_t bh = smem_desc(sw + b * kStage, NT * 16, 128);
                const uint64_t bl = smem_desc(sw + b * kStage + NT * GT_KC * 2, NT * 16, 128);
                mma_ts_f16_x4(tD, tA, bh, step, idesc, cs > 0);             // hi . hi
                mma_ts_f16_x4(tD, tA + 32, bh, step, idesc, 1);             // lo . hi
                mma_ts_f16_x4(tD, tA, bl, step, idesc, 1);                  // hi . lo
                mma_commit(&empty[b]);
                if (cs == GT_PC - 1 || it.c == kcn - 1) { mma_commit(&seg_full[db]); ++sg; }
            }
        }
    } else if (warp == 9) {
        // ================= weight-chunk copies (cp.async.bulk) ============================================
        if (lane == 0) {
            for (GtIter it{(int)blockIdx.x, 0u, 0}; it.tile < tiles; it.next(kcn, gridDim.x)) {
                const int b = (int)(it.g & 1), nt = it.tile % n_nt;
                mbar_wait(&empty[b], (uint32_t)((it.g >> 1) & 1) ^ 1u);
                mbar_expect_tx(&b_full[b], kStage);
                bulk_g2s(wst + b * kStage, a.Wp + (long long)(it.tile / per_batch) * a.sW + ((size_t)nt * kcn + it.c) * (size_t)(2 * NT * GT_KC), kStage,
                         &b_full[b]);
            }
        }
    } else {
        // ================= accumulate + epilogue: 2 warps per lane quarter, NT/4 + NT/4 columns each =========
        constexpr int W4 = NT / 4;
        const int q = warp & 3, hset = (warp - 10) >> 2;
        const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
        const int nseg = (kcn + GT_PC - 1) / GT_PC;
        unsigned sg = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
            float aa[W4], bb[W4];     // local columns hset*W4 + i  and  NT/2 + hset*W4 + i  (GroupSort partners)
            for (int s_ = 0; s_ < nseg; ++s_, ++sg) {
                const int db = (int)(sg & 1);
                mbar_wait(&seg_full[db], (uint32_t)((sg >> 1) & 1));
                fence_after_sync();
                if (s_ == 0) {        // first segment: straight into the sums, ONE TMEM round trip for all NT/2 columns
#pragma unroll
                    for (int i = 0; i < W4; i += 8) {
                        tmem_ld8f(lane_base + db * NT + hset * W4 + i, aa + i);
                        tmem_ld8f(lane_base + db * NT + NT / 2 + hset * W4 + i, bb + i);
                    }
                    tmem_wait_ld();
                } else
#pragma unroll
                for (int i = 0; i < W4; i += 8) {
                    uint32_t ua[8], ub[8];
                    tmem_ld8(lane_base + db * NT + hset * W4 + i, ua);
                    tmem_ld8(lane_base + db * NT + NT / 2 + hset * W4 + i, ub);
                    tmem_wait_ld();
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        aa[i + j] += __uint_as_float(ua[j]);
                        bb[i + j] += __uint_as_float(ub[j]);
                    }
                }
                fence_before_sync();
                mbar_arrive(&seg_empty[db]);
            }
            // ---- epilogue: / 64, + bias, GroupSort (layers.py:127-131), store ------------------------------
            const int fb = tile / per_batch, rem = tile - fb * per_batch, mti = rem / n_nt, nt = rem - mti * n_nt;
            const long long mt = mti;
            const long long m = mt * 128 + q * 32 + lane;
            if (LOAD != kLoadRow && m >= a.M) continue;   // row-major: guarded at the stores (the TR path needs the whole warp)
            const int ca = a.group_sort ? nt * (NT / 2) + hset * W4 : nt * NT + hset * W4;
            const int cb = a.group_sort ? ca + a.N / 2 : ca + NT / 2;
#pragma unroll
            for (int i = 0; i < W4; ++i) {
                const float u = fmaf(aa[i], kGtWSInv, a.bias ? __ldg(a.bias + ca + i) : 0.f);
                const float w = fmaf(bb[i], kGtWSInv, a.bias ? __ldg(a.bias + cb + i) : 0.f);
                aa[i] = a.group_sort ? fmaxf(u, w) : u;
                bb[i] = a.group_sort ? fminf(u, w) : w;
            }
            if (LOAD == kLoadRow && TR) {
                // thread = row stores 16 bytes per lane at a row stride: 32 half-filled sectors per request (ncu r01d: 18.9 M
                // sectors written for 9.4 M sectors of output).  Here one instruction writes 32 / W4 whole row segments.
                float* sg = reinterpret_cast<float*>(wst + 2 * kStage + 256) + (warp - 10) * 32 * (W4 + 1);
                const long long m0 = mt * 128 + q * 32;
                constexpr int RPI = 32 / W4;               // rows per store instruction
                const int rl = lane / W4, cl = lane % W4;
                const size_t ldo = (size_t)a.ldo;
                const int nrow = (int)(a.M - m0 < 32 ? a.M - m0 : 32);
                float* const obase = a.out + fb * a.sO + (m0 + rl) * (long long)ldo + cl;
                const float* const sp = sg + rl * (W4 + 1) + cl;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < W4; ++i) sg[lane * (W4 + 1) + i] = half ? bb[i] : aa[i];
                    __syncwarp();
                    float* op = obase + (half ? cb : ca);
                    if (nrow == 32) {
#pragma unroll
                        for (int r0 = 0; r0 < 32; r0 += 8 * RPI) {
                            float v[8];
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[j] = sp[(r0 + j * RPI) * (W4 + 1)];
#pragma unroll
                            for (int j = 0; j < 8; ++j) { *op = v[j]; op += ldo * RPI; }
                        }
                    } else {
                        for (int r = rl; r < nrow; r += RPI) { *op = sp[(r - rl) * (W4 + 1)]; op += ldo * RPI; }
                    }
                }
            } else if (LOAD == kLoadRow) {
                if (m < a.M) {
                    float* o = a.out + fb * a.sO + m * a.ldo;
#pragma unroll
                    for (int i = 0; i < W4; i += 4) {
                        *reinterpret_cast<float4*>(o + ca + i) = make_float4(aa[i], aa[i + 1], aa[i + 2], aa[i + 3]);
                        *reinterpret_cast<float4*>(o + cb + i) = make_float4(bb[i], bb[i + 1], bb[i + 2], bb[i + 3]);
                    }
                }
            } else {
                const long long b = m / P;
                float* o = a.out + (size_t)b * a.N * P + (m - b * P);
#pragma unroll
                for (int i = 0; i < W4; ++i) {
                    o[(size_t)(ca + i) * P] = aa[i];
                    o[(size_t)(cb + i) * P] = bb[i];
                }
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 8) tmem_dealloc(tmem, kTmemCols);
}
}  // namespace fiode

namespace fiode {

size_t gemm_tc_packed_bytes(int N, int K) { return (size_t)N * K * 2 * sizeof(__half); }
bool gemm_tc_supported(int N, int K, int gs) {
    return K % GT_KC == 0 && N % 32 == 0 && (!gs || N % 64 == 0);
}
// tile width (32, 64 or 128 columns per CTA)
int gemm_tc_tile(int N, int gs) {
    if (N == 32 || N == 64 || N == 128) return N;
    if (N % 128 == 0) return 128;
    if (N % 64 == 0) return 64;
    return 32;
}

int launch_pack_w(const float* W, int N, int K, int NT, int gs, void* Wp, cudaStream_t st, int nbatch) {
    const long long total = (long long)N * K * nbatch;
    const int grid = (int)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096);
    pack_w_kernel<<<grid, 256, 0, st>>>(W, N, K, NT, gs, reinterpret_cast<__half*>(Wp), nbatch);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}

// Transposing epilogue for row-major outputs; FIODE_GEMM_ST=0 selects the thread-per-row stores (measured on the 144-batch
// mixing of 8192 images: 0.284 ms -> 0.273 ms; the epilogue is not what bounds that launch)
static bool gemm_tr_stores() {
    static int v = [] { const char* e = getenv("FIODE_GEMM_ST"); return e ? atoi(e) : 1; }();
    return v != 0;
}
template <int LOAD, int NT, bool TR>
static int launch_gemm_tc_ntr(const GemmTcArgs& a, cudaStream_t st) {
    const size_t smem = (size_t)4 * GT_RAW + (size_t)2 * NT * 2 * GT_KC * 2 + 256 +
                        (TR ? (size_t)8 * 32 * (NT / 4 + 1) * sizeof(float) : 0);   // + epilogue transpose tiles
    FIODE_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<LOAD, NT, TR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long tiles = ((a.M + 127) / 128) * (a.N / NT) * (a.nbatch > 0 ? a.nbatch : 1);
    const int grid = (int)(tiles < sm_count() ? tiles : sm_count());       // one persistent CTA per SM
    gemm_tc_kernel<LOAD, NT, TR><<<grid, GT_THREADS, smem, st>>>(a);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
template <int LOAD, int NT>
static int launch_gemm_tc_nt(const GemmTcArgs& a, cudaStream_t st) {
    if (LOAD == kLoadRow && gemm_tr_stores()) return launch_gemm_tc_ntr<LOAD, NT, LOAD == kLoadRow>(a, st);
    return launch_gemm_tc_ntr<LOAD, NT, false>(a, st);
}
template <int LOAD>
static int launch_gemm_tc_t(const GemmTcArgs& a, cudaStream_t st) {
    switch (a.NT) {
        case 32: return launch_gemm_tc_nt<LOAD, 32>(a, st);
        case 64: return launch_gemm_tc_nt<LOAD, 64>(a, st);
        case 128: return launch_gemm_tc_nt<LOAD, 128>(a, st);
    }
    set_error("gemm_tc: unsupported tile width %d", a.NT);
    return -1;
}

// y (M, N) = X (M, K) . W^T + bias (+ GroupSort); Wp from launch_pack_w(W, N, K, gemm_tc_tile(N, gs), gs)
int launch_linear_tc(const float* X, long long lda, const void* Wp, const float* bias, long long M, int N, int K, int gs,
                     float* y, long long ldy, cudaStream_t st) {
    GemmTcArgs a{};
    a.A = X; a.lda = lda; a.M = M; a.N = N; a.K = K; a.NT = gemm_tc_tile(N, gs);
    a.Wp = reinterpret_cast<const __half*>(Wp); a.bias = bias; a.group_sort = gs; a.out = y; a.ldo = ldy;
    return launch_gemm_tc_t<kLoadRow>(a, st);
}
// nbatch independent products y_f (M, N) = X_f (M, K) . W_f^T (no bias): the per-frequency channel mixing of a CayleyConv
int launch_linear_tc_batched(const float* X, long long lda, long long sA, const void* Wp, long long M, int N, int K, int nbatch,
                             float* y, long long ldy, long long sO, cudaStream_t st) {
    GemmTcArgs a{};
    a.A = X; a.lda = lda; a.M = M; a.N = N; a.K = K; a.NT = gemm_tc_tile(N, 0);
    a.Wp = reinterpret_cast<const __half*>(Wp); a.out = y; a.ldo = ldy;
    a.nbatch = nbatch; a.sA = sA; a.sO = sO; a.sW = (long long)N * K * 2;
    return launch_gemm_tc_t<kLoadRow>(a, st);
}
// stride-2 CayleyConv as a 1x1 conv on the space-to-depth tensor: x (B, cin, n, n) -> y (B, cout, n/2, n/2)
int launch_conv1x1_tc(const float* x, const void* Wp, const float* bias, int B, int cin, int cout, int n, int gs, float* y,
                      cudaStream_t st) {
    GemmTcArgs a{};
    a.A = x; a.cin = cin; a.n = n; a.M = (long long)B * (n / 2) * (n / 2); a.N = cout; a.K = 4 * cin;
    a.NT = gemm_tc_tile(cout, gs);
    a.Wp = reinterpret_cast<const __half*>(Wp); a.bias = bias; a.group_sort = gs; a.out = y;
    return launch_gemm_tc_t<kLoadS2d>(a, st);
}

}  // namespace fiode
