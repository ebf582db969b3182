// fiode_b200/csrc/cayley.cu -- Cayley transform of the layer weights (weight preparation, run once
// per weight version and cached by the host; the reference redoes it on EVERY forward).
//
// Replaces libs/ortho_conv/layers.py:50-60 `cayley`, as used by
//   * CayleyLinear.forward :103-104   Q = cayley(alpha * W / ||W||_F)           (real, any shape)
//   * CayleyConv.forward   :74-82     Q[phi] = cayley(alpha * wfft[phi] / ||wfft||_F) per frequency
// and the lazily created CayleyConv.alpha = ||wfft||_F (:80-81).
//
//   W (R x n after the plain transpose of :54-55 when cin > cout), U = W[:n], V = W[n:]
//   A = U - U^H + V^H V ;  M = (I + A)^-1 ;  Q = [ M (I - A) ; -2 V M ]
//
// I + A has the positive-definite Hermitian part I + V^H V, so Gauss-Jordan elimination needs no
// pivoting.  Real path: GEMMs (linear.cu) + a ping-pong Gauss-Jordan kernel (cooperative grid for
// n = 512, one CTA for n <= 128).  Complex path: one CTA per frequency, everything in shared
// memory; the weight spectrum is the direct 9-tap sum of SURVEY App. C-1 (no FFT of the weights).
#include "common.cuh"
#include <cstdlib>
#include "fft.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace fiode {

int gemm_nt(const float* A, int lda, const float* Bm, int ldb, int M, int N, int K, const float* bias, float alpha,
            int group_sort, float* y, int ldy, float* ws, size_t ws_bytes, cudaStream_t st);
int gemm_nt_tall(const float* A, int lda, const float* Bm, int ldb, int M, int N, int K, float alpha, float* y, int ldy,
                 float* ws, size_t ws_bytes, cudaStream_t st);
size_t gemm_nt_workspace(int M, int N, int K);
int gemm_nt_gram(const float* A, int lda, int n, int K, float* y, int ldy, float* ws, size_t ws_bytes, cudaStream_t st);

// ---- Frobenius norm, deterministic two-stage --------------------------------------------------
constexpr int kNormBlocks = 128;
__global__ void __launch_bounds__(256) sumsq_partial_kernel(const float* __restrict__ x, size_t n,
                                                            double* __restrict__ part) {
    __shared__ double sh[8];
    double acc = 0;
    for (size_t e = blockIdx.x * 256 + threadIdx.x; e < n; e += (size_t)gridDim.x * 256) {
        const double v = x[e];
        acc += v * v;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int i = 0; i < 8; ++i) s += sh[i];
        part[blockIdx.x] = s;
    }
}
// scale[0] = alpha / (sqrt(sum part) * mult) ; scale[1] = sqrt(sum part) * mult
__global__ void norm_finalize_kernel(const double* __restrict__ part, int np, const float* __restrict__ alpha,
                                     float mult, float* __restrict__ scale) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double s = 0;
        for (int i = 0; i < np; ++i) s += part[i];
        const float nrm = (float)sqrt(s) * mult;
        scale[1] = nrm;
        scale[0] = alpha ? alpha[0] / nrm : 1.f / nrm;
    }
}

// Ws (R x n) = scale * (tr ? W^T : W) ; Vt (n x (R-n)) = transpose of Ws[n:]
__global__ void cayley_prep1_kernel(const float* __restrict__ W, int cout, int cin, int tr,
                                    const float* __restrict__ scale, float* __restrict__ Ws,
                                    float* __restrict__ Vt) {
    const int R = tr ? cin : cout, n = tr ? cout : cin;
    const float sc = scale[0];
    const size_t total = (size_t)R * n;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int r = (int)(e / n), c = (int)(e - (size_t)r * n);
        const float v = sc * (tr ? W[(size_t)c * cin + r] : W[(size_t)r * cin + c]);
        Ws[e] = v;
        if (r >= n) Vt[(size_t)c * (R - n) + (r - n)] = v;
    }
}
// The same two outputs through a 32 x 32 shared-memory tile (block 32 x 8): out1 (R x n) = f * (tr ? S^T : S), out2 (n x (R-n)) = the
// transpose of out1's rows n.. -- every global access runs along a contiguous axis (the element-per-thread kernels above and
// below read or write one of the three arrays with a stride of a whole row).  f = scale[0] when scale is given, else 1.
__global__ void __launch_bounds__(256) cayley_tile_prep_kernel(const float* __restrict__ S, int cout, int cin, int tr,
                                                               const float* __restrict__ scale, float* __restrict__ out1,
                                                               float* __restrict__ out2) {
    __shared__ float tile[32][33];
    const int R = tr ? cin : cout, n = tr ? cout : cin;
    const float f = scale ? scale[0] : 1.f;
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32, x = threadIdx.x;
    // tile[rl][cl] = f * value at (r0 + rl, c0 + cl)
    if (tr) {
        for (int cl = threadIdx.y; cl < 32; cl += 8)
            if (c0 + cl < n && r0 + x < R) tile[x][cl] = f * S[(size_t)(c0 + cl) * cin + r0 + x];
    } else {
        for (int rl = threadIdx.y; rl < 32; rl += 8)
            if (r0 + rl < R && c0 + x < n) tile[rl][x] = f * S[(size_t)(r0 + rl) * cin + c0 + x];
    }
    __syncthreads();
    for (int rl = threadIdx.y; rl < 32; rl += 8)
        if (r0 + rl < R && c0 + x < n) out1[(size_t)(r0 + rl) * n + c0 + x] = tile[rl][x];
    for (int cl = threadIdx.y; cl < 32; cl += 8)
        if (c0 + cl < n && r0 + x < R && r0 + x >= n) out2[(size_t)(c0 + cl) * (R - n) + (r0 + x - n)] = tile[x][cl];
}
static void launch_tile_prep(const float* S, int cout, int cin, int tr, const float* scale, float* out1, float* out2,
                             cudaStream_t st) {
    const int R = tr ? cin : cout, n = tr ? cout : cin;
    cayley_tile_prep_kernel<<<dim3((n + 31) / 32, (R + 31) / 32), dim3(32, 8), 0, st>>>(S, cout, cin, tr, scale, out1, out2);
}
// IpA = I + U - U^T + G ; ImAT = (I - A)^T = I - (U^T - U + G^T)
__global__ void cayley_prep2_kernel(const float* __restrict__ Ws, const float* __restrict__ G, int n, int has_v,
                                    float* __restrict__ IpA, float* __restrict__ ImAT) {
    const int total = n * n;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
        const int i = e / n, j = e - i * n;
        const float uij = Ws[i * n + j], uji = Ws[j * n + i];
        const float gij = has_v ? G[i * n + j] : 0.f, gji = has_v ? G[j * n + i] : 0.f;
        const float d = (i == j) ? 1.f : 0.f;
        IpA[e] = d + ((uij - uji) + gij);
        ImAT[e] = d - ((uji - uij) + gji);
    }
}
__global__ void transpose_kernel(const float* __restrict__ x, int rows, int cols, float* __restrict__ y) {
    __shared__ float tile[32][33];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        if (r < rows && c < cols) tile[i][threadIdx.x] = x[(size_t)r * cols + c];
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (r < rows && c < cols) y[(size_t)c * rows + r] = tile[threadIdx.x][i];
    }
}

// Ping-pong Gauss-Jordan inversion without pivoting: after n steps the inverse is in (n even ? X : Y).
// Cooperative launch (grid.sync per pivot).  out[i][j] depends on in[i][j], in[i][k], in[k][j], in[k][k].
__global__ void __launch_bounds__(256) gj_inverse_kernel(float* X, float* Y, int n) {
    cg::grid_group grid = cg::this_grid();
    const int total = n * n;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    for (int k = 0; k < n; ++k) {
        const float* in = (k & 1) ? Y : X;
        float* out = (k & 1) ? X : Y;
        const float ip = 1.f / __ldcg(in + k * n + k);
        for (int e = tid; e < total; e += nth) {
            const int i = e / n, j = e - i * n;
            float v;
            if (i == k) v = (j == k) ? ip : __ldcg(in + e) * ip;
            else {
                const float f = __ldcg(in + i * n + k);
                v = (j == k) ? -f * ip : fmaf(-f, __ldcg(in + k * n + j) * ip, __ldcg(in + e));
            }
            out[e] = v;
        }
        grid.sync();
    }
}

// n <= 128: the whole elimination in ONE CTA, the matrix in REGISTERS: warp w owns rows w + 32 a, lane l columns l + 32 c
// (16 elements per thread).  Per pivot k the owner warp publishes the pivot row through a double-buffered 128-float line
// (ONE block barrier per pivot), the column-k multipliers come from lane k % 32 of the own warp by shuffle.  Same operations
// per element as the shared-memory ping-pong it replaces (bit-identical inverse), ~3.5x faster (no smem round trip per element).
__global__ void __launch_bounds__(1024) gj_inverse_smem_kernel(const float* __restrict__ Xin, float* __restrict__ out, int n) {
    __shared__ float prow[2][128];
    const int t = threadIdx.x, w = t >> 5, l = t & 31;
    float v[4][4];                                     // v[a][c] = M[w + 32 a][l + 32 c]
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int i = w + 32 * a, j = l + 32 * c;
            v[a][c] = (i < n && j < n) ? Xin[i * n + j] : 0.f;
        }
#pragma unroll
    for (int ka = 0; ka < 4; ++ka) {                   // static ka (= k / 32): register-array indices are compile-time constants
        for (int kk = 0; kk < 32; ++kk) {
            const int k = 32 * ka + kk;
            if (k >= n) break;                         // uniform
            float* pr = prow[k & 1];
            if (w == kk) {
#pragma unroll
                for (int c = 0; c < 4; ++c) pr[l + 32 * c] = v[ka][c];
            }
            __syncthreads();
            const float ip = 1.f / pr[k];
            float pk[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) pk[c] = pr[l + 32 * c] * ip;
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const int i = w + 32 * a;
                const float f = __shfl_sync(0xffffffffu, v[a][ka], kk);        // M[i][k]: column k = kk + 32 ka
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int j = l + 32 * c;
                    float r;
                    if (i == k) r = (j == k) ? ip : pk[c];
                    else r = (j == k) ? -f * ip : fmaf(-f, pk[c], v[a][c]);
                    v[a][c] = r;
                }
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int i = w + 32 * a, j = l + 32 * c;
            if (i < n && j < n) out[i * n + j] = v[a][c];
        }
}

// n a multiple of 64: blocked Gauss-Jordan, one pivot block (64 x 64) per step, out of place X -> Y:
//   Y_kk = D^-1, Y_kj = D^-1 X_kj, Y_ik = -X_ik D^-1, Y_ij = X_ij - X_ik D^-1 X_kj      (D = X_kk)
// panel kernel: CTA j inverts D in shared memory (redundantly, 64 block barriers) and writes block row k of Y;
// update kernel: CTA (i, j), i != k, applies the rank-64 update with Y's block row k.  2 plain launches per step.
constexpr int GB = 64;
__global__ void __launch_bounds__(256) gj_block_panel_kernel(const float* __restrict__ X, float* __restrict__ Y, int n, int kb) {
    __shared__ float D0[GB * GB];
    __shared__ float D1[GB * GB];
    __shared__ float P[GB * GB];
    const int t = threadIdx.x, jb = blockIdx.x;
    for (int e = t; e < GB * GB; e += 256) {
        const int i = e / GB, j = e - i * GB;
        D0[e] = X[(size_t)(kb * GB + i) * n + kb * GB + j];
        P[e] = X[(size_t)(kb * GB + i) * n + jb * GB + j];
    }
    __syncthreads();
    // the pivot block is inverted in registers: warp w owns rows w + 8 a, lane l columns l, l + 32; per pivot the owner warp
    // publishes the pivot row (double-buffered line, one block barrier), multipliers by shuffle; same operations per element as
    // the shared-memory ping-pong this replaces
    const int w = t >> 5, l = t & 31;
    float v0[GB / 8], v1[GB / 8];
#pragma unroll
    for (int a = 0; a < GB / 8; ++a) { v0[a] = D0[(w + 8 * a) * GB + l]; v1[a] = D0[(w + 8 * a) * GB + l + 32]; }
    float* prow = D1;                                  // two 64-float lines
#pragma unroll
    for (int ka = 0; ka < GB / 8; ++ka) {              // static ka: every register-array index below is a compile-time constant
        for (int kk = 0; kk < 8; ++kk) {
            const int k = 8 * ka + kk;
            float* pr = prow + (k & 1) * GB;
            if (w == kk) { pr[l] = v0[ka]; pr[l + 32] = v1[ka]; }
            __syncthreads();
            const float ip = 1.f / pr[k];
            const float pk0 = pr[l] * ip, pk1 = pr[l + 32] * ip;
#pragma unroll
            for (int a = 0; a < GB / 8; ++a) {
                const int i = w + 8 * a;
                const float f = __shfl_sync(0xffffffffu, ka < 4 ? v0[a] : v1[a], k & 31);
                float r0, r1;
                if (i == k) { r0 = (l == k) ? ip : pk0; r1 = (l + 32 == k) ? ip : pk1; }
                else {
                    r0 = (l == k) ? -f * ip : fmaf(-f, pk0, v0[a]);
                    r1 = (l + 32 == k) ? -f * ip : fmaf(-f, pk1, v1[a]);
                }
                v0[a] = r0; v1[a] = r1;
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int a = 0; a < GB / 8; ++a) { D0[(w + 8 * a) * GB + l] = v0[a]; D0[(w + 8 * a) * GB + l + 32] = v1[a]; }
    __syncthreads();
    const float* Dinv = D0;
    if (jb == kb) {
        for (int e = t; e < GB * GB; e += 256) Y[(size_t)(kb * GB + e / GB) * n + jb * GB + e % GB] = Dinv[e];
        return;
    }
    // Y_kj = D^-1 X_kj: thread tile 4 x 4 (rows ty + 16 a, columns tx + 16 b) like the update kernel -- 8 loads per 16 FMAs
    const int tx = t & 15, ty = t >> 4;
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
#pragma unroll 4
    for (int r = 0; r < GB; ++r) {
        float dl[4], pq[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) { dl[a] = Dinv[(ty + 16 * a) * GB + r]; pq[a] = P[r * GB + tx + 16 * a]; }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(dl[a], pq[b], acc[a][b]);
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) Y[(size_t)(kb * GB + ty + 16 * a) * n + jb * GB + tx + 16 * b] = acc[a][b];
}
__global__ void __launch_bounds__(256) gj_block_update_kernel(const float* __restrict__ X, float* __restrict__ Y, int n, int kb) {
    __shared__ float Ls[GB * (GB + 1)];               // X_ik
    __shared__ float Rs[GB * GB];                      // Y_kj
    const int t = threadIdx.x, jb = blockIdx.x;
    const int ib = blockIdx.y + (blockIdx.y >= kb ? 1 : 0);          // skip the pivot block row
    for (int e = t; e < GB * GB; e += 256) {
        const int i = e / GB, j = e - i * GB;
        Ls[i * (GB + 1) + j] = X[(size_t)(ib * GB + i) * n + kb * GB + j];
        Rs[e] = Y[(size_t)(kb * GB + i) * n + jb * GB + j];
    }
    __syncthreads();
    const int tx = t & 15, ty = t >> 4;                // thread tile 4 x 4: rows ty + 16 a, columns tx + 16 b
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
#pragma unroll 4
    for (int r = 0; r < GB; ++r) {
        float l[4], q[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) { l[a] = Ls[(ty + 16 * a) * (GB + 1) + r]; q[a] = Rs[r * GB + tx + 16 * a]; }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(l[a], q[b], acc[a][b]);
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const size_t o = (size_t)(ib * GB + ty + 16 * a) * n + jb * GB + tx + 16 * b;
            Y[o] = (jb == kb ? 0.f : X[o]) - acc[a][b];
        }
}

// The panel and the update kernel above as ONE launch per block step: CTA (ib, jb) inverts D = X_kk itself (every CTA does, in
// parallel: the inversion is a latency chain, not work), forms T = D^-1 X_kj in registers and then writes its own block --
// T for the pivot row, X_ij - X_ik T elsewhere -- without the round trip of block row k through global memory and the second
// launch.  Per element the same operations in the same order as panel + update: bit-identical results.
constexpr int GLS = GB + 4;              // row stride of the X_ik tile: float4 rows, and ty / ty + 1 land on different banks
constexpr size_t kGjStepSmem = (size_t)(2 * GB * GB + GB * GLS + 2 * GB) * sizeof(float);
__global__ void __launch_bounds__(256) gj_block_step_kernel(const float* __restrict__ X, float* __restrict__ Y, int n, int kb) {
    extern __shared__ __align__(16) float gj_sm[];
    float* D0 = gj_sm;                       // D, then D^-1
    float* P = D0 + GB * GB;                 // X_kj, then T = D^-1 X_kj
    float* Ls = P + GB * GB;                 // X_ik, row stride GLS
    float* prow = Ls + GB * GLS;             // two 64-float lines
    const int t = threadIdx.x, jb = blockIdx.x, ib = blockIdx.y;
    {   // the three 64 x 64 tiles: all twelve 16-byte loads of a thread in flight before the first store (the scalar version
        // spent a quarter of the kernel waiting for them one after the other)
        float4 d4[4], p4[4], l4[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int e4 = t + 256 * q, i = e4 >> 4, j = (e4 & 15) * 4;
            d4[q] = *reinterpret_cast<const float4*>(X + (size_t)(kb * GB + i) * n + kb * GB + j);
            p4[q] = *reinterpret_cast<const float4*>(X + (size_t)(kb * GB + i) * n + jb * GB + j);
            if (ib != kb) l4[q] = *reinterpret_cast<const float4*>(X + (size_t)(ib * GB + i) * n + kb * GB + j);
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int e4 = t + 256 * q, i = e4 >> 4, j = (e4 & 15) * 4;
            *reinterpret_cast<float4*>(D0 + i * GB + j) = d4[q];
            *reinterpret_cast<float4*>(P + i * GB + j) = p4[q];
            if (ib != kb) *reinterpret_cast<float4*>(Ls + i * GLS + j) = l4[q];
        }
    }
    __syncthreads();
    const int w = t >> 5, l = t & 31;
    float v0[GB / 8], v1[GB / 8];
#pragma unroll
    for (int a = 0; a < GB / 8; ++a) { v0[a] = D0[(w + 8 * a) * GB + l]; v1[a] = D0[(w + 8 * a) * GB + l + 32]; }
#pragma unroll
    for (int ka = 0; ka < GB / 8; ++ka) {
        for (int kk = 0; kk < 8; ++kk) {
            const int k = 8 * ka + kk;
            float* pr = prow + (k & 1) * GB;
            if (w == kk) { pr[l] = v0[ka]; pr[l + 32] = v1[ka]; }
            __syncthreads();
            const float ip = 1.f / pr[k];
            const float pk0 = pr[l] * ip, pk1 = pr[l + 32] * ip;
#pragma unroll
            for (int a = 0; a < GB / 8; ++a) {
                const int i = w + 8 * a;
                const float f = __shfl_sync(0xffffffffu, ka < 4 ? v0[a] : v1[a], k & 31);
                float r0, r1;
                if (i == k) { r0 = (l == k) ? ip : pk0; r1 = (l + 32 == k) ? ip : pk1; }
                else {
                    r0 = (l == k) ? -f * ip : fmaf(-f, pk0, v0[a]);
                    r1 = (l + 32 == k) ? -f * ip : fmaf(-f, pk1, v1[a]);
                }
                v0[a] = r0; v1[a] = r1;
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int a = 0; a < GB / 8; ++a) { D0[(w + 8 * a) * GB + l] = v0[a]; D0[(w + 8 * a) * GB + l + 32] = v1[a]; }
    __syncthreads();
    const int tx = t & 15, ty = t >> 4;                // thread tile 4 x 4: rows ty + 16 a, columns tx + 16 b
    float acc[4][4];
    if (jb == kb) {
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = D0[(ty + 16 * a) * GB + tx + 16 * b];
    } else {
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
#pragma unroll 4
        for (int r = 0; r < GB; ++r) {
            float dl[4], pq[4];
#pragma unroll
            for (int a = 0; a < 4; ++a) { dl[a] = D0[(ty + 16 * a) * GB + r]; pq[a] = P[r * GB + tx + 16 * a]; }
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(dl[a], pq[b], acc[a][b]);
        }
    }
    if (ib == kb) {                                    // pivot block row: Y_kj = T (D^-1 itself when jb = k)
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) Y[(size_t)(kb * GB + ty + 16 * a) * n + jb * GB + tx + 16 * b] = acc[a][b];
        return;
    }
    __syncthreads();                                   // everyone has read P
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) P[(ty + 16 * a) * GB + tx + 16 * b] = acc[a][b];
    __syncthreads();
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
#pragma unroll 4
    for (int r = 0; r < GB; ++r) {
        float lv[4], q[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) { lv[a] = Ls[(ty + 16 * a) * GLS + r]; q[a] = P[r * GB + tx + 16 * a]; }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(lv[a], q[b], acc[a][b]);
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const size_t o = (size_t)(ib * GB + ty + 16 * a) * n + jb * GB + tx + 16 * b;
            Y[o] = (jb == kb ? 0.f : X[o]) - acc[a][b];
        }
}
static bool gj_fused_step() {
    static const bool v = [] { const char* e = getenv("FIODE_GJ_FUSED"); return !(e && e[0] == '0'); }();
    return v;
}

static int gj_inverse(float* X, float* Y, int n, float** result, cudaStream_t st) {
    if (n <= 128) {
        gj_inverse_smem_kernel<<<1, 1024, 0, st>>>(X, Y, n);
        count_launch(1);
        *result = Y;
        return 0;
    }
    if (n % GB == 0 && gj_fused_step()) {
        static std::atomic<bool> attr[kMaxDevices];
        if (first_use_on_device(attr))
            FIODE_CUDA(cudaFuncSetAttribute(gj_block_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGjStepSmem));
        const int nb = n / GB;
        float *in = X, *out = Y;
        for (int kb = 0; kb < nb; ++kb) {
            gj_block_step_kernel<<<dim3(nb, nb), 256, kGjStepSmem, st>>>(in, out, n, kb);
            float* tmp = in; in = out; out = tmp;
        }
        count_launch(nb);
        *result = in;
        return 0;
    }
    if (n % GB == 0) {
        const int nb = n / GB;
        float *in = X, *out = Y;
        for (int kb = 0; kb < nb; ++kb) {
            gj_block_panel_kernel<<<nb, 256, 0, st>>>(in, out, n, kb);
            gj_block_update_kernel<<<dim3(nb, nb - 1), 256, 0, st>>>(in, out, n, kb);
            float* tmp = in; in = out; out = tmp;
        }
        count_launch(2 * nb);
        *result = in;
        return 0;
    }
    int grid = (n * n + 256 * 4 - 1) / (256 * 4);
    const int cap = sm_count() * 2;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    void* args[] = {&X, &Y, &n};
    cudaError_t e = cudaLaunchCooperativeKernel((void*)gj_inverse_kernel, dim3(grid), dim3(256), args, 0, st);
    if (e != cudaSuccess) return cuda_fail(e, "gj_inverse launch");
    count_launch(1);
    *result = (n & 1) ? Y : X;
    return 0;
}

// ---- small layers, fused: n = min(cout, cin) <= 64 and R n <= 16384 (R = max): ONE CTA does the whole transform ------------
// (the dynamics' 128x10 / 10x128 layers, fc 10x512, the stride-2 convs' 32x128 and 64x256 matrices).  The staged version above
// costs such a layer ~12 launches forward and ~25 backward, each a few microseconds of latency on an otherwise idle GPU: 7 of
// the FI step's 11 layers are this size.  Same formulas as the staged path (libs/ortho_conv/layers.py:50-60 and its adjoint).
constexpr int kSmallThreads = 1024;
__device__ __forceinline__ bool cayley_small_ok(int R, int n) { return n <= 64 && (size_t)R * n <= 16384; }
static bool cayley_small_host_ok(int cout, int cin) {
    const int R = cout > cin ? cout : cin, n = cout > cin ? cin : cout;
    static const bool off = [] { const char* e = getenv("FIODE_CAYLEY_FUSED"); return e && e[0] == '0'; }();
    return !off && n <= 64 && (size_t)R * n <= 16384;
}
struct SmallSm { float* Ws; float* Gs; float* A0; float* A1; float* A2; float* A3; float* Mi; float* prow; double* red; float* sc; };
__device__ __forceinline__ SmallSm small_carve(float* base, int R, int n, bool with_g) {
    SmallSm s; float* p = base;
    const int ld = n + 1;                           // odd-ish leading dimension: column walks (stride ld) hit different banks
    s.Ws = p; p += (size_t)R * ld;
    s.Gs = p; if (with_g) p += (size_t)R * ld;
    s.A0 = p; p += n * ld; s.A1 = p; p += n * ld; s.A2 = p; p += n * ld; s.A3 = p; p += n * ld; s.Mi = p; p += n * ld;
    s.prow = p; p += 256;
    s.red = reinterpret_cast<double*>(p + (((size_t)(p - base)) & 1)); p += 2 * 34;
    s.sc = p;
    return s;
}
static size_t small_smem_bytes(int R, int n, bool with_g) { return ((size_t)R * (n + 1) * (with_g ? 2 : 1) + 5 * (size_t)n * (n + 1) + 256 + 72 + 8) * sizeof(float); }

// ||W||_F (double accumulation), scale -> sc[0] = alpha / (||W|| mult), sc[1] = ||W|| mult
__device__ __forceinline__ void small_norm(const float* __restrict__ W, size_t total, const float* alpha, float mult, const SmallSm& S) {
    double acc = 0;
    for (size_t e = threadIdx.x; e < total; e += kSmallThreads) { const double v = W[e]; acc += v * v; }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) S.red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        for (int i = 0; i < kSmallThreads / 32; ++i) t += S.red[i];
        const float nrm = (float)sqrt(t) * mult;
        S.sc[1] = nrm;
        S.sc[0] = alpha ? alpha[0] / nrm : 1.f / nrm;
    }
    __syncthreads();
}
// Ws, X = I + A (A0), I - A (A1); A = U - U^T + V^T V
__device__ __forceinline__ void small_build(const float* __restrict__ W, int cout, int cin, int tr, int R, int n, const SmallSm& S) {
    const int t = threadIdx.x, ld = n + 1;
    const float sc = S.sc[0];
    for (int e = t; e < cout * cin; e += kSmallThreads) {   // global order (coalesced); (r, c) = position in the R x n working matrix
        const int o = e / cin, i2 = e - o * cin;
        const int r = tr ? i2 : o, c = tr ? o : i2;
        S.Ws[r * ld + c] = sc * W[e];
    }
    __syncthreads();
    for (int e = t; e < n * n; e += kSmallThreads) {
        const int i = e / n, j = e - i * n;
        float g = 0.f;
        for (int r = n; r < R; ++r) g = fmaf(S.Ws[r * ld + i], S.Ws[r * ld + j], g);
        const float a = (S.Ws[i * ld + j] - S.Ws[j * ld + i]) + g;
        const float d = (i == j) ? 1.f : 0.f;
        S.A0[i * ld + j] = d + a;
        S.A1[i * ld + j] = d - a;
    }
    __syncthreads();
}
// out (smem) = X^-1 for n <= 64 ... 128: Gauss-Jordan without pivoting in registers (see gj_inverse_smem_kernel)
__device__ __forceinline__ void small_inverse(const float* X, float* out, int n, int ld, float* prow2) {
    const int t = threadIdx.x, w = t >> 5, l = t & 31;
    float v[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int i = w + 32 * a, j = l + 32 * c;
            v[a][c] = (i < n && j < n) ? X[i * ld + j] : 0.f;
        }
#pragma unroll
    for (int ka = 0; ka < 4; ++ka) {
        for (int kk = 0; kk < 32; ++kk) {
            const int k = 32 * ka + kk;
            if (k >= n) break;
            float* pr = prow2 + (k & 1) * 128;
            if (w == kk) {
#pragma unroll
                for (int c = 0; c < 4; ++c) pr[l + 32 * c] = v[ka][c];
            }
            __syncthreads();
            const float ip = 1.f / pr[k];
            float pk[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) pk[c] = pr[l + 32 * c] * ip;
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const int i = w + 32 * a;
                const float f = __shfl_sync(0xffffffffu, v[a][ka], kk);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int j = l + 32 * c;
                    float r;
                    if (i == k) r = (j == k) ? ip : pk[c];
                    else r = (j == k) ? -f * ip : fmaf(-f, pk[c], v[a][c]);
                    v[a][c] = r;
                }
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int i = w + 32 * a, j = l + 32 * c;
            if (i < n && j < n) out[i * ld + j] = v[a][c];
        }
    __syncthreads();
}

extern __shared__ __align__(16) float sm_small[];
__device__ __forceinline__ void cayley_small_fwd_body(const float* __restrict__ W, const float* __restrict__ alpha, int cout, int cin,
                                                      float mult, float* __restrict__ Q, float* __restrict__ M_save) {
    const int tr = cin > cout, R = tr ? cin : cout, n = tr ? cout : cin, t = threadIdx.x, ld = n + 1;
    const SmallSm S = small_carve(sm_small, R, n, false);
    small_norm(W, (size_t)cout * cin, alpha, mult, S);
    small_build(W, cout, cin, tr, R, n, S);
    small_inverse(S.A0, S.Mi, n, ld, S.prow);
    if (M_save) for (int e = t; e < n * n; e += kSmallThreads) M_save[e] = S.Mi[(e / n) * ld + e % n];
    // Q_top = M (I - A), Q_bottom = -2 V M; Q = tr ? Qfull^T : Qfull
    for (int e = t; e < cout * cin; e += kSmallThreads) {   // global order: coalesced stores
        const int o = e / cin, i2 = e - o * cin;
        const int r = tr ? i2 : o, j = tr ? o : i2;
        float acc = 0.f;
        if (r < n) { for (int k = 0; k < n; ++k) acc = fmaf(S.Mi[r * ld + k], S.A1[k * ld + j], acc); }
        else { for (int k = 0; k < n; ++k) acc = fmaf(S.Ws[r * ld + k], S.Mi[k * ld + j], acc); acc *= -2.f; }
        Q[e] = acc;
    }
}

__global__ void __launch_bounds__(kSmallThreads) cayley_small_fwd_kernel(const float* __restrict__ W, const float* __restrict__ alpha,
                                                                         int cout, int cin, float mult, float* __restrict__ Q,
                                                                         float* __restrict__ M_save) {
    cayley_small_fwd_body(W, alpha, cout, cin, mult, Q, M_save);
}

__device__ __forceinline__ void cayley_small_bwd_body(const float* __restrict__ W, const float* __restrict__ alpha, int cout, int cin,
                                                      float mult, const float* __restrict__ gQ, const float* __restrict__ M_saved,
                                                      float* __restrict__ gW, float* __restrict__ galpha) {
    const int tr = cin > cout, R = tr ? cin : cout, n = tr ? cout : cin, t = threadIdx.x, ld = n + 1;
    const SmallSm S = small_carve(sm_small, R, n, true);
    for (int e = t; e < cout * cin; e += kSmallThreads) {   // incoming gradient in the working orientation (R x n), staged once
        const int o = e / cin, i2 = e - o * cin;
        S.Gs[(tr ? i2 : o) * ld + (tr ? o : i2)] = gQ[e];
    }
    auto gq = [&](int r, int c) { return S.Gs[r * ld + c]; };
    small_norm(W, (size_t)cout * cin, alpha, mult, S);
    small_build(W, cout, cin, tr, R, n, S);
    if (M_saved) { for (int e = t; e < n * n; e += kSmallThreads) S.Mi[(e / n) * ld + e % n] = M_saved[e]; __syncthreads(); }
    else small_inverse(S.A0, S.Mi, n, ld, S.prow);
    // gM = 2 G1 - 2 V^T G2   (A0)
    for (int e = t; e < n * n; e += kSmallThreads) {
        const int i = e / n, j = e - i * n;
        float pacc = 0.f;
        for (int r = n; r < R; ++r) pacc = fmaf(S.Ws[r * ld + i], gq(r, j), pacc);
        S.A0[i * ld + j] = 2.f * gq(i, j) - 2.f * pacc;
    }
    __syncthreads();
    // T1t = M gM^T   (A1)
    for (int e = t; e < n * n; e += kSmallThreads) {
        const int i = e / n, j = e - i * n;
        float acc = 0.f;
        for (int k = 0; k < n; ++k) acc = fmaf(S.Mi[i * ld + k], S.A0[j * ld + k], acc);
        S.A1[i * ld + j] = acc;
    }
    __syncthreads();
    // gA^T = -T1t M   (A2)
    for (int e = t; e < n * n; e += kSmallThreads) {
        const int i = e / n, j = e - i * n;
        float acc = 0.f;
        for (int k = 0; k < n; ++k) acc = fmaf(S.A1[i * ld + k], S.Mi[k * ld + j], acc);
        S.A2[i * ld + j] = -acc;
    }
    __syncthreads();
    // S = gA + gA^T (A3), gU = gA - gA^T (A0)
    for (int e = t; e < n * n; e += kSmallThreads) {
        const int i = e / n, j = e - i * n;
        const float a = S.A2[j * ld + i], at = S.A2[i * ld + j];
        S.A3[i * ld + j] = a + at;
        S.A0[i * ld + j] = a - at;
    }
    __syncthreads();
    // gWn = un-transposed [gU ; -2 G2 M^T + V S] -> gW (temporarily), dot = <gWn, W>
    double dacc = 0;
    for (int e = t; e < cout * cin; e += kSmallThreads) {   // global order
        const int o_ = e / cin, i2 = e - o_ * cin;
        const int r = tr ? i2 : o_, c = tr ? o_ : i2;
        float g;
        if (r < n) g = S.A0[r * ld + c];
        else {
            float a1 = 0.f, a2 = 0.f;
            for (int k = 0; k < n; ++k) { a1 = fmaf(gq(r, k), S.Mi[c * ld + k], a1); a2 = fmaf(S.Ws[r * ld + k], S.A3[c * ld + k], a2); }
            g = -2.f * a1 + a2;
        }
        gW[e] = g;
        dacc += (double)g * (double)W[e];
    }
    dacc = warp_sum(dacc);
    __syncthreads();
    if ((t & 31) == 0) S.red[t >> 5] = dacc;
    __syncthreads();
    if (t == 0) {
        double d = 0;
        for (int i = 0; i < kSmallThreads / 32; ++i) d += S.red[i];
        S.red[33] = d;
    }
    __syncthreads();
    const double dot = S.red[33];
    const float sc = S.sc[0], nrm_m = S.sc[1], nrm = nrm_m / mult;
    const float c2 = (float)(dot / ((double)nrm * nrm));
    for (size_t e = t; e < (size_t)cout * cin; e += kSmallThreads) gW[e] = sc * (gW[e] - W[e] * c2);
    if (t == 0) galpha[0] = (float)(dot / nrm_m);
}

__global__ void __launch_bounds__(kSmallThreads) cayley_small_bwd_kernel(const float* __restrict__ W, const float* __restrict__ alpha,
                                                                         int cout, int cin, float mult, const float* __restrict__ gQ,
                                                                         const float* __restrict__ M_saved, float* __restrict__ gW,
                                                                         float* __restrict__ galpha) {
    cayley_small_bwd_body(W, alpha, cout, cin, mult, gQ, M_saved, gW, galpha);
}

// Several small layers in ONE launch, one CTA per layer: the transforms of a step's layers do not depend on each other (forward:
// weights only; adjoint: each layer's dL/dQ), so a training step prepares all of them up front and differentiates all of them at
// the end instead of paying each single-CTA kernel's latency in sequence.
constexpr int kSmallBatchMax = 16;
struct SmallDesc {
    const float* W; const float* alpha; float* Q; float* M_save;          // forward
    const float* gQ; const float* M_saved; float* gW; float* galpha;      // adjoint
    int cout, cin; float mult;
};
struct SmallBatch { SmallDesc d[kSmallBatchMax]; };
__global__ void __launch_bounds__(kSmallThreads) cayley_small_fwd_batch_kernel(const __grid_constant__ SmallBatch b) {
    const SmallDesc& d = b.d[blockIdx.x];
    cayley_small_fwd_body(d.W, d.alpha, d.cout, d.cin, d.mult, d.Q, d.M_save);
}
__global__ void __launch_bounds__(kSmallThreads) cayley_small_bwd_batch_kernel(const __grid_constant__ SmallBatch b) {
    const SmallDesc& d = b.d[blockIdx.x];
    cayley_small_bwd_body(d.W, d.alpha, d.cout, d.cin, d.mult, d.gQ, d.M_saved, d.gW, d.galpha);
}

static int launch_cayley_small_fwd(const float* W, const float* alpha, int cout, int cin, float mult, float* Q, float* M_save,
                                   cudaStream_t st) {
    const int R = cout > cin ? cout : cin, n = cout > cin ? cin : cout;
    const size_t smem = small_smem_bytes(R, n, false);
    static std::atomic<bool> attr[kMaxDevices];
    if (first_use_on_device(attr))
        FIODE_CUDA(cudaFuncSetAttribute(cayley_small_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)small_smem_bytes(256, 64, false)));
    cayley_small_fwd_kernel<<<1, kSmallThreads, smem, st>>>(W, alpha, cout, cin, mult, Q, M_save);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
static int launch_cayley_small_bwd(const float* W, const float* alpha, int cout, int cin, float mult, const float* gQ,
                                   const float* M_saved, float* gW, float* galpha, cudaStream_t st) {
    const int R = cout > cin ? cout : cin, n = cout > cin ? cin : cout;
    const size_t smem = small_smem_bytes(R, n, true);
    static std::atomic<bool> attr[kMaxDevices];
    if (first_use_on_device(attr))
        FIODE_CUDA(cudaFuncSetAttribute(cayley_small_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)small_smem_bytes(256, 64, true)));
    cayley_small_bwd_kernel<<<1, kSmallThreads, smem, st>>>(W, alpha, cout, cin, mult, gQ, M_saved, gW, galpha);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

size_t cayley_linear_ws(int cout, int cin) {
    const size_t R = cout > cin ? cout : cin, n = cout > cin ? cin : cout;
    size_t b = 0;
    b += align256(kNormBlocks * sizeof(double)) + 256;   // partials + scale
    b += align256(R * n * 4);                            // Ws
    b += align256(n * (R - n) * 4 + 4);                  // Vt
    b += 6 * align256(n * n * 4);                        // G, IpA(X), Y, ImAT, Mt, spare
    b += align256(R * n * 4);                            // Qfull
    b += align256(gemm_nt_workspace((int)R, (int)n, (int)R));
    return b;
}

// Q = cayley(alpha * W / (||W||_F * norm_mult))
// What a training step keeps from the forward transform for its adjoint (fiode_cayley_linear_fwd_ctx / _bwd_ctx): one caller
// buffer holding the scale pair, (I + A)^-1, its transpose, the scaled working matrix Ws and V^T.  The forward call writes them in
// place of its workspace copies (no extra traffic); the adjoint then starts at the incoming gradient instead of re-deriving them
// (norm, staging, transposes: ~0.1 ms of dependent launches for the 4096 -> 512 layer).  Small layers keep (I + A)^-1 only.
struct CayleyCtx { float* scale; float* Minv; float* Mt; float* Ws; float* Vt; };
static size_t ctx_round(size_t floats) { return (floats + 63) & ~(size_t)63; }
static size_t cayley_ctx_floats(int cout, int cin) {
    const size_t R = cout > cin ? cout : cin, n = cout > cin ? cin : cout;
    if (cayley_small_host_ok(cout, cin)) return n * n;
    return 64 + 2 * ctx_round(n * n) + ctx_round(R * n) + ctx_round(n * (R - n) + 4);
}
static CayleyCtx cayley_ctx_carve(float* base, int cout, int cin) {
    const size_t R = cout > cin ? cout : cin, n = cout > cin ? cin : cout;
    CayleyCtx c;
    c.scale = base; base += 64;
    c.Minv = base; base += ctx_round(n * n);
    c.Mt = base; base += ctx_round(n * n);
    c.Ws = base; base += ctx_round(R * n);
    c.Vt = base;
    return c;
}

int cayley_linear_impl(const float* W, const float* alpha, int cout, int cin, float norm_mult, float* Q, void* ws,
                       size_t ws_bytes, cudaStream_t st, float* M_save, float* ctx_base = nullptr) {
    if (ws_bytes < cayley_linear_ws(cout, cin)) { set_error("cayley_linear: workspace too small"); return -1; }
    if (cayley_small_host_ok(cout, cin))
        return launch_cayley_small_fwd(W, alpha, cout, cin, norm_mult, Q, ctx_base ? ctx_base : M_save, st);
    const int tr = cin > cout;
    const int R = tr ? cin : cout, n = tr ? cout : cin;
    char* p = reinterpret_cast<char*>(ws);
    auto take = [&](size_t bytes) { void* r = p; p += align256(bytes); return r; };
    double* part = (double*)take(kNormBlocks * sizeof(double));
    float* scale = (float*)take(256);
    float* Ws = (float*)take((size_t)R * n * 4);
    float* Vt = (float*)take((size_t)n * (R - n) * 4 + 4);
    float* G = (float*)take((size_t)n * n * 4);
    float* X = (float*)take((size_t)n * n * 4);
    float* Y = (float*)take((size_t)n * n * 4);
    float* ImAT = (float*)take((size_t)n * n * 4);
    float* Mt = (float*)take((size_t)n * n * 4);
    take((size_t)n * n * 4);
    float* Qfull = (float*)take((size_t)R * n * 4);
    if (ctx_base) {                    // the adjoint's inputs are produced straight into the caller's context buffer
        const CayleyCtx c = cayley_ctx_carve(ctx_base, cout, cin);
        scale = c.scale; Ws = c.Ws; Vt = c.Vt; Mt = c.Mt; M_save = c.Minv;
    }
    float* gws = (float*)p;
    const size_t gws_bytes = ws_bytes - (size_t)(p - reinterpret_cast<char*>(ws));

    const size_t total = (size_t)cout * cin;
    sumsq_partial_kernel<<<kNormBlocks, 256, 0, st>>>(W, total, part);
    norm_finalize_kernel<<<1, 32, 0, st>>>(part, kNormBlocks, alpha, norm_mult, scale);
    const unsigned blocks = (unsigned)((total + 255) / 256 < 2048 ? (total + 255) / 256 : 2048);
    launch_tile_prep(W, cout, cin, tr, scale, Ws, Vt, st);
    count_launch(3);
    int rc;
    const int has_v = R > n;
    if (has_v && (rc = gemm_nt_gram(Vt, R - n, n, R - n, G, n, gws, gws_bytes, st))) return rc;
    cayley_prep2_kernel<<<(n * n + 255) / 256, 256, 0, st>>>(Ws, G, n, has_v, X, ImAT);
    count_launch(1);
    float* M = nullptr;
    if ((rc = gj_inverse(X, Y, n, &M, st))) return rc;
    if (M_save) FIODE_CUDA(cudaMemcpyAsync(M_save, M, (size_t)n * n * sizeof(float), cudaMemcpyDeviceToDevice, st));   // for the adjoint
    float* out_top = tr ? Qfull : Q;
    if ((rc = gemm_nt(M, n, ImAT, n, n, n, n, nullptr, 1.f, 0, out_top, n, gws, gws_bytes, st))) return rc;
    if (has_v || ctx_base) {
        transpose_kernel<<<dim3((n + 31) / 32, (n + 31) / 32), dim3(32, 8), 0, st>>>(M, n, n, Mt);
        count_launch(1);
    }
    if (has_v) {
        if ((rc = gemm_nt(Ws + (size_t)n * n, n, Mt, n, R - n, n, n, nullptr, -2.f, 0, out_top + (size_t)n * n, n, gws,
                          gws_bytes, st)))
            return rc;
    }
    if (tr) {   // Q (cout x cin) = Qfull^T, plain transpose (layers.py:54-55)
        transpose_kernel<<<dim3((n + 31) / 32, (R + 31) / 32), dim3(32, 8), 0, st>>>(Qfull, R, n, Q);
        count_launch(1);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "cayley_linear");
    return 0;
}

// ---- convolution weight spectrum ----------------------------------------------------------------
// wfft[phi][o][i] = sum_{a,b<k} w[o,i,a,b] exp(+2 pi i (u (a-s) + v (b-s)) / n), s = (k-1)/2  (App. C-1)
__device__ __forceinline__ float2 conv_spectrum_at(const float* __restrict__ w, int k, int n, int u, int v) {
    const int s = (k - 1) / 2;
    float2 acc = make_float2(0.f, 0.f);
    for (int a = 0; a < k; ++a)
        for (int b = 0; b < k; ++b) {
            int m = (u * (a - s) + v * (b - s)) % n;
            if (m < 0) m += n;
            float sn, cs;
            sincospif(2.f * (float)m / (float)n, &sn, &cs);
            const float wv = w[a * k + b];
            acc.x = fmaf(wv, cs, acc.x);
            acc.y = fmaf(wv, sn, acc.y);
        }
    return acc;
}

__global__ void __launch_bounds__(256) conv_spectrum_sumsq_kernel(const float* __restrict__ w, int cout, int cin,
                                                                  int k, int n, double* __restrict__ part) {
    __shared__ double sh[8];
    const int phi = blockIdx.x, H = n / 2 + 1, u = phi / H, v = phi - u * H;
    double acc = 0;
    for (int e = threadIdx.x; e < cout * cin; e += 256) {
        const float2 z = conv_spectrum_at(w + (size_t)e * k * k, k, n, u, v);
        acc += (double)z.x * z.x + (double)z.y * z.y;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int i = 0; i < 8; ++i) s += sh[i];
        part[phi] = s;
    }
}
__global__ void conv_alpha_finalize_kernel(const double* __restrict__ part, int F, float* __restrict__ alpha_out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double s = 0;
        for (int i = 0; i < F; ++i) s += part[i];
        alpha_out[0] = (float)sqrt(s);
    }
}

// One CTA per frequency: Q[phi] = cayley(alpha * wfft[phi] / ||wfft||).  cout >= cin.
// smem (float2): Wf[cout][cin], X[cin][cin], Y[cin][cin], A[cin][cin]
__global__ void __launch_bounds__(1024) cayley_conv_kernel(const float* __restrict__ w, const float* __restrict__ alpha,
                                                          const double* __restrict__ part, int F, int cout, int cin,
                                                          int k, int n, float2* __restrict__ Qf) {
    extern __shared__ float4 smem4[];
    float2* Wf = reinterpret_cast<float2*>(smem4);
    float2* X = Wf + cout * cin;
    float2* Y = X + cin * cin;
    float2* A = Y + cin * cin;
    __shared__ float sc_sh;
    const int nt = blockDim.x;   // 128 ... 1024 threads: one frequency per CTA, its matrix elements spread over the threads
    const int t = threadIdx.x, phi = blockIdx.x, H = n / 2 + 1, u = phi / H, v = phi - u * H;
    if (t < 32) {                               // the F per-frequency partials, by one warp (a single thread walking them
        double s = 0;                           // cost every CTA ~8 us before it could start)
        for (int i = t; i < F; i += 32) s += part[i];
        s = warp_sum(s);
        if (t == 0) sc_sh = alpha[0] / (float)sqrt(s);
    }
    __syncthreads();
    const float sc = sc_sh;
    for (int e = t; e < cout * cin; e += nt) {
        const float2 z = conv_spectrum_at(w + (size_t)e * k * k, k, n, u, v);
        Wf[e] = make_float2(sc * z.x, sc * z.y);
    }
    __syncthreads();
    const int nv = cout - cin;
    // A = U - U^H + V^H V ; X = I + A
    for (int e = t; e < cin * cin; e += nt) {
        const int i = e / cin, j = e - i * cin;
        const float2 uij = Wf[i * cin + j], uji = Wf[j * cin + i];
        float2 g = make_float2(0.f, 0.f);
        for (int r = 0; r < nv; ++r) cfma(g, cconj(Wf[(cin + r) * cin + i]), Wf[(cin + r) * cin + j]);
        const float2 a = make_float2(uij.x - uji.x + g.x, uij.y + uji.y + g.y);
        A[e] = a;
        X[e] = make_float2((i == j ? 1.f : 0.f) + a.x, a.y);
    }
    __syncthreads();
    // ping-pong Gauss-Jordan (complex, no pivoting)
    for (int kk = 0; kk < cin; ++kk) {
        const float2* in = (kk & 1) ? Y : X;
        float2* out = (kk & 1) ? X : Y;
        const float2 p = in[kk * cin + kk];
        const float den = 1.f / (p.x * p.x + p.y * p.y);
        const float2 ip = make_float2(p.x * den, -p.y * den);
        for (int e = t; e < cin * cin; e += nt) {
            const int i = e / cin, j = e - i * cin;
            float2 r;
            if (i == kk) r = (j == kk) ? ip : cmul(in[e], ip);
            else {
                const float2 f = in[i * cin + kk];
                if (j == kk) { const float2 q = cmul(f, ip); r = make_float2(-q.x, -q.y); }
                else { const float2 q = cmul(f, cmul(in[kk * cin + j], ip)); r = csub(in[e], q); }
            }
            out[e] = r;
        }
        __syncthreads();
    }
    const float2* M = (cin & 1) ? Y : X;
    float2* dst = Qf + (size_t)phi * cout * cin;
    // top: M (I - A) ; bottom: -2 V M
    for (int e = t; e < cout * cin; e += nt) {
        const int i = e / cin, j = e - i * cin;
        float2 acc = make_float2(0.f, 0.f);
        if (i < cin) {
            for (int r = 0; r < cin; ++r) {
                const float2 a = A[r * cin + j];
                const float2 ima = make_float2((r == j ? 1.f : 0.f) - a.x, -a.y);
                cfma(acc, M[i * cin + r], ima);
            }
        } else {
            for (int r = 0; r < cin; ++r) cfma(acc, Wf[i * cin + r], M[r * cin + j]);
            acc = make_float2(-2.f * acc.x, -2.f * acc.y);
        }
        dst[e] = acc;
    }
}


// ================================================================================================
// Backward of the Cayley weight preparation (training: libs/ortho_conv/layers.py:50-60 under autograd,
// reached from loss.backward() at pl_modules.py:504-506).  Forward intermediates are recomputed.
//   Q = [2M - I ; -2 V M],  M = (I + A)^-1,  A = U - U^H + V^H V,  [U; V] = sc * W,  sc = alpha / (||W|| mult)
//   gM = 2 G1 - 2 V^H G2 ;  gA = -M^H gM M^H ;  gU = gA - gA^H ;  gV = -2 G2 M^H + V (gA + gA^H)
//   gW = sc (gWn - W <gWn, W> / ||W||^2) ;  galpha = <gWn, W> / (||W|| mult)
// ================================================================================================
// G (R x n) = tr ? gQ^T : gQ ; Gt2 (n x (R-n)) = transpose of G[n:]
__global__ void cayley_bwd_prep_kernel(const float* __restrict__ gQ, int cout, int cin, int tr, float* __restrict__ G,
                                       float* __restrict__ Gt2) {
    const int R = tr ? cin : cout, n = tr ? cout : cin;
    const size_t total = (size_t)R * n;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int r = (int)(e / n), c = (int)(e - (size_t)r * n);
        const float v = tr ? gQ[(size_t)c * cin + r] : gQ[(size_t)r * cin + c];
        G[e] = v;
        if (r >= n) Gt2[(size_t)c * (R - n) + (r - n)] = v;
    }
}
// gM = 2 G1 - 2 P
__global__ void cayley_bwd_gm_kernel(const float* __restrict__ G1, const float* __restrict__ P, int nn, int has_v,
                                     float* __restrict__ gM) {
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < nn; e += gridDim.x * blockDim.x)
        gM[e] = 2.f * G1[e] - (has_v ? 2.f * P[e] : 0.f);
}
// from gAt (= gA^T): S = gA + gA^T, gU = gA - gA^T
__global__ void cayley_bwd_sym_kernel(const float* __restrict__ gAt, int n, float* __restrict__ S, float* __restrict__ gU) {
    const int total = n * n;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
        const int i = e / n, j = e - i * n;
        const float a = gAt[j * n + i], at = gAt[e];
        S[e] = a + at;
        gU[e] = a - at;
    }
}
// gWn (cout x cin) = un-transposed [gU ; gV1 + gV2], and per-block partial sums of <gWn, W>
__global__ void __launch_bounds__(256) cayley_bwd_assemble_kernel(const float* __restrict__ gU, const float* __restrict__ gV1,
                                                                  const float* __restrict__ gV2, const float* __restrict__ W,
                                                                  int cout, int cin, int tr, float* __restrict__ gWn,
                                                                  double* __restrict__ part) {
    __shared__ double sh[8];
    const int R = tr ? cin : cout, n = tr ? cout : cin;
    const size_t total = (size_t)cout * cin;
    double acc = 0;
    for (size_t e = blockIdx.x * 256 + threadIdx.x; e < total; e += (size_t)gridDim.x * 256) {
        const int o = (int)(e / cin), i = (int)(e - (size_t)o * cin);
        const int r = tr ? i : o, c = tr ? o : i;       // position inside the R x n working matrix
        float g;
        if (r < n) g = gU[(size_t)r * n + c];
        else { const size_t q = (size_t)(r - n) * n + c; g = gV1[q] + gV2[q]; }
        gWn[e] = g;
        acc += (double)g * (double)W[e];
    }
    (void)R;
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int i = 0; i < 8; ++i) s += sh[i];
        part[blockIdx.x] = s;
    }
}
// the same for a transposed layer (cin > cout: gWn = [gU ; gV1 + gV2]^T) through 32 x 32 tiles, so that the reads of gU / gV and the
// writes of gWn both run along contiguous rows; CTA b takes tiles b, b + gridDim.x, ... and leaves one partial of <gWn, W>
__global__ void __launch_bounds__(1024) cayley_bwd_assemble_tr_kernel(const float* __restrict__ gU, const float* __restrict__ gV1,
                                                                     const float* __restrict__ gV2, const float* __restrict__ W,
                                                                     int R, int n, float* __restrict__ gWn,
                                                                     double* __restrict__ part) {
    __shared__ float tile[32][33];
    __shared__ double sh[32];
    const int tc = (n + 31) / 32, tr_ = (R + 31) / 32, x = threadIdx.x & 31, y = threadIdx.x >> 5;   // 32 x 32 threads: one tile element each
    double acc = 0;
    for (int tI = blockIdx.x; tI < tc * tr_; tI += gridDim.x) {
        const int r0 = (tI / tc) * 32, c0 = (tI % tc) * 32;
        for (int rl = y; rl < 32; rl += 32) {
            const int r = r0 + rl, c = c0 + x;
            if (r < R && c < n) {
                float g;
                if (r < n) g = gU[(size_t)r * n + c];
                else { const size_t q = (size_t)(r - n) * n + c; g = gV1[q] + gV2[q]; }
                tile[rl][x] = g;
            }
        }
        __syncthreads();
        for (int cl = y; cl < 32; cl += 32) {
            const int c = c0 + cl, r = r0 + x;             // gWn is (cout = n) x (cin = R): row c, column r
            if (c < n && r < R) {
                const float g = tile[x][cl];
                const size_t e = (size_t)c * R + r;
                gWn[e] = g;
                acc += (double)g * (double)W[e];
            }
        }
        __syncthreads();
    }
    acc = warp_sum(acc);
    if (x == 0) sh[y] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s2 = 0;
        for (int i = 0; i < 32; ++i) s2 += sh[i];
        part[blockIdx.x] = s2;
    }
}
// gW = sc (gWn - W dot / ||W||^2) ; galpha = dot / (||W|| mult)      scale[0] = alpha / (||W|| mult), scale[1] = ||W|| mult
__global__ void cayley_bwd_finalize_kernel(const float* __restrict__ gWn, const float* __restrict__ W, size_t total,
                                           const double* __restrict__ part, int np, const float* __restrict__ scale,
                                           float mult, float* __restrict__ gW, float* __restrict__ galpha) {
    double dot = 0;
    for (int i = 0; i < np; ++i) dot += part[i];
    const float sc = scale[0], nrm_m = scale[1], nrm = nrm_m / mult;
    const float c2 = (float)(dot / ((double)nrm * nrm));
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x)
        gW[e] = sc * (gWn[e] - W[e] * c2);
    if (blockIdx.x == 0 && threadIdx.x == 0) galpha[0] = (float)(dot / nrm_m);
}

size_t cayley_linear_bwd_ws(int cout, int cin) {
    const size_t R = cout > cin ? cout : cin, n = cout > cin ? cin : cout;
    size_t b = cayley_linear_ws(cout, cin);
    b += 2 * align256(R * n * 4);                         // G, gWn
    b += align256(n * (R - n) * 4 + 4);                   // Gt2
    b += 2 * align256((R - n) * n * 4 + 4);               // gV1, gV2
    b += 8 * align256(n * n * 4);                         // P, gM, T1t, gAt, S, gU, Mt, spare
    b += align256(gemm_nt_workspace((int)R, (int)n, (int)n));   // split-K partials of the product that runs on the branch
    return b;
}

int cayley_linear_bwd_impl(const float* W, const float* alpha, int cout, int cin, float norm_mult, const float* gQ,
                           float* gW, float* galpha, void* ws, size_t ws_bytes, cudaStream_t st, const float* M_saved,
                           const float* ctx_base = nullptr) {
    if (ws_bytes < cayley_linear_bwd_ws(cout, cin)) { set_error("cayley_linear_bwd: workspace too small"); return -1; }
    if (cayley_small_host_ok(cout, cin))
        return launch_cayley_small_bwd(W, alpha, cout, cin, norm_mult, gQ, ctx_base ? ctx_base : M_saved, gW, galpha, st);
    const int tr = cin > cout;
    const int R = tr ? cin : cout, n = tr ? cout : cin;
    char* p = reinterpret_cast<char*>(ws);
    auto take = [&](size_t bytes) { void* r = p; p += align256(bytes); return r; };
    double* part = (double*)take(kNormBlocks * sizeof(double));
    float* scale = (float*)take(256);
    float* Ws = (float*)take((size_t)R * n * 4);
    float* Vt = (float*)take((size_t)n * (R - n) * 4 + 4);
    float* Gm = (float*)take((size_t)n * n * 4);          // V^T V
    float* X = (float*)take((size_t)n * n * 4);
    float* Y = (float*)take((size_t)n * n * 4);
    float* ImAT = (float*)take((size_t)n * n * 4);
    float* Mt = (float*)take((size_t)n * n * 4);
    float* G = (float*)take((size_t)R * n * 4);
    float* gWn = (float*)take((size_t)R * n * 4);
    float* Gt2 = (float*)take((size_t)n * (R - n) * 4 + 4);
    float* gV1 = (float*)take((size_t)(R - n) * n * 4 + 4);
    float* gV2 = (float*)take((size_t)(R - n) * n * 4 + 4);
    float* P = (float*)take((size_t)n * n * 4);
    float* gM = (float*)take((size_t)n * n * 4);
    float* T1t = (float*)take((size_t)n * n * 4);
    float* gAt = (float*)take((size_t)n * n * 4);
    float* S = (float*)take((size_t)n * n * 4);
    float* gU = (float*)take((size_t)n * n * 4);
    const size_t gws2_bytes = align256(gemm_nt_workspace(R, n, n));
    float* gws2 = (float*)take(gws2_bytes);
    float* gws = (float*)p;
    const size_t gws_bytes = ws_bytes - (size_t)(p - reinterpret_cast<char*>(ws));
    if (gws_bytes < gemm_nt_workspace(R, n, R)) { set_error("cayley_linear_bwd: gemm workspace too small"); return -1; }

    // ---- recompute the forward up to M (or take it from the forward call's context buffer) -----------------------
    const size_t total = (size_t)cout * cin;
    const unsigned blocks = (unsigned)((total + 255) / 256 < 2048 ? (total + 255) / 256 : 2048);
    if (ctx_base) {
        const CayleyCtx c = cayley_ctx_carve(const_cast<float*>(ctx_base), cout, cin);
        scale = c.scale; Ws = c.Ws; Vt = c.Vt; Mt = c.Mt; M_saved = c.Minv;
    } else {
        sumsq_partial_kernel<<<kNormBlocks, 256, 0, st>>>(W, total, part);
        norm_finalize_kernel<<<1, 32, 0, st>>>(part, kNormBlocks, alpha, norm_mult, scale);
        launch_tile_prep(W, cout, cin, tr, scale, Ws, Vt, st);
        count_launch(3);
    }
    int rc;
    const int has_v = R > n;
    const int nb = (n * n + 255) / 256;
    const float* M = M_saved;                  // (I + A)^-1 kept by the forward call, else inverted again here
    if (!M) {                                  // (A = U - U^T + V^T V is needed for that only)
        if (has_v && (rc = gemm_nt(Vt, R - n, Vt, R - n, n, n, R - n, nullptr, 1.f, 0, Gm, n, gws, gws_bytes, st))) return rc;
        cayley_prep2_kernel<<<nb, 256, 0, st>>>(Ws, Gm, n, has_v, X, ImAT);
        count_launch(1);
        float* Minv = nullptr;
        if ((rc = gj_inverse(X, Y, n, &Minv, st))) return rc;
        M = Minv;
    }
    if (!ctx_base) {
        transpose_kernel<<<dim3((n + 31) / 32, (n + 31) / 32), dim3(32, 8), 0, st>>>(M, n, n, Mt);
        count_launch(1);
    }
    // ---- backward -----------------------------------------------------------------------------------
    launch_tile_prep(gQ, cout, cin, tr, nullptr, G, Gt2, st);
    count_launch(1);
    const float* V = Ws + (size_t)n * n;
    const float* G2 = G + (size_t)n * n;
    // -2 G2 M^T needs the incoming gradient and M only: on a branch, beside the chain P -> gM -> gA -> V (gA + gA^T)
    cudaStream_t sd = st;
    if (has_v) {
        sd = fork_side(st);
        if ((rc = gemm_nt_tall(G2, n, M, n, R - n, n, n, -2.f, gV1, n, gws2, gws2_bytes, sd))) return rc;
    }
    if (has_v && (rc = gemm_nt(Vt, R - n, Gt2, R - n, n, n, R - n, nullptr, 1.f, 0, P, n, gws, gws_bytes, st))) return rc;
    cayley_bwd_gm_kernel<<<nb, 256, 0, st>>>(G, P, n * n, has_v, gM);
    count_launch(1);
    if ((rc = gemm_nt(M, n, gM, n, n, n, n, nullptr, 1.f, 0, T1t, n, gws, gws_bytes, st))) return rc;       // (gM M^T)^T
    if ((rc = gemm_nt(T1t, n, Mt, n, n, n, n, nullptr, -1.f, 0, gAt, n, gws, gws_bytes, st))) return rc;    // gA^T = -T1^T M
    cayley_bwd_sym_kernel<<<nb, 256, 0, st>>>(gAt, n, S, gU);
    count_launch(1);
    if (has_v) {
        if ((rc = gemm_nt_tall(V, n, S, n, R - n, n, n, 1.f, gV2, n, gws, gws_bytes, st))) return rc;     // V (gA + gA^T)
        if (sd != st && (rc = join_side(st))) return rc;
    }
    if (tr) cayley_bwd_assemble_tr_kernel<<<kNormBlocks, 1024, 0, st>>>(gU, gV1, gV2, W, R, n, gWn, part);
    else cayley_bwd_assemble_kernel<<<kNormBlocks, 256, 0, st>>>(gU, gV1, gV2, W, cout, cin, tr, gWn, part);
    cayley_bwd_finalize_kernel<<<blocks, 256, 0, st>>>(gWn, W, total, part, kNormBlocks, scale, norm_mult, gW, galpha);
    count_launch(2);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "cayley_linear_bwd");
    return 0;
}

// ---- convolution (k > 1): one CTA per frequency, everything in shared memory ---------------------------
// Writes gWn[phi] (cout x cin complex, gradient w.r.t. the scaled spectrum sc * wfft) and dotp[phi] = Re <gWn[phi], wfft[phi]>.
__global__ void __launch_bounds__(1024) cayley_conv_bwd_kernel(const float* __restrict__ w, const float* __restrict__ alpha,
                                                              const double* __restrict__ part, int F, int cout, int cin,
                                                              int k, int n, const float2* __restrict__ gQf,
                                                              float2* __restrict__ gWn, double* __restrict__ dotp) {
    extern __shared__ float4 smem4[];
    float2* Wf = reinterpret_cast<float2*>(smem4);       // [cout][cin] scaled spectrum
    float2* X = Wf + cout * cin;
    float2* Y = X + cin * cin;
    float2* Gq = Y + cin * cin;                           // [cout][cin] incoming gradient
    float2* gM = Gq + cout * cin;
    float2* T = gM + cin * cin;
    float2* gA = T + cin * cin;
    __shared__ float sc_sh;
    __shared__ double red[32];
    const int nt = blockDim.x;   // 128 ... 1024 threads: one frequency per CTA, its matrix elements spread over the threads
    const int t = threadIdx.x, phi = blockIdx.x, H = n / 2 + 1, u = phi / H, v = phi - u * H;
    if (t < 32) {                               // the F per-frequency partials, by one warp (a single thread walking them
        double s = 0;                           // cost every CTA ~8 us before it could start)
        for (int i = t; i < F; i += 32) s += part[i];
        s = warp_sum(s);
        if (t == 0) sc_sh = alpha[0] / (float)sqrt(s);
    }
    __syncthreads();
    const float sc = sc_sh;
    const float2* gsrc = gQf + (size_t)phi * cout * cin;
    for (int e = t; e < cout * cin; e += nt) {
        const float2 z = conv_spectrum_at(w + (size_t)e * k * k, k, n, u, v);
        Wf[e] = make_float2(sc * z.x, sc * z.y);
        Gq[e] = gsrc[e];
    }
    __syncthreads();
    const int nv = cout - cin;
    for (int e = t; e < cin * cin; e += nt) {
        const int i = e / cin, j = e - i * cin;
        const float2 uij = Wf[i * cin + j], uji = Wf[j * cin + i];
        float2 g = make_float2(0.f, 0.f);
        for (int r = 0; r < nv; ++r) cfma(g, cconj(Wf[(cin + r) * cin + i]), Wf[(cin + r) * cin + j]);
        X[e] = make_float2((i == j ? 1.f : 0.f) + uij.x - uji.x + g.x, uij.y + uji.y + g.y);
    }
    __syncthreads();
    for (int kk = 0; kk < cin; ++kk) {
        const float2* in = (kk & 1) ? Y : X;
        float2* out = (kk & 1) ? X : Y;
        const float2 pv = in[kk * cin + kk];
        const float den = 1.f / (pv.x * pv.x + pv.y * pv.y);
        const float2 ip = make_float2(pv.x * den, -pv.y * den);
        for (int e = t; e < cin * cin; e += nt) {
            const int i = e / cin, j = e - i * cin;
            float2 r;
            if (i == kk) r = (j == kk) ? ip : cmul(in[e], ip);
            else {
                const float2 f = in[i * cin + kk];
                if (j == kk) { const float2 q = cmul(f, ip); r = make_float2(-q.x, -q.y); }
                else { const float2 q = cmul(f, cmul(in[kk * cin + j], ip)); r = csub(in[e], q); }
            }
            out[e] = r;
        }
        __syncthreads();
    }
    const float2* M = (cin & 1) ? Y : X;
    const float2* V = Wf + cin * cin;
    const float2* G2 = Gq + cin * cin;
    // gM = 2 G1 - 2 V^H G2
    for (int e = t; e < cin * cin; e += nt) {
        const int i = e / cin, j = e - i * cin;
        float2 acc = make_float2(0.f, 0.f);
        for (int r = 0; r < nv; ++r) cfma(acc, cconj(V[r * cin + i]), G2[r * cin + j]);
        gM[e] = make_float2(2.f * (Gq[e].x - acc.x), 2.f * (Gq[e].y - acc.y));
    }
    __syncthreads();
    // T = gM M^H.  Consecutive threads hold consecutive j, so M[j][r] at a common r is a walk down a column: one bank for the
    // whole warp.  Each thread therefore starts its sum at r = j and wraps around (the lanes then touch 32 different columns);
    // the order of the sum differs per element but is fixed.
    for (int e = t; e < cin * cin; e += nt) {
        const int i = e / cin, j = e - i * cin;
        float2 acc = make_float2(0.f, 0.f);
        int r = j;
        for (int q = 0; q < cin; ++q) {
            cfma(acc, gM[i * cin + r], cconj(M[j * cin + r]));
            if (++r == cin) r = 0;
        }
        T[e] = acc;
    }
    __syncthreads();
    // gA = -M^H T
    for (int e = t; e < cin * cin; e += nt) {
        const int i = e / cin, j = e - i * cin;
        float2 acc = make_float2(0.f, 0.f);
        for (int r = 0; r < cin; ++r) cfma(acc, cconj(M[r * cin + i]), T[r * cin + j]);
        gA[e] = make_float2(-acc.x, -acc.y);
    }
    __syncthreads();
    // gU = gA - gA^H ; gV = -2 G2 M^H + V (gA + gA^H) ; dot with the unscaled spectrum
    double dacc = 0;
    float2* dst = gWn + (size_t)phi * cout * cin;
    const float inv_sc = 1.f / sc;
    for (int e = t; e < cout * cin; e += nt) {
        const int i = e / cin, j = e - i * cin;
        float2 g;
        if (i < cin) {
            const float2 a = gA[i * cin + j], b = gA[j * cin + i];
            g = make_float2(a.x - b.x, a.y + b.y);
        } else {
            const int r = i - cin;
            float2 a1 = make_float2(0.f, 0.f), a2 = make_float2(0.f, 0.f);
            for (int q = 0; q < cin; ++q) {
                cfma(a1, G2[r * cin + q], cconj(M[j * cin + q]));
                const float2 x = gA[q * cin + j], y = gA[j * cin + q];
                cfma(a2, V[r * cin + q], make_float2(x.x + y.x, x.y - y.y));
            }
            g = make_float2(a2.x - 2.f * a1.x, a2.y - 2.f * a1.y);
        }
        dst[e] = g;
        dacc += ((double)g.x * Wf[e].x + (double)g.y * Wf[e].y) * inv_sc;
    }
    dacc = warp_sum(dacc);
    if ((t & 31) == 0) red[t >> 5] = dacc;
    __syncthreads();
    if (t == 0) {
        double s = 0;
        for (int i = 0; i < nt / 32; ++i) s += red[i];
        dotp[phi] = s;
    }
}
// gw[o,i,a,b] = sum_phi Re( conj(e^{i theta}) * sc (gWn - wfft D / ||wfft||^2) ) ; galpha = D / ||wfft||
__global__ void __launch_bounds__(128) cayley_conv_bwd_taps_kernel(const float* __restrict__ w, const float* __restrict__ alpha,
                                                                   const double* __restrict__ part,
                                                                   const double* __restrict__ dotp, int F, int cout, int cin,
                                                                   int k, int n, const float2* __restrict__ gWn,
                                                                   float* __restrict__ gw, float* __restrict__ galpha) {
    __shared__ double sh[2];
    __shared__ double tap[4][9];
    const int t = threadIdx.x, e = blockIdx.x, H = n / 2 + 1, s = (k - 1) / 2;   // e = o * cin + i
    if (t < 2) {
        double a = 0;
        const double* src = t ? dotp : part;
        for (int i = 0; i < F; ++i) a += src[i];
        sh[t] = a;
    }
    __syncthreads();
    const double nrm2 = sh[0], D = sh[1];
    const float sc = alpha[0] / (float)sqrt(nrm2);
    const float c2 = (float)(D / nrm2);
    double acc[9];
    for (int q = 0; q < 9; ++q) acc[q] = 0;
    for (int phi = t; phi < F; phi += 128) {
        const int u = phi / H, v = phi - u * H;
        const float2 z = conv_spectrum_at(w + (size_t)e * k * k, k, n, u, v);
        const float2 g0 = gWn[(size_t)phi * cout * cin + e];
        const float gx = sc * (g0.x - z.x * c2), gy = sc * (g0.y - z.y * c2);
        for (int a = 0; a < k; ++a)
            for (int b = 0; b < k; ++b) {
                int m = (u * (a - s) + v * (b - s)) % n;
                if (m < 0) m += n;
                float sn, cs;
                sincospif(2.f * (float)m / (float)n, &sn, &cs);
                acc[a * k + b] += (double)(gx * cs + gy * sn);
            }
    }
    for (int q = 0; q < k * k; ++q) {
        const double r = warp_sum(acc[q]);
        if ((t & 31) == 0) tap[t >> 5][q] = r;
    }
    __syncthreads();
    if (t < k * k) gw[(size_t)e * k * k + t] = (float)(tap[0][t] + tap[1][t] + tap[2][t] + tap[3][t]);
    if (e == 0 && t == 0) galpha[0] = (float)(D / sqrt(nrm2));
}

}  // namespace fiode

using namespace fiode;

extern "C" size_t fiode_cayley_linear_workspace_bytes(int cout, int cin) { return cayley_linear_ws(cout, cin); }

extern "C" int fiode_cayley_linear(const float* W, const float* alpha, int cout, int cin, float* Q, void* ws,
                                   size_t ws_bytes, void* stream) {
    if (!W || !alpha || !Q || !ws) { set_error("fiode_cayley_linear: null pointer"); return -1; }
    if (cout <= 0 || cin <= 0) { set_error("fiode_cayley_linear: bad shape"); return -1; }
    return cayley_linear_impl(W, alpha, cout, cin, 1.f, Q, ws, ws_bytes, (cudaStream_t)stream, nullptr);
}
extern "C" int fiode_cayley_linear_save(const float* W, const float* alpha, int cout, int cin, float* Q, float* Minv,
                                        void* ws, size_t ws_bytes, void* stream) {
    if (!W || !alpha || !Q || !Minv || !ws) { set_error("fiode_cayley_linear_save: null pointer"); return -1; }
    if (cout <= 0 || cin <= 0) { set_error("fiode_cayley_linear_save: bad shape"); return -1; }
    return cayley_linear_impl(W, alpha, cout, cin, 1.f, Q, ws, ws_bytes, (cudaStream_t)stream, Minv);
}

static size_t conv_part_bytes(int n) { return align256((size_t)n * (n / 2 + 1) * sizeof(double)); }

extern "C" int fiode_conv_alpha_init(const float* w, int cout, int cin, int k, int n, float* alpha_out, void* ws,
                                     size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!w || !alpha_out || !ws) { set_error("fiode_conv_alpha_init: null pointer"); return -1; }
    if (ws_bytes < conv_part_bytes(n)) { set_error("fiode_conv_alpha_init: workspace too small"); return -1; }
    const int F = n * (n / 2 + 1);
    conv_spectrum_sumsq_kernel<<<F, 256, 0, st>>>(w, cout, cin, k, n, (double*)ws);
    conv_alpha_finalize_kernel<<<1, 32, 0, st>>>((const double*)ws, F, alpha_out);
    FIODE_CUDA(cudaGetLastError());
    count_launch(2);
    return 0;
}

// threads per frequency CTA of the two kernels above: the 64 x 64 layer's CTA fills an SM's shared memory (seven complex
// matrices), so how long it sits there decides how long every other kernel of the step is kept off that SM
static int conv_cayley_threads(int cout, int cin) {
    const int e = cout * cin;
    return e >= 4096 ? 1024 : e >= 1024 ? 512 : e >= 256 ? 256 : 128;
}
extern "C" size_t fiode_cayley_conv_workspace_bytes(int cout, int cin, int k, int n) {
    if (k == 1) return cayley_linear_ws(cout, cin);
    return conv_part_bytes(n);
}

extern "C" int fiode_cayley_conv(const float* w, const float* alpha, int cout, int cin, int k, int n, float* Qf,
                                 void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!w || !alpha || !Qf || !ws) { set_error("fiode_cayley_conv: null pointer"); return -1; }
    const int F = n * (n / 2 + 1);
    if (k == 1) {
        // wfft[phi] = w[:, :, 0, 0] for every phi  =>  ||wfft||_F = sqrt(F) ||w||_F, one real Q (App. C-1)
        return cayley_linear_impl(w, alpha, cout, cin, sqrtf((float)F), Qf, ws, ws_bytes, st, nullptr);
    }
    if (cout < cin) { set_error("fiode_cayley_conv: k>1 needs cout >= cin (got %d < %d)", cout, cin); return -1; }
    if (ws_bytes < conv_part_bytes(n)) { set_error("fiode_cayley_conv: workspace too small"); return -1; }
    const size_t smem = (size_t)(cout * cin + 3 * cin * cin) * sizeof(float2);
    if (smem > 200 * 1024) { set_error("fiode_cayley_conv: layer too large for the in-smem solver"); return -1; }
    FIODE_CUDA(cudaFuncSetAttribute(cayley_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    conv_spectrum_sumsq_kernel<<<F, 256, 0, st>>>(w, cout, cin, k, n, (double*)ws);
    cayley_conv_kernel<<<F, conv_cayley_threads(cout, cin), smem, st>>>(w, alpha, (const double*)ws, F, cout, cin, k, n,
                                             reinterpret_cast<float2*>(Qf));
    FIODE_CUDA(cudaGetLastError());
    count_launch(2);
    return 0;
}

// ---- backward of the weight preparation (training) ---------------------------------------------------
extern "C" size_t fiode_cayley_linear_bwd_workspace_bytes(int cout, int cin) { return cayley_linear_bwd_ws(cout, cin); }

extern "C" int fiode_cayley_linear_bwd(const float* W, const float* alpha, int cout, int cin, const float* gQ, float* gW,
                                       float* galpha, void* ws, size_t ws_bytes, void* stream) {
    if (!W || !alpha || !gQ || !gW || !galpha || !ws) { set_error("fiode_cayley_linear_bwd: null pointer"); return -1; }
    if (cout <= 0 || cin <= 0) { set_error("fiode_cayley_linear_bwd: bad shape"); return -1; }
    return cayley_linear_bwd_impl(W, alpha, cout, cin, 1.f, gQ, gW, galpha, ws, ws_bytes, (cudaStream_t)stream, nullptr);
}
extern "C" int fiode_cayley_linear_bwd_saved(const float* W, const float* alpha, int cout, int cin, const float* gQ,
                                             const float* Minv, float* gW, float* galpha, void* ws, size_t ws_bytes,
                                             void* stream) {
    if (!W || !alpha || !gQ || !Minv || !gW || !galpha || !ws) { set_error("fiode_cayley_linear_bwd_saved: null pointer"); return -1; }
    if (cout <= 0 || cin <= 0) { set_error("fiode_cayley_linear_bwd_saved: bad shape"); return -1; }
    return cayley_linear_bwd_impl(W, alpha, cout, cin, 1.f, gQ, gW, galpha, ws, ws_bytes, (cudaStream_t)stream, Minv);
}

extern "C" size_t fiode_cayley_conv_bwd_workspace_bytes(int cout, int cin, int k, int n) {
    if (k == 1) return cayley_linear_bwd_ws(cout, cin);
    const size_t F = (size_t)n * (n / 2 + 1);
    return 2 * conv_part_bytes(n) + align256(F * cout * cin * sizeof(float2));
}

extern "C" int fiode_cayley_conv_bwd(const float* w, const float* alpha, int cout, int cin, int k, int n, const float* gQf,
                                     float* gw, float* galpha, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!w || !alpha || !gQf || !gw || !galpha || !ws) { set_error("fiode_cayley_conv_bwd: null pointer"); return -1; }
    const int F = n * (n / 2 + 1);
    if (k == 1) return cayley_linear_bwd_impl(w, alpha, cout, cin, sqrtf((float)F), gQf, gw, galpha, ws, ws_bytes, st, nullptr);
    if (cout < cin || k * k > 9) { set_error("fiode_cayley_conv_bwd: needs cout >= cin and k <= 3"); return -1; }
    if (ws_bytes < fiode_cayley_conv_bwd_workspace_bytes(cout, cin, k, n)) { set_error("fiode_cayley_conv_bwd: workspace too small"); return -1; }
    char* p = reinterpret_cast<char*>(ws);
    double* part = (double*)p; p += conv_part_bytes(n);
    double* dotp = (double*)p; p += conv_part_bytes(n);
    float2* gWn = (float2*)p;
    const size_t smem = (size_t)(2 * cout * cin + 5 * cin * cin) * sizeof(float2);
    if (smem > 200 * 1024) { set_error("fiode_cayley_conv_bwd: layer too large for the in-smem solver"); return -1; }
    FIODE_CUDA(cudaFuncSetAttribute(cayley_conv_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    conv_spectrum_sumsq_kernel<<<F, 256, 0, st>>>(w, cout, cin, k, n, part);
    cayley_conv_bwd_kernel<<<F, conv_cayley_threads(cout, cin), smem, st>>>(w, alpha, part, F, cout, cin, k, n, reinterpret_cast<const float2*>(gQf), gWn, dotp);
    cayley_conv_bwd_taps_kernel<<<cout * cin, 128, 0, st>>>(w, alpha, part, dotp, F, cout, cin, k, n, gWn, gw, galpha);
    FIODE_CUDA(cudaGetLastError());
    count_launch(3);
    return 0;
}

// ---- batched small layers (one launch for all of a step's small Cayley transforms / adjoints) -----------------------------
extern "C" int fiode_cayley_small_ok(int cout, int cin) { return cayley_small_host_ok(cout, cin) ? 1 : 0; }

static int small_batch_check(const char* who, int count, const float* const* W, const int* cout, const int* cin) {
    if (count <= 0 || !W || !cout || !cin) { set_error("%s: bad arguments", who); return -1; }
    for (int i = 0; i < count; ++i)
        if (!W[i] || cout[i] <= 0 || cin[i] <= 0 || !cayley_small_host_ok(cout[i], cin[i])) {
            set_error("%s: layer %d is not a small layer (min(cout, cin) <= 64, max * min <= 16384)", who, i);
            return -1;
        }
    return 0;
}

extern "C" int fiode_cayley_small_batch_fwd(int count, const float* const* W, const float* const* alpha, const int* cout,
                                            const int* cin, const float* mult, float* const* Q, float* const* Minv,
                                            void* stream) {
    if (small_batch_check("fiode_cayley_small_batch_fwd", count, W, cout, cin)) return -1;
    if (!Q) { set_error("fiode_cayley_small_batch_fwd: null output list"); return -1; }
    cudaStream_t st = (cudaStream_t)stream;
    static std::atomic<bool> attr[kMaxDevices];
    if (first_use_on_device(attr))
        FIODE_CUDA(cudaFuncSetAttribute(cayley_small_fwd_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)small_smem_bytes(256, 64, false)));
    for (int base = 0; base < count; base += kSmallBatchMax) {
        const int m = count - base < kSmallBatchMax ? count - base : kSmallBatchMax;
        SmallBatch b{};
        size_t smem = 0;
        for (int i = 0; i < m; ++i) {
            const int j = base + i;
            if (!Q[j]) { set_error("fiode_cayley_small_batch_fwd: null output"); return -1; }
            SmallDesc& d = b.d[i];
            d.W = W[j]; d.alpha = alpha ? alpha[j] : nullptr; d.Q = Q[j]; d.M_save = Minv ? Minv[j] : nullptr;
            d.cout = cout[j]; d.cin = cin[j]; d.mult = mult ? mult[j] : 1.f;
            const int R = d.cout > d.cin ? d.cout : d.cin, n = d.cout > d.cin ? d.cin : d.cout;
            const size_t s = small_smem_bytes(R, n, false);
            if (s > smem) smem = s;
        }
        cayley_small_fwd_batch_kernel<<<m, kSmallThreads, smem, st>>>(b);
        FIODE_CUDA(cudaGetLastError());
        count_launch(1);
    }
    return 0;
}

extern "C" int fiode_cayley_small_batch_bwd(int count, const float* const* W, const float* const* alpha, const int* cout,
                                            const int* cin, const float* mult, const float* const* gQ,
                                            const float* const* Minv, float* const* gW, float* const* galpha, void* stream) {
    if (small_batch_check("fiode_cayley_small_batch_bwd", count, W, cout, cin)) return -1;
    if (!gQ || !gW || !galpha) { set_error("fiode_cayley_small_batch_bwd: null list"); return -1; }
    cudaStream_t st = (cudaStream_t)stream;
    static std::atomic<bool> attr[kMaxDevices];
    if (first_use_on_device(attr))
        FIODE_CUDA(cudaFuncSetAttribute(cayley_small_bwd_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)small_smem_bytes(256, 64, true)));
    for (int base = 0; base < count; base += kSmallBatchMax) {
        const int m = count - base < kSmallBatchMax ? count - base : kSmallBatchMax;
        SmallBatch b{};
        size_t smem = 0;
        for (int i = 0; i < m; ++i) {
            const int j = base + i;
            if (!gQ[j] || !gW[j] || !galpha[j]) { set_error("fiode_cayley_small_batch_bwd: null pointer"); return -1; }
            SmallDesc& d = b.d[i];
            d.W = W[j]; d.alpha = alpha ? alpha[j] : nullptr; d.gQ = gQ[j]; d.M_saved = Minv ? Minv[j] : nullptr;
            d.gW = gW[j]; d.galpha = galpha[j];
            d.cout = cout[j]; d.cin = cin[j]; d.mult = mult ? mult[j] : 1.f;
            const int R = d.cout > d.cin ? d.cout : d.cin, n = d.cout > d.cin ? d.cin : d.cout;
            const size_t s = small_smem_bytes(R, n, true);
            if (s > smem) smem = s;
        }
        cayley_small_bwd_batch_kernel<<<m, kSmallThreads, smem, st>>>(b);
        FIODE_CUDA(cudaGetLastError());
        count_launch(1);
    }
    return 0;
}

// ---- forward transform that keeps its intermediates for the adjoint (training) ----------------------------------------
extern "C" size_t fiode_cayley_linear_ctx_floats(int cout, int cin) {
    if (cout <= 0 || cin <= 0) return 0;
    return cayley_ctx_floats(cout, cin);
}
extern "C" int fiode_cayley_linear_fwd_ctx(const float* W, const float* alpha, int cout, int cin, float* Q, float* ctx, void* ws,
                                           size_t ws_bytes, void* stream) {
    if (!W || !alpha || !Q || !ctx || !ws) { set_error("fiode_cayley_linear_fwd_ctx: null pointer"); return -1; }
    if (cout <= 0 || cin <= 0) { set_error("fiode_cayley_linear_fwd_ctx: bad shape"); return -1; }
    if (reinterpret_cast<uintptr_t>(ctx) & 15) { set_error("fiode_cayley_linear_fwd_ctx: ctx must be 16-byte aligned"); return -1; }
    return cayley_linear_impl(W, alpha, cout, cin, 1.f, Q, ws, ws_bytes, (cudaStream_t)stream, nullptr, ctx);
}
extern "C" int fiode_cayley_linear_bwd_ctx(const float* W, const float* alpha, int cout, int cin, const float* gQ,
                                           const float* ctx, float* gW, float* galpha, void* ws, size_t ws_bytes, void* stream) {
    if (!W || !alpha || !gQ || !ctx || !gW || !galpha || !ws) { set_error("fiode_cayley_linear_bwd_ctx: null pointer"); return -1; }
    if (cout <= 0 || cin <= 0) { set_error("fiode_cayley_linear_bwd_ctx: bad shape"); return -1; }
    return cayley_linear_bwd_impl(W, alpha, cout, cin, 1.f, gQ, gW, galpha, ws, ws_bytes, (cudaStream_t)stream, nullptr, ctx);
}
