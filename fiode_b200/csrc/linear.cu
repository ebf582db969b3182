// fiode_b200/csrc/linear.cu -- fp32 "NT" GEMM  C[m][n] = sum_k A[m][k] * B[n][k]  (both K-major)
// with deterministic split-K, + bias, + GroupSort epilogue.
//
// Replaces F.linear(X, Q, b) at libs/ortho_conv/layers.py:105 (CayleyLinear.forward) for the
// backbone head 4096->512->512->10, and serves the Cayley weight preparation (cayley.cu).
// FP32 FFMA on purpose: SURVEY section 8(d) asks tensor cores only where 3xTF32 holds rtol 1e-4 -- the
// tcgen05 3xTF32 variant is the planned replacement for the M>=256 head GEMM.
//
// CTA tile 64x64, 128 threads, thread tile 8(m, strided by 8) x 4(n, strided by 16); smem tiles keep
// K contiguous ([row][32+4]) so global->smem is a straight 16-byte cp.async and both operands
// are read back as float4 along K (12 LDS.128 per 128 FFMA).
#include "common.cuh"
#include "gemm_tc.cuh"
#include <cuda_pipeline.h>
#include <cstdlib>

namespace fiode {

constexpr int GT = 64, GK = 32, GS = GK + 4, GTHREADS = 128;

// DIRECT (one split, no GroupSort): the tile goes straight to y with alpha and bias applied -- no partials, no second launch
template <bool DIRECT>
__global__ void __launch_bounds__(GTHREADS) gemm_nt_kernel(const float* __restrict__ A, int lda,
                                                           const float* __restrict__ Bm, int ldb, int M, int N, int K,
                                                           int k_per_split, int vec_ok, float* __restrict__ part /*[splits][M][N]*/,
                                                           const float* __restrict__ bias = nullptr, float alpha = 1.f, int ldy = 0) {
    __shared__ __align__(16) float As[2][GT * GS];
    __shared__ __align__(16) float Bs[2][GT * GS];
    const int t = threadIdx.x;
    const int m0 = blockIdx.y * GT, n0 = blockIdx.x * GT;
    // bit 1 of vec_ok: the product is a Gram matrix (A == B, M == N): only the tiles on and above the diagonal are computed,
    // the rest is their mirror image -- exact, because sum_k a_ik a_jk runs over the same products in the same order
    const bool sym = (vec_ok & 2) != 0;
    vec_ok &= 1;
    if (sym && n0 < m0) return;
    const int k_begin = blockIdx.z * k_per_split, k_end = min(K, k_begin + k_per_split);
    const int tn = t & 15, tm = t >> 4;
    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    auto load_tile = [&](int buf, int k0) {
        // 64 rows x 8 float4 per operand = 512 float4 -> 4 per thread per operand
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int e = t + i * GTHREADS, r = e >> 3, q = e & 7;
            const int k = k0 + q * 4;
            float* da = &As[buf][r * GS + q * 4];
            float* db = &Bs[buf][r * GS + q * 4];
            if (vec_ok && m0 + r < M && k + 3 < k_end) __pipeline_memcpy_async(da, A + (size_t)(m0 + r) * lda + k, 16);
            else {
#pragma unroll
                for (int c = 0; c < 4; ++c) da[c] = (m0 + r < M && k + c < k_end) ? A[(size_t)(m0 + r) * lda + k + c] : 0.f;
            }
            if (vec_ok && n0 + r < N && k + 3 < k_end) __pipeline_memcpy_async(db, Bm + (size_t)(n0 + r) * ldb + k, 16);
            else {
#pragma unroll
                for (int c = 0; c < 4; ++c) db[c] = (n0 + r < N && k + c < k_end) ? Bm[(size_t)(n0 + r) * ldb + k + c] : 0.f;
            }
        }
        __pipeline_commit();
    };

    const int n_tiles = (k_end - k_begin + GK - 1) / GK;
    if (n_tiles > 0) load_tile(0, k_begin);
    for (int it = 0; it < n_tiles; ++it) {
        const int buf = it & 1;
        if (it + 1 < n_tiles) { load_tile(buf ^ 1, k_begin + (it + 1) * GK); __pipeline_wait_prior(1); }
        else __pipeline_wait_prior(0);
        __syncthreads();
#pragma unroll
        for (int q = 0; q < GK / 4; ++q) {
            float4 a[8], b[4];
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = *reinterpret_cast<const float4*>(&As[buf][(tm + 8 * i) * GS + q * 4]);
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = *reinterpret_cast<const float4*>(&Bs[buf][(tn + 16 * j) * GS + q * 4]);
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    acc[i][j] = fmaf(a[i].x, b[j].x, acc[i][j]);
                    acc[i][j] = fmaf(a[i].y, b[j].y, acc[i][j]);
                    acc[i][j] = fmaf(a[i].z, b[j].z, acc[i][j]);
                    acc[i][j] = fmaf(a[i].w, b[j].w, acc[i][j]);
                }
        }
        __syncthreads();
    }
    float* dst = part + (size_t)blockIdx.z * M * N;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + tm + 8 * i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tn + 16 * j;
            if (n < N) {
                if (DIRECT) part[(size_t)m * ldy + n] = alpha * acc[i][j] + (bias ? bias[n] : 0.f);
                else {
                    dst[(size_t)m * N + n] = acc[i][j];
                    if (sym && n0 > m0) dst[(size_t)n * N + m] = acc[i][j];
                }
            }
        }
    }
}

// y[m][n] = alpha * sum_s part[s][m][n] (+ bias[n]) (+ beta * addend[m][n]), optional GroupSort over n halves
__global__ void gemm_epilogue_kernel(const float* __restrict__ part, int splits, int M, int N,
                                     const float* __restrict__ bias, float alpha, int group_sort,
                                     float* __restrict__ y, int ldy) {
    const size_t MN = (size_t)M * N;
    if (group_sort) {
        const int half = N / 2;
        const size_t total = (size_t)M * half;
        for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
            const size_t m = e / half; const int n = (int)(e - m * half);
            float u = 0.f, w = 0.f;
            int s = 0;
            for (; s + 4 <= splits; s += 4) {      // loads in flight together, added in split order
                float a[4], b[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) { a[i] = part[(s + i) * MN + m * N + n]; b[i] = part[(s + i) * MN + m * N + n + half]; }
#pragma unroll
                for (int i = 0; i < 4; ++i) { u += a[i]; w += b[i]; }
            }
            for (; s < splits; ++s) { u += part[s * MN + m * N + n]; w += part[s * MN + m * N + n + half]; }
            u = alpha * u + (bias ? bias[n] : 0.f);
            w = alpha * w + (bias ? bias[n + half] : 0.f);
            y[m * ldy + n] = fmaxf(u, w);
            y[m * ldy + n + half] = fminf(u, w);
        }
    } else {
        for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < MN; e += (size_t)gridDim.x * blockDim.x) {
            const size_t m = e / N; const int n = (int)(e - m * N);
            float u = 0.f;
            int s = 0;
            for (; s + 8 <= splits; s += 8) {
                float a[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) a[i] = part[(s + i) * MN + e];
#pragma unroll
                for (int i = 0; i < 8; ++i) u += a[i];
            }
            for (; s < splits; ++s) u += part[s * MN + e];
            y[m * ldy + n] = alpha * u + (bias ? bias[n] : 0.f);
        }
    }
}

static int gemm_splits(int M, int N, int K) {
    const int tiles = ((M + GT - 1) / GT) * ((N + GT - 1) / GT);
    int splits = 1;
    const int target = 4 * sm_count();
    while (splits < 32 && tiles * splits < target && K / (splits * 2) >= 64) splits *= 2;
    return splits;
}
size_t gemm_nt_workspace(int M, int N, int K) {
    return (size_t)gemm_splits(M, N, K) * M * N * sizeof(float) + 256;
}

// C = alpha * A B^T (+bias) (+GroupSort), deterministic split-K sized to fill the GPU
static int gemm_nt_impl(const float* A, int lda, const float* Bm, int ldb, int M, int N, int K, const float* bias, float alpha,
                        int group_sort, float* y, int ldy, float* ws, size_t ws_bytes, cudaStream_t st, int splits);
int gemm_nt(const float* A, int lda, const float* Bm, int ldb, int M, int N, int K, const float* bias, float alpha,
            int group_sort, float* y, int ldy, float* ws, size_t ws_bytes, cudaStream_t st) {
    return gemm_nt_impl(A, lda, Bm, ldb, M, N, K, bias, alpha, group_sort, y, ldy, ws, ws_bytes, st, gemm_splits(M, N, K));
}
// The same product for the adjoints' tall operands (3584 x 512 x 512: 448 tiles of 64 x 64, three per SM): one pass over K that
// writes y directly instead of two K halves plus the reduction launch that re-reads both (18 of 90 us).  Not used on the forward
// path: summing K in one chain instead of two halves changes the last bits of Q.
int gemm_nt_tall(const float* A, int lda, const float* Bm, int ldb, int M, int N, int K, float alpha, float* y, int ldy,
                 float* ws, size_t ws_bytes, cudaStream_t st) {
    const int tiles = ((M + GT - 1) / GT) * ((N + GT - 1) / GT);
    const int splits = tiles >= 2 * sm_count() ? 1 : gemm_splits(M, N, K);
    return gemm_nt_impl(A, lda, Bm, ldb, M, N, K, nullptr, alpha, 0, y, ldy, ws, ws_bytes, st, splits);
}
// G = A A^T (n x n, the V^T V of the Cayley transform): same split-K partials and reduction as gemm_nt(A, A), bit for bit, with the
// tiles below the diagonal mirrored instead of computed
int gemm_nt_gram(const float* A, int lda, int n, int K, float* y, int ldy, float* ws, size_t ws_bytes, cudaStream_t st) {
    static const bool off = [] { const char* e = getenv("FIODE_GRAM"); return e && e[0] == '0'; }();   // A/B: the plain product
    const int splits = gemm_splits(n, n, K);
    return gemm_nt_impl(A, lda, A, lda, n, n, K, nullptr, 1.f, 0, y, ldy, ws, ws_bytes, st, (splits == 1 || off) ? splits : -splits);
}
static int gemm_nt_impl(const float* A, int lda, const float* Bm, int ldb, int M, int N, int K, const float* bias, float alpha,
                        int group_sort, float* y, int ldy, float* ws, size_t ws_bytes, cudaStream_t st, int splits) {
    const int sym = splits < 0 ? 2 : 0;           // negative split count: Gram product (see gemm_nt_gram)
    if (splits < 0) splits = -splits;
    if ((size_t)splits * M * N * sizeof(float) > ws_bytes) { set_error("gemm_nt: workspace too small"); return -1; }
    int kps = (K + splits - 1) / splits;
    kps = (kps + GK - 1) / GK * GK;
    splits = (K + kps - 1) / kps;
    dim3 grid((N + GT - 1) / GT, (M + GT - 1) / GT, splits);
    // 16-byte cp.async needs 16-byte aligned rows on both operands; otherwise scalar loads
    const int vec_ok = ((lda & 3) == 0) && ((ldb & 3) == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0) &&
                       ((reinterpret_cast<uintptr_t>(Bm) & 15) == 0);
    if (splits == 1 && !group_sort) {
        gemm_nt_kernel<true><<<grid, GTHREADS, 0, st>>>(A, lda, Bm, ldb, M, N, K, kps, vec_ok, y, bias, alpha, ldy);
        cudaError_t e1 = cudaGetLastError();
        if (e1 != cudaSuccess) return cuda_fail(e1, "gemm_nt launch");
        count_launch(1);
        return 0;
    }
    gemm_nt_kernel<false><<<grid, GTHREADS, 0, st>>>(A, lda, Bm, ldb, M, N, K, kps, vec_ok | sym, ws);
    const size_t total = group_sort ? (size_t)M * (N / 2) : (size_t)M * N;
    gemm_epilogue_kernel<<<(unsigned)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096), 256, 0, st>>>(
        ws, splits, M, N, bias, alpha, group_sort, y, ldy);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "gemm_nt launch");
    count_launch(2);
    return 0;
}


// ---- "TN" GEMM: C[m][n] = sum_k A[k][m] * B[k][n]  (reductions over rows: weight gradients) -----
// split-K over gridDim.z with deterministic partials, same epilogue as gemm_nt.
__global__ void __launch_bounds__(256) gemm_tn_kernel(const float* __restrict__ A, int lda, const float* __restrict__ Bm,
                                                      int ldb, int M, int N, long long K, long long k_per_split,
                                                      float* __restrict__ part) {
    __shared__ float As[16][64 + 4];
    __shared__ float Bs[16][64 + 4];
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    const long long k_begin = blockIdx.z * k_per_split, k_end = min(K, k_begin + k_per_split);
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (long long k0 = k_begin; k0 < k_end; k0 += 16) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int e = t + i * 256, kk = e >> 6, c = e & 63;
            const long long k = k0 + kk;
            As[kk][c] = (k < k_end && m0 + c < M) ? A[k * lda + m0 + c] : 0.f;
            Bs[kk][c] = (k < k_end && n0 + c < N) ? Bm[k * ldb + n0 + c] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
    float* dst = part + (size_t)blockIdx.z * M * N;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int m = m0 + ty * 4 + i, n = n0 + tx * 4 + j;
            if (m < M && n < N) dst[(size_t)m * N + n] = acc[i][j];
        }
}

static int gemm_tn_splits(int M, int N, long long K) {
    const int tiles = ((M + 63) / 64) * ((N + 63) / 64);
    long long splits = (2ll * sm_count() + tiles - 1) / tiles;
    const long long max_by_k = (K + 255) / 256;
    if (splits > max_by_k) splits = max_by_k;
    if (splits > 256) splits = 256;
    if (splits < 1) splits = 1;
    return (int)splits;
}
size_t gemm_tn_workspace(int M, int N, long long K) { return (size_t)gemm_tn_splits(M, N, K) * M * N * sizeof(float) + 256; }

int gemm_tn(const float* A, int lda, const float* Bm, int ldb, int M, int N, long long K, float alpha, float* y, int ldy,
            float* ws, size_t ws_bytes, cudaStream_t st) {
    int splits = gemm_tn_splits(M, N, K);
    if ((size_t)splits * M * N * sizeof(float) > ws_bytes) { set_error("gemm_tn: workspace too small"); return -1; }
    long long kps = (K + splits - 1) / splits;
    kps = (kps + 15) / 16 * 16;
    splits = (int)((K + kps - 1) / kps);
    dim3 grid((N + 63) / 64, (M + 63) / 64, splits);
    gemm_tn_kernel<<<grid, 256, 0, st>>>(A, lda, Bm, ldb, M, N, K, kps, ws);
    const size_t total = (size_t)M * N;
    gemm_epilogue_kernel<<<(unsigned)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096), 256, 0, st>>>(
        ws, splits, M, N, nullptr, alpha, 0, y, ldy);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "gemm_tn launch");
    count_launch(2);
    return 0;
}

}  // namespace fiode

using namespace fiode;

extern "C" size_t fiode_linear_workspace_bytes(int M, int N, int K) {
    const size_t a = gemm_nt_workspace(M, N, K), b = gemm_tc_packed_bytes(N, K) + 256;
    return a > b ? a : b;
}

extern "C" int fiode_linear_fwd(const float* X, const float* Q, const float* bias, int M, int N, int K,
                                int group_sort, float* y, void* ws, size_t ws_bytes, void* stream) {
    if (!X || !Q || !y || !ws) { set_error("fiode_linear_fwd: null pointer"); return -1; }
    if (M <= 0 || N <= 0 || K <= 0) { set_error("fiode_linear_fwd: bad sizes"); return -1; }
    if (group_sort && (N & 1)) { set_error("fiode_linear_fwd: GroupSort needs an even width"); return -1; }
    cudaStream_t st = (cudaStream_t)stream;
    // the tensor-core kernel tiles M by 128 rows per CTA and walks all of K in each: at a training batch (M = 128) it is 4 CTAs
    // deep in a latency chain plus a weight re-pack per call, where the split-K FFMA product fills the GPU -- rows >= 1024 only
    if (mlp_path() == 0 && M >= 1024 && gemm_tc_supported(N, K, group_sort) && ws_bytes >= gemm_tc_packed_bytes(N, K)) {
        // tensor cores (fp16x2 split): split + tile the weights into the workspace, then one persistent GEMM
        int rc = launch_pack_w(Q, N, K, gemm_tc_tile(N, group_sort), group_sort, ws, st);
        if (rc) return rc;
        return launch_linear_tc(X, K, ws, bias, M, N, K, group_sort, y, N, st);
    }
    return gemm_nt(X, K, Q, K, M, N, K, bias, 1.f, group_sort, y, N, reinterpret_cast<float*>(ws), ws_bytes, st);
}

// ---- fiode_linear_fwd without a caller-supplied workspace -------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot allocate (code under
// CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The scratch of fiode_linear_fwd -- the split-K
// partial sums of the FFMA product, or the fp16-split re-tiled weights of the tensor-core product -- lives in a file-scope device
// array: every GPU that loads the library holds kLinearStaticBytes (16 MiB) until the process exits.  Nothing in it has to start
// at zero (both products write every scratch element they later read), so it is not cleared; it keeps the last call's contents,
// and because every call on a device shares it, two calls that overlap on the GPU (different streams, or graph branches) race.
constexpr size_t kLinearStaticBytes = (size_t)16 << 20;
static __device__ __align__(256) unsigned char g_linear_static_scratch[kLinearStaticBytes];

// In host code the array's name is a host-side stand-in, not a device address: resolved once per device, with that device current.
static void* linear_static_scratch_ptr() {
    static std::atomic<void*> ptrs[kMaxDevices];
    int dev = -1;
    const cudaError_t eg = cudaGetDevice(&dev);
    if (eg != cudaSuccess) { set_error("fiode_linear_fwd_static: cudaGetDevice failed: %s", cudaGetErrorString(eg)); return nullptr; }
    if (dev < 0 || dev >= kMaxDevices) { set_error("fiode_linear_fwd_static: device index %d out of range", dev); return nullptr; }
    void* p = ptrs[dev].load(std::memory_order_acquire);
    if (p) return p;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_linear_static_scratch);
    if (e != cudaSuccess || !p) {
        set_error("fiode_linear_fwd_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
        return nullptr;
    }
    ptrs[dev].store(p, std::memory_order_release);      // every thread resolves the same address: a repeated store is harmless
    return p;
}

extern "C" size_t fiode_linear_static_max_bytes(void) { return kLinearStaticBytes; }

extern "C" int fiode_linear_fwd_static(const float* X, const float* Q, const float* bias, int M, int N, int K,
                                       int group_sort, float* y, void* stream) {
    if (!X || !Q || !y) { set_error("fiode_linear_fwd_static: null pointer"); return -1; }
    if (M <= 0 || N <= 0 || K <= 0) { set_error("fiode_linear_fwd_static: bad sizes"); return -1; }
    if (group_sort && (N & 1)) { set_error("fiode_linear_fwd_static: GroupSort needs an even width"); return -1; }
    // same routing as fiode_linear_fwd; the scratch this call needs is that of the product it takes
    const bool tcp = mlp_path() == 0 && M >= 1024 && gemm_tc_supported(N, K, group_sort);
    const size_t need = tcp ? gemm_tc_packed_bytes(N, K) : gemm_nt_workspace(M, N, K);
    if (need > kLinearStaticBytes) {
        set_error("fiode_linear_fwd_static: M=%d N=%d K=%d needs %zu bytes of scratch, the built-in array holds %zu; "
                  "use fiode_linear_fwd with a caller workspace", M, N, K, need, kLinearStaticBytes);
        return -1;
    }
    void* ws = linear_static_scratch_ptr();
    if (!ws) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    if (tcp) {
        int rc = launch_pack_w(Q, N, K, gemm_tc_tile(N, group_sort), group_sort, ws, st);
        if (rc) return rc;
        return launch_linear_tc(X, K, ws, bias, M, N, K, group_sort, y, N, st);
    }
    return gemm_nt(X, K, Q, K, M, N, K, bias, 1.f, group_sort, y, N, reinterpret_cast<float*>(ws), kLinearStaticBytes, st);
}
