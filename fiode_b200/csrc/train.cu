// fiode_b200/csrc/train.cu -- backward of the backbone layers for the FI training step.
//
// The reference trains through autograd (pl_modules.py:504-506 `loss.backward()`): the FI loss reaches the
// backbone through the static feature s = param_map(x) (pl_modules.py:395-400) and each layer's orthogonal
// matrix through `cayley` (libs/ortho_conv/layers.py:50-60).  These kernels are the hand-written adjoints of
//   * CayleyLinear.forward   layers.py:103-105   y = X Q^T + b
//   * CayleyConv.forward     layers.py:72-86     y = irfft2(Q_phi rfft2(x)) + b           (k = 3)
//   * the emulated stride-2 conv, layers.py:24-48 (space-to-depth + 1x1 mixing)
//   * GroupSort.forward      layers.py:127-131   (max, min) of the two channel halves
//   * Normalize.forward      models.py:23-26     (forward only: the input carries no gradient)
// The adjoint of cayley itself is in cayley.cu (fiode_cayley_linear_bwd / fiode_cayley_conv_bwd).
// Sums over the batch are ordered (no float atomics): gradients are bit-reproducible run to run.
#include "common.cuh"
#include "fft.cuh"

namespace fiode {

int gemm_nt(const float* A, int lda, const float* Bm, int ldb, int M, int N, int K, const float* bias, float alpha,
            int group_sort, float* y, int ldy, float* ws, size_t ws_bytes, cudaStream_t st);
size_t gemm_nt_workspace(int M, int N, int K);
int gemm_tn(const float* A, int lda, const float* Bm, int ldb, int M, int N, long long K, float alpha, float* y, int ldy,
            float* ws, size_t ws_bytes, cudaStream_t st);
size_t gemm_tn_workspace(int M, int N, long long K);
int launch_rfft2(const float* x, int n_img, int C, int n, const float* mu, const float* sd, float2* Xf, cudaStream_t st,
                 bool freq_major, bool x_u8 = false, int xrow = 0);
int launch_irfft2(const float2* Yf, int B, int C, int n, const float* bias, int gs, float* y, cudaStream_t st,
                  int freq_major);

static size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
static unsigned nblocks(size_t total, unsigned cap = 65535) {
    const size_t b = (total + 255) / 256;
    return (unsigned)(b < cap ? (b ? b : 1) : cap);
}

// ---- Normalize (models.py:23-26) ------------------------------------------------------------------
__global__ void normalize_kernel(const float* __restrict__ x, size_t total, int C, int inner, const float* __restrict__ mu,
                                 const float* __restrict__ sd, float* __restrict__ y) {
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)((e / inner) % C);
        y[e] = (x[e] - mu[c]) / sd[c];
    }
}

// ---- GroupSort adjoint: route (g_max, g_min) back to whichever half held the max / min; a tie splits evenly
// (torch.max / torch.min of two tensors, derivatives.yaml `maximum`) -----------------------------------------
__global__ void group_sort_bwd_kernel(const float* __restrict__ x, const float* __restrict__ gy, int B, int C, int inner,
                                      float* __restrict__ gx) {
    const size_t half = (size_t)(C / 2) * inner, total = (size_t)B * half;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t b = e / half, r = e - b * half;
        const size_t ia = b * 2 * half + r, ib = ia + half;
        const float u = x[ia], w = x[ib], gmax = gy[ia], gmin = gy[ib];
        float ga, gb;
        if (u > w) { ga = gmax; gb = gmin; }
        else if (u < w) { ga = gmin; gb = gmax; }
        else { ga = gb = 0.5f * (gmax + gmin); }
        gx[ia] = ga;
        gx[ib] = gb;
    }
}

// ---- bias gradient: gb[c] = sum_{b, r} gy[b][c][r], one CTA per channel, ordered --------------------------
// 32 warps, a warp per image with the lanes along the contiguous inner axis (float4 when it divides): the channel count is
// far below the SM count (32 / 64 CTAs), so each CTA has to keep many loads in flight on its own
constexpr int kBiasThreads = 1024;
__global__ void __launch_bounds__(kBiasThreads) bias_grad_kernel(const float* __restrict__ gy, int B, int C, int inner,
                                                                 float* __restrict__ gb) {
    __shared__ double sh[kBiasThreads / 32];
    const int c = blockIdx.x, w = threadIdx.x >> 5, l = threadIdx.x & 31;
    double acc = 0;
    if (inner >= 32) {
        for (int b = w; b < B; b += kBiasThreads / 32) {
            const float* g = gy + ((size_t)b * C + c) * inner;
            float a = 0.f;
            if ((inner & 127) == 0) {
                const float4* g4 = reinterpret_cast<const float4*>(g);
#pragma unroll 4
                for (int r = l; r < inner / 4; r += 32) { const float4 v = g4[r]; a += (v.x + v.y) + (v.z + v.w); }
            } else {
                for (int r = l; r < inner; r += 32) a += g[r];
            }
            acc += (double)a;
        }
    } else {
        const size_t per = (size_t)B * inner;
        for (size_t e = threadIdx.x; e < per; e += kBiasThreads) {
            const size_t b = e / inner, r = e - b * inner;
            acc += (double)gy[(b * C + c) * inner + r];
        }
    }
    acc = warp_sum(acc);
    if (l == 0) sh[w] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int i = 0; i < kBiasThreads / 32; ++i) s += sh[i];
        gb[c] = (float)s;
    }
}

__global__ void transpose2d_kernel(const float* __restrict__ x, int rows, int cols, float* __restrict__ y) {
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

// ---- stride-2 conv as 1x1 on space-to-depth: adjoints ----------------------------------------------------
// xs[b][ci][w][h] = x[b][c][2w + k1][2h + k2], ci = (c * 2 + k1) * 2 + k2          (layers.py:45-48)
__device__ __forceinline__ size_t s2d_src(int b, int ci, int p, int cin, int n) {
    const int n2 = n / 2, c = ci >> 2, k1 = (ci >> 1) & 1, k2 = ci & 1, w = p / n2, h = p - w * n2;
    return (((size_t)b * cin + c) * n + 2 * w + k1) * n + 2 * h + k2;
}
// gx (un-space-to-depth'd) = Q^T gy
__global__ void conv1x1_bwd_data_kernel(const float* __restrict__ Q, const float* __restrict__ gy, int B, int cin, int cout,
                                        int n, float* __restrict__ gx) {
    // one thread = the 2 x 2 input pixels of one (image, channel, output position): one gy load and one float4 of Q per output
    // channel feed four sums, and the results leave as two float2 rows (the scalar version was bound by its two loads per FMA)
    const int n2 = n / 2, P = n2 * n2, cin4 = 4 * cin;
    const size_t total = (size_t)B * cin * P;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int p = (int)(e % P), c = (int)((e / P) % cin), b = (int)(e / ((size_t)P * cin));
        const float* g = gy + (size_t)b * cout * P + p;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        for (int co = 0; co < cout; ++co) {
            const float4 q = __ldg(reinterpret_cast<const float4*>(Q + (size_t)co * cin4 + 4 * c));
            const float gv = g[(size_t)co * P];
            a0 = fmaf(q.x, gv, a0); a1 = fmaf(q.y, gv, a1); a2 = fmaf(q.z, gv, a2); a3 = fmaf(q.w, gv, a3);
        }
        const int w = p / n2, h = p - w * n2;                       // ci = 4c + 2 k1 + k2 -> pixel (2w + k1, 2h + k2), as s2d_src
        float* o = gx + (((size_t)b * cin + c) * n + 2 * w) * n + 2 * h;
        *reinterpret_cast<float2*>(o) = make_float2(a0, a1);
        *reinterpret_cast<float2*>(o + n) = make_float2(a2, a3);
    }
}
// part[split][co][ci] = sum_{b in split} sum_p gy[b][co][p] xs[b][ci][p] ; CTA tile 32 co x 32 ci, thread 2 x 2
__global__ void __launch_bounds__(256) conv1x1_bwd_weight_kernel(const float* __restrict__ x, const float* __restrict__ gy,
                                                                 int B, int cin, int cout, int n, int b_per_split,
                                                                 float* __restrict__ part) {
    constexpr int PT = 64, PS = PT + 4;             // row stride 68 floats: 16-byte aligned rows, 8 consecutive rows cover all 32 banks
    __shared__ __align__(16) float Gs[32][PS];
    __shared__ __align__(16) float Xs[32][PS];
    const int P = (n / 2) * (n / 2), cin4 = 4 * cin;
    const int ci0 = blockIdx.x * 32, co0 = blockIdx.y * 32;
    const int b_begin = blockIdx.z * b_per_split, b_end = min(B, b_begin + b_per_split);
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    float acc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
    for (int b = b_begin; b < b_end; ++b)
        for (int p0 = 0; p0 < P; p0 += PT) {
            for (int e = t; e < 32 * PT; e += 256) {
                const int r = e / PT, q = e - r * PT, p = p0 + q;
                Gs[r][q] = (co0 + r < cout && p < P) ? gy[((size_t)b * cout + co0 + r) * P + p] : 0.f;
                Xs[r][q] = (ci0 + r < cin4 && p < P) ? x[s2d_src(b, ci0 + r, p, cin, n)] : 0.f;
            }
            __syncthreads();
            // four positions per step: 4 LDS.128 feed 16 FMAs (the scalar version was bound by its 4 loads per 4 FMAs)
#pragma unroll 4
            for (int q = 0; q < PT; q += 4) {
                const float4 g0 = *reinterpret_cast<const float4*>(&Gs[ty][q]), g1 = *reinterpret_cast<const float4*>(&Gs[ty + 16][q]);
                const float4 x0 = *reinterpret_cast<const float4*>(&Xs[tx][q]), x1 = *reinterpret_cast<const float4*>(&Xs[tx + 16][q]);
                acc[0][0] = fmaf(g0.x, x0.x, acc[0][0]); acc[0][0] = fmaf(g0.y, x0.y, acc[0][0]); acc[0][0] = fmaf(g0.z, x0.z, acc[0][0]); acc[0][0] = fmaf(g0.w, x0.w, acc[0][0]);
                acc[0][1] = fmaf(g0.x, x1.x, acc[0][1]); acc[0][1] = fmaf(g0.y, x1.y, acc[0][1]); acc[0][1] = fmaf(g0.z, x1.z, acc[0][1]); acc[0][1] = fmaf(g0.w, x1.w, acc[0][1]);
                acc[1][0] = fmaf(g1.x, x0.x, acc[1][0]); acc[1][0] = fmaf(g1.y, x0.y, acc[1][0]); acc[1][0] = fmaf(g1.z, x0.z, acc[1][0]); acc[1][0] = fmaf(g1.w, x0.w, acc[1][0]);
                acc[1][1] = fmaf(g1.x, x1.x, acc[1][1]); acc[1][1] = fmaf(g1.y, x1.y, acc[1][1]); acc[1][1] = fmaf(g1.z, x1.z, acc[1][1]); acc[1][1] = fmaf(g1.w, x1.w, acc[1][1]);
            }
            __syncthreads();
        }
    float* dst = part + (size_t)blockIdx.z * cout * cin4;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int co = co0 + ty + 16 * i, ci = ci0 + tx + 16 * j;
            if (co < cout && ci < cin4) dst[(size_t)co * cin4 + ci] = acc[i][j];
        }
}
// out[e] = sum_k part[k][e]: a CTA takes 32 consecutive outputs, its eight warps an eighth of the splits each (loads of a whole
// chunk in flight, added in split order), then the eight sub-sums are added in warp order -- deterministic, and the 128 splits
// of the stride-2 conv's weight gradient no longer sit in one thread's dependent chain
__global__ void __launch_bounds__(256) sum_partials_kernel(const float* __restrict__ part, int splits, size_t total,
                                                           float* __restrict__ out) {
    __shared__ float sh[8][32];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    const int chunk = (splits + 7) / 8, k0 = w * chunk, k1 = min(splits, k0 + chunk);
    for (size_t base = (size_t)blockIdx.x * 32; base < total; base += (size_t)gridDim.x * 32) {
        const size_t e = base + l;
        float s = 0.f;
        if (e < total) {
            int k = k0;
            for (; k + 8 <= k1; k += 8) {
                float v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = part[(size_t)(k + i) * total + e];
#pragma unroll
                for (int i = 0; i < 8; ++i) s += v[i];
            }
            for (; k < k1; ++k) s += part[(size_t)k * total + e];
        }
        sh[w][l] = s;
        __syncthreads();
        if (w == 0 && e < total) {
            float r = sh[0][l];
#pragma unroll
            for (int i = 1; i < 8; ++i) r += sh[i][l];
            out[e] = r;
        }
        __syncthreads();
    }
}

// ---- Fourier-domain conv adjoints ---------------------------------------------------------------------
// gQf[phi][co][ci] = (c_phi / n^2) sum_b GYf[b][co][phi] conj(Xf[b][ci][phi]),  c_phi = 1 on the self-conjugate columns
// v = 0, n/2 of the half spectrum and 2 elsewhere (adjoint of irfft2, which reads each interior bin twice)
__global__ void __launch_bounds__(256) conv_bwd_weight_kernel(const float2* __restrict__ GYf, const float2* __restrict__ Xf, int B,
                                                              int cin, int cout, int n, float2* __restrict__ gQf) {
    // one thread = 2 output x 2 input channels of one frequency: four loads feed four complex FMAs per image (the one-entry
    // version re-read every spectrum value cin resp. cout times through L2 and was bound by that); each sum still runs over
    // the images in order
    const int H = n / 2 + 1, F = n * H, cin2 = (cin + 1) / 2, cout2 = (cout + 1) / 2;
    const size_t total = (size_t)cout2 * cin2 * F;
    const float inv = 1.f / (float)(n * n);
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int phi = (int)(e % F), ci = 2 * (int)((e / F) % cin2), co = 2 * (int)(e / ((size_t)F * cin2));
        const bool i1 = ci + 1 < cin, o1 = co + 1 < cout;
        const float2* gp0 = GYf + (size_t)co * F + phi;
        const float2* gp1 = GYf + (size_t)(o1 ? co + 1 : co) * F + phi;
        const float2* xp0 = Xf + (size_t)ci * F + phi;
        const float2* xp1 = Xf + (size_t)(i1 ? ci + 1 : ci) * F + phi;
        const size_t gs = (size_t)cout * F, xs = (size_t)cin * F;
        float2 a00 = make_float2(0.f, 0.f), a01 = a00, a10 = a00, a11 = a00;
#pragma unroll 8
        for (int b = 0; b < B; ++b) {
            const float2 g0 = gp0[b * gs], g1 = gp1[b * gs], x0 = cconj(xp0[b * xs]), x1 = cconj(xp1[b * xs]);
            cfma(a00, g0, x0); cfma(a01, g0, x1); cfma(a10, g1, x0); cfma(a11, g1, x1);
        }
        const int v = phi % H;
        const float c = ((v == 0 || v == n / 2) ? 1.f : 2.f) * inv;
        float2* o = gQf + ((size_t)phi * cout + co) * cin + ci;
        o[0] = make_float2(a00.x * c, a00.y * c);
        if (i1) o[1] = make_float2(a01.x * c, a01.y * c);
        if (o1) {
            o[cin] = make_float2(a10.x * c, a10.y * c);
            if (i1) o[cin + 1] = make_float2(a11.x * c, a11.y * c);
        }
    }
}
// GXf[b][ci][phi] = sum_co conj(Qf[phi][co][ci]) GYf[b][co][phi]
// consecutive threads = consecutive ci (Qf is (phi, co, ci): coalesced; the GYf value is one broadcast per warp), four images
// per thread so that each Qf entry is loaded once for four sums
constexpr int kBwdDataImgs = 4;
__global__ void __launch_bounds__(256) conv_bwd_data_kernel(const float2* __restrict__ Qf, const float2* __restrict__ GYf, int B,
                                                            int cin, int cout, int F, float2* __restrict__ GXf) {
    const int Bg = (B + kBwdDataImgs - 1) / kBwdDataImgs;
    const size_t total = (size_t)Bg * cin * F;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int ci = (int)(e % cin), phi = (int)((e / cin) % F), b0 = kBwdDataImgs * (int)(e / ((size_t)F * cin));
        float2 acc[kBwdDataImgs];
        const float2* gp[kBwdDataImgs];
#pragma unroll
        for (int i = 0; i < kBwdDataImgs; ++i) {
            acc[i] = make_float2(0.f, 0.f);
            gp[i] = GYf + (size_t)(b0 + i < B ? b0 + i : b0) * cout * F + phi;
        }
        const float2* qp = Qf + (size_t)phi * cout * cin + ci;
#pragma unroll 4
        for (int co = 0; co < cout; ++co) {
            const float2 q = cconj(__ldg(qp + (size_t)co * cin));
#pragma unroll
            for (int i = 0; i < kBwdDataImgs; ++i) cfma(acc[i], q, gp[i][(size_t)co * F]);
        }
#pragma unroll
        for (int i = 0; i < kBwdDataImgs; ++i)
            if (b0 + i < B) GXf[((size_t)(b0 + i) * cin + ci) * F + phi] = acc[i];
    }
}

}  // namespace fiode

using namespace fiode;

extern "C" int fiode_normalize(const float* x, int B, int C, int inner, const float* mu, const float* std, float* y,
                               void* stream) {
    if (!x || !mu || !std || !y || B <= 0 || C <= 0 || inner <= 0) { set_error("fiode_normalize: bad args"); return -1; }
    const size_t total = (size_t)B * C * inner;
    normalize_kernel<<<nblocks(total), 256, 0, (cudaStream_t)stream>>>(x, total, C, inner, mu, std, y);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}

extern "C" int fiode_group_sort_bwd(const float* x_pre, const float* gy, int B, int C, int inner, float* gx, void* stream) {
    if (!x_pre || !gy || !gx || (C & 1)) { set_error("fiode_group_sort_bwd: bad args"); return -1; }
    group_sort_bwd_kernel<<<nblocks((size_t)B * (C / 2) * inner), 256, 0, (cudaStream_t)stream>>>(x_pre, gy, B, C, inner, gx);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}

extern "C" int fiode_bias_grad(const float* gy, int B, int C, int inner, float* gb, void* stream) {
    if (!gy || !gb || B <= 0 || C <= 0 || inner <= 0) { set_error("fiode_bias_grad: bad args"); return -1; }
    bias_grad_kernel<<<C, kBiasThreads, 0, (cudaStream_t)stream>>>(gy, B, C, inner, gb);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}

extern "C" size_t fiode_linear_bwd_workspace_bytes(int M, int N, int K) {
    const size_t a = gemm_nt_workspace(M, K, N), b = gemm_tn_workspace(N, K, M);
    return al256((size_t)N * K * sizeof(float)) + al256(a) + al256(b) + 256;      // the two products run side by side
}

// y = X Q^T + b  (X: M x K, Q: N x K)  ->  gX = gY Q (optional), gQ = gY^T X, gb = column sums of gY (optional)
extern "C" int fiode_linear_bwd(const float* X, const float* Q, const float* gY, int M, int N, int K, float* gX, float* gQ,
                                float* gb, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!X || !Q || !gY || !gQ || !ws) { set_error("fiode_linear_bwd: null pointer"); return -1; }
    if (ws_bytes < fiode_linear_bwd_workspace_bytes(M, N, K)) { set_error("fiode_linear_bwd: workspace too small"); return -1; }
    float* Qt = reinterpret_cast<float*>(ws);
    char* wp = reinterpret_cast<char*>(ws) + al256((size_t)N * K * sizeof(float));
    const size_t a_bytes = al256(gemm_nt_workspace(M, K, N));
    float* gws_x = reinterpret_cast<float*>(wp);
    float* gws_q = reinterpret_cast<float*>(wp + a_bytes);
    const size_t gws_q_bytes = ws_bytes - al256((size_t)N * K * sizeof(float)) - a_bytes;
    int rc;
    // the weight and bias gradients (leaves of the backward pass) on a branch beside the data gradient the next layer waits for
    cudaStream_t sd = gX ? fork_side(st) : st;
    if ((rc = gemm_tn(gY, N, X, K, N, K, M, 1.f, gQ, K, gws_q, gws_q_bytes, sd))) return rc;
    if (gb) {
        bias_grad_kernel<<<N, kBiasThreads, 0, sd>>>(gY, M, N, 1, gb);
        count_launch(1);
    }
    if (gX) {
        transpose2d_kernel<<<dim3((K + 31) / 32, (N + 31) / 32), dim3(32, 8), 0, st>>>(Q, N, K, Qt);   // Qt: K x N
        count_launch(1);
        if ((rc = gemm_nt(gY, N, Qt, N, M, K, N, nullptr, 1.f, 0, gX, K, gws_x, a_bytes, st))) return rc;
        if (sd != st && (rc = join_side(st))) return rc;
    }
    FIODE_CUDA(cudaGetLastError());
    return 0;
}

// one split per image up to 128: the weight gradient's grid is (4cin/32) x (cout/32) x splits CTAs, and at 16 splits the 32 x 128
// matrix of the first stride-2 conv was 64 CTAs doing eight images each (0.27 ms on an otherwise idle GPU)
static int conv1x1_splits(int B) { return B < 128 ? B : 128; }

extern "C" size_t fiode_conv1x1_s2d_bwd_workspace_bytes(int B, int cin, int cout, int n) {
    (void)n;
    return (size_t)conv1x1_splits(B) * cout * 4 * cin * sizeof(float) + 256;
}

// y = Q space_to_depth(x) + b  (x: B x cin x n x n, Q: cout x 4cin)  ->  gx (optional), gQ, gb (optional)
extern "C" int fiode_conv1x1_s2d_bwd(const float* x, const float* Q, const float* gy, int B, int cin, int cout, int n,
                                     float* gx, float* gQ, float* gb, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!x || !Q || !gy || !gQ || !ws || (n & 1)) { set_error("fiode_conv1x1_s2d_bwd: bad args"); return -1; }
    if (ws_bytes < fiode_conv1x1_s2d_bwd_workspace_bytes(B, cin, cout, n)) { set_error("fiode_conv1x1_s2d_bwd: workspace too small"); return -1; }
    const int P = (n / 2) * (n / 2), cin4 = 4 * cin;
    cudaStream_t sd = gx ? fork_side(st) : st;      // weight + bias gradients beside the data gradient
    const int splits = conv1x1_splits(B), bps = (B + splits - 1) / splits;
    float* part = reinterpret_cast<float*>(ws);
    conv1x1_bwd_weight_kernel<<<dim3((cin4 + 31) / 32, (cout + 31) / 32, splits), 256, 0, sd>>>(x, gy, B, cin, cout, n, bps, part);
    sum_partials_kernel<<<nblocks(((size_t)cout * cin4 + 31) / 32 * 256), 256, 0, sd>>>(part, splits, (size_t)cout * cin4, gQ);
    count_launch(2);
    if (gb) {
        bias_grad_kernel<<<cout, kBiasThreads, 0, sd>>>(gy, B, cout, P, gb);
        count_launch(1);
    }
    if (gx) {
        conv1x1_bwd_data_kernel<<<nblocks((size_t)B * cin * P), 256, 0, st>>>(Q, gy, B, cin, cout, n, gx);
        count_launch(1);
        int rc;
        if (sd != st && (rc = join_side(st))) return rc;
    }
    FIODE_CUDA(cudaGetLastError());
    return 0;
}

extern "C" size_t fiode_cayley_conv_layer_bwd_workspace_bytes(int B, int cin, int cout, int n) {
    const size_t F = (size_t)n * (n / 2 + 1);
    return (size_t)B * (2 * cin + cout) * F * sizeof(float2) + 1024;
}

// y = irfft2(Qf[phi] rfft2(x)) + b  (x: B x cin x n x n, Qf: F x cout x cin complex)  ->  gx (optional), gQf, gb (optional)
extern "C" int fiode_cayley_conv_layer_bwd(const float* x, const float* Qf, const float* gy, int B, int cin, int cout, int n,
                                           float* gx, float* gQf, float* gb, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!x || !Qf || !gy || !gQf || !ws) { set_error("fiode_cayley_conv_layer_bwd: null pointer"); return -1; }
    if (ws_bytes < fiode_cayley_conv_layer_bwd_workspace_bytes(B, cin, cout, n)) { set_error("fiode_cayley_conv_layer_bwd: workspace too small"); return -1; }
    const int F = n * (n / 2 + 1);
    float2* Xf = reinterpret_cast<float2*>(ws);
    float2* GYf = Xf + (size_t)B * cin * F;
    float2* GXf = GYf + (size_t)B * cout * F;
    int rc;
    // branch: the input's spectrum, then the weight and bias gradients; the caller's stream: the incoming gradient's spectrum,
    // then the data gradient the previous layer waits for
    cudaStream_t sd = fork_side(st);
    if ((rc = launch_rfft2(x, B * cin, cin, n, nullptr, nullptr, Xf, sd, false))) return rc;
    if (gb) {
        bias_grad_kernel<<<cout, kBiasThreads, 0, sd>>>(gy, B, cout, n * n, gb);
        count_launch(1);
    }
    if ((rc = launch_rfft2(gy, B * cout, cout, n, nullptr, nullptr, GYf, st, false))) return rc;
    if (sd != st) sd = fork_side(st);                 // the branch waits for GYf
    conv_bwd_weight_kernel<<<nblocks((size_t)((cout + 1) / 2) * ((cin + 1) / 2) * F), 256, 0, sd>>>(GYf, Xf, B, cin, cout, n,
                                                                                                 reinterpret_cast<float2*>(gQf));
    count_launch(1);
    if (gx) {
        conv_bwd_data_kernel<<<nblocks((size_t)((B + kBwdDataImgs - 1) / kBwdDataImgs) * cin * F), 256, 0, st>>>(
            reinterpret_cast<const float2*>(Qf), GYf, B, cin, cout, F, GXf);
        count_launch(1);
        // no bias, no GroupSort: the channel index is irrelevant to the inverse FFT, so a channel count the kernel does not
        // tile (the 3-channel input layer) is regrouped as (B * cin / CH) images of CH channels
        const int CH = n == 32 ? 8 : 16;
        int Bv = B, Cv = cin;
        if (cin % CH) {
            if (((long long)B * cin) % CH) { set_error("fiode_cayley_conv_layer_bwd: B * cin = %d * %d is not a multiple of %d", B, cin, CH); return -1; }
            Bv = B * cin / CH; Cv = CH;
        }
        if ((rc = launch_irfft2(GXf, Bv, Cv, n, nullptr, 0, gx, st, false))) return rc;
    }
    if (sd != st && (rc = join_side(st))) return rc;
    FIODE_CUDA(cudaGetLastError());
    return 0;
}
