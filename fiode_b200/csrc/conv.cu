// fiode_b200/csrc/conv.cu -- Cayley orthogonal convolutions, forward (eval mode).
//
// Replaces libs/ortho_conv/layers.py:72-86 (CayleyConv.forward: rfft2 -> per-frequency orthogonal
// channel mixing -> irfft2 -> +bias), :24-48 (StridedConv space-to-depth pre-hook) and :127-131
// (GroupSort) with the per-frequency matrices Qf prepared once per weight version (cayley.cu).
//
//   k=3 layers:  rfft2_kernel -> mix_kernel -> irfft2_kernel(+bias, +GroupSort)
//   k=1 stride-2 layers (frequency-independent real Q):
//                conv1x1_s2d_kernel = space-to-depth + real 1x1 GEMM + bias + GroupSort, no FFT
//
// Spectra are kept in torch's rfft2 layout [b][c][u][v], v in [0, n/2], phi = u*(n/2+1)+v.
#include "common.cuh"
#include <cuda_pipeline.h>
#include <cstdlib>
#include "fft.cuh"

namespace fiode {

// ------------------------------------------------------------------------------------------------
// forward real 2-D FFT, n x n, one thread per row pair then one thread per spectrum column
// ------------------------------------------------------------------------------------------------
template <int n> struct FftCfg {
    static constexpr int H = n / 2 + 1;          // half-spectrum width
    static constexpr int RS = H + 1;             // padded smem row stride (complex)
    static constexpr int CH = (n == 32) ? 8 : 16;  // channel-images per CTA
    // 160 threads: the CH * H column tasks (136 at n = 32, 144 at n = 16) are one round instead of 128 + a nearly empty second one
    static constexpr int THREADS = 160;
};

// FM ("frequency-major"): the spectrum is written as Xt[phi][b][c] instead of torch's [b][c][phi], so that the channel
// mixing of one frequency is a plain row-major (batch x 2C) GEMM operand for the tensor-core kernel (gemm_tc.cu).
// U8: x holds 8-bit pixels; the load applies torchvision's ToTensor (float(p) / 255, IEEE division) before Normalize.
template <int n, bool FM, bool U8 = false>
__global__ void __launch_bounds__(FftCfg<n>::THREADS) rfft2_kernel(const float* __restrict__ x, int n_img, int C,
                                                    const float* __restrict__ mu, const float* __restrict__ sd,
                                                    float2* __restrict__ Xf, int xrow) {   // xrow: row stride of the [b][c][u][v] output (>= n/2 + 1)
    using Cf = FftCfg<n>;
    constexpr int RSI = n * Cf::RS + 1;           // per-image stride, odd: lanes may run along the image index
    __shared__ float2 R[Cf::CH * RSI];
    const int img0 = blockIdx.x * Cf::CH;
    // step 1: rows.  task = (channel-image ci, row pair p)
    for (int task = threadIdx.x; task < Cf::CH * (n / 2); task += Cf::THREADS) {
        const int ci = task / (n / 2), p = task - ci * (n / 2);
        const int img = img0 + ci;
        if (img >= n_img) continue;
        float m = 0.f, inv = 1.f;
        if (mu) { const int c = img % C; m = mu[c]; inv = sd[c]; }
        const float4* src = reinterpret_cast<const float4*>(x + ((size_t)img * n + 2 * p) * n);
        const uchar4* src8 = reinterpret_cast<const uchar4*>(reinterpret_cast<const unsigned char*>(x) + ((size_t)img * n + 2 * p) * n);
        float2 z[n];
#pragma unroll
        for (int q = 0; q < n / 4; ++q) {
            float4 a, b;
            if (U8) {
                const uchar4 ua = src8[q], ub = src8[n / 4 + q];
                a = make_float4(__fdiv_rn((float)ua.x, 255.f), __fdiv_rn((float)ua.y, 255.f), __fdiv_rn((float)ua.z, 255.f), __fdiv_rn((float)ua.w, 255.f));
                b = make_float4(__fdiv_rn((float)ub.x, 255.f), __fdiv_rn((float)ub.y, 255.f), __fdiv_rn((float)ub.z, 255.f), __fdiv_rn((float)ub.w, 255.f));
            } else { a = src[q]; b = src[n / 4 + q]; }
            if (mu) {   // models.py:23-26 Normalize: (x - mu) / std
                z[4 * q + 0] = make_float2((a.x - m) / inv, (b.x - m) / inv);
                z[4 * q + 1] = make_float2((a.y - m) / inv, (b.y - m) / inv);
                z[4 * q + 2] = make_float2((a.z - m) / inv, (b.z - m) / inv);
                z[4 * q + 3] = make_float2((a.w - m) / inv, (b.w - m) / inv);
            } else {
                z[4 * q + 0] = make_float2(a.x, b.x); z[4 * q + 1] = make_float2(a.y, b.y);
                z[4 * q + 2] = make_float2(a.z, b.z); z[4 * q + 3] = make_float2(a.w, b.w);
            }
        }
        fft_dif<n, false>(z);
        float2* ra = R + ci * RSI + 2 * p * Cf::RS;
        float2* rb = ra + Cf::RS;
#pragma unroll
        for (int v = 0; v <= n / 2; ++v) {
            const float2 zk = z[bitrev<n>(v)];
            const float2 zn = cconj(z[bitrev<n>((n - v) % n)]);
            ra[v] = make_float2(0.5f * (zk.x + zn.x), 0.5f * (zk.y + zn.y));
            const float2 d = csub(zk, zn);                 // B = -i (zk - zn) / 2
            rb[v] = make_float2(0.5f * d.y, -0.5f * d.x);
        }
    }
    __syncthreads();
    // step 2: columns.  task = (ci, v)
    // (rows padded to xrow > H: the pad columns are written too, as zeros, so that whole sectors leave the SM)
    const int vcols = FM ? Cf::H : xrow;
    for (int task = threadIdx.x; task < Cf::CH * vcols; task += Cf::THREADS) {
        const int ci = FM ? task % Cf::CH : task / vcols, v = FM ? task / Cf::CH : task - ci * vcols;
        const int img = img0 + ci;
        if (img >= n_img) continue;
        float2 z[n];
        if (v < Cf::H) {
#pragma unroll
            for (int a = 0; a < n; ++a) z[a] = R[ci * RSI + a * Cf::RS + v];
            fft_dif<n, false>(z);
        } else {
#pragma unroll
            for (int a = 0; a < n; ++a) z[a] = make_float2(0.f, 0.f);
        }
        if (FM) {                                  // lanes = consecutive channels of one image: contiguous runs
            float2* dst = Xf + (size_t)v * n_img + img;        // [(u H + v)][b][c], b C + c == img
#pragma unroll
            for (int u = 0; u < n; ++u) dst[(size_t)u * Cf::H * n_img] = z[bitrev<n>(u)];
        } else {
            float2* dst = Xf + (size_t)img * n * xrow + v;
#pragma unroll
            for (int u = 0; u < n; ++u) dst[u * xrow] = z[bitrev<n>(u)];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// inverse: Yf [img][n][H] -> y [img][n][n] (+bias[c]), optional GroupSort over channel halves
// CTA handles CH channel-images: with GroupSort they are CH/2 (c, c + C/2) pairs of one batch item.
// ------------------------------------------------------------------------------------------------
// FM: 0 torch layout [b][c][phi]; 1 frequency-major [phi][b][c]; 2 image-major, channels innermost [b][phi][c] (what the
// batched mixing GEMM writes with ldo = F * C: the 144 pieces a CTA gathers then sit in ONE 74 KB block per image)
template <int n, int FM>
__global__ void __launch_bounds__(FftCfg<n>::THREADS) irfft2_kernel(const float2* __restrict__ Yf, int B, int C,
                                                     const float* __restrict__ bias, int group_sort,
                                                     float* __restrict__ y) {
    using Cf = FftCfg<n>;
    constexpr int CH = Cf::CH;
    constexpr int RSI = n * Cf::RS + 1;
    extern __shared__ float4 smem4[];
    float2* T = reinterpret_cast<float2*>(smem4);                 // [CH][n][RS] (+1 per image)
    float* out = reinterpret_cast<float*>(T + CH * RSI + 1);      // [CH][n][n+1]
    // channel-image list of this CTA
    const int groups_per_b = C / CH;              // C is a multiple of CH (32, 64)
    const int b = blockIdx.x / groups_per_b, g = blockIdx.x - b * groups_per_b;
    auto chan = [&](int ci) {                     // ci in [0, CH)
        if (group_sort) return (ci < CH / 2) ? g * (CH / 2) + ci : C / 2 + g * (CH / 2) + (ci - CH / 2);
        return g * CH + ci;
    };
    if (b >= B) return;
    // step 1: columns (inverse along u)
    for (int task = threadIdx.x; task < CH * Cf::H; task += Cf::THREADS) {
        const int ci = FM ? task % CH : task / Cf::H, v = FM ? task / CH : task - ci * Cf::H;
        float2 z[n];
        if (FM == 1) {                             // Yt[(u H + v)][b][co]: lanes read consecutive channels
            const float2* src = Yf + ((size_t)v * B + b) * C + chan(ci);
#pragma unroll
            for (int u = 0; u < n; ++u) z[u] = src[(size_t)u * Cf::H * B * C];
        } else if (FM == 2) {                      // Yi[b][(u H + v)][co]
            const float2* src = Yf + ((size_t)b * n * Cf::H + v) * C + chan(ci);
#pragma unroll
            for (int u = 0; u < n; ++u) z[u] = src[(size_t)u * Cf::H * C];
        } else {
            const float2* src = Yf + ((size_t)b * C + chan(ci)) * n * Cf::H + v;
#pragma unroll
            for (int u = 0; u < n; ++u) z[u] = src[u * Cf::H];
        }
        fft_dif<n, true>(z);
#pragma unroll
        for (int a = 0; a < n; ++a) T[ci * RSI + a * Cf::RS + v] = z[bitrev<n>(a)];
    }
    __syncthreads();
    // step 2: rows, two at a time (C2R: imag of DC / Nyquist ignored)
    const float scale = 1.0f / (float)(n * n);
    for (int task = threadIdx.x; task < CH * (n / 2); task += Cf::THREADS) {
        const int ci = task / (n / 2), p = task - ci * (n / 2);
        const float2* ta = T + ci * RSI + 2 * p * Cf::RS;
        const float2* tb = ta + Cf::RS;
        float2 z[n];
#pragma unroll
        for (int k = 0; k <= n / 2; ++k) {
            float2 A = ta[k], Bq = tb[k];
            if (k == 0 || k == n / 2) { A.y = 0.f; Bq.y = 0.f; }
            z[k] = make_float2(A.x - Bq.y, A.y + Bq.x);                     // A + iB
            if (k > 0 && k < n / 2) z[n - k] = make_float2(A.x + Bq.y, -A.y + Bq.x);  // conj(A) + i conj(B)
        }
        fft_dif<n, true>(z);
        const float bs = bias ? bias[chan(ci)] : 0.f;
        float* oa = out + (ci * n + 2 * p) * (n + 1);
        float* ob = oa + (n + 1);
#pragma unroll
        for (int k = 0; k < n; ++k) {
            const float2 r = z[bitrev<n>(k)];
            oa[k] = r.x * scale + bs;
            ob[k] = r.y * scale + bs;
        }
    }
    __syncthreads();
    if (group_sort) {
        for (int e = threadIdx.x; e < (CH / 2) * n * n; e += Cf::THREADS) {
            const int ci = e / (n * n), px = e - ci * (n * n), a = px / n, k = px - a * n;
            const float u = out[(ci * n + a) * (n + 1) + k], w = out[((ci + CH / 2) * n + a) * (n + 1) + k];
            y[((size_t)b * C + chan(ci)) * n * n + px] = fmaxf(u, w);
            y[((size_t)b * C + chan(ci + CH / 2)) * n * n + px] = fminf(u, w);
        }
    } else {
        for (int e = threadIdx.x; e < CH * n * n; e += Cf::THREADS) {
            const int ci = e / (n * n), px = e - ci * (n * n), a = px / n, k = px - a * n;
            y[((size_t)b * C + chan(ci)) * n * n + px] = out[(ci * n + a) * (n + 1) + k];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// model.0 fused: per-frequency mixing (3 -> C channels) + inverse 2-D FFT + bias + GroupSort in one kernel.
// The 3-channel input spectrum is tiny (13 KB / image), the mixed spectrum is not (139 KB / image): computing
// Yf[b][co][phi] = sum_c Q[phi][co][c] Xf[b][c][phi] on the fly in the column pass removes its HBM round trip.
// CTA = IM images x 8 output channels (4 GroupSort pairs); Qt is the spectrum re-laid as [co][c][phi] so that
// lanes (consecutive phi) read consecutive addresses.  Row-pass results overwrite the T rows they came from.
// ------------------------------------------------------------------------------------------------
__global__ void transpose_q_kernel(const float2* __restrict__ Qf, int F, int CC, float2* __restrict__ Qt) {
    const long long total = (long long)F * CC;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const long long cc = e / F, phi = e - cc * F;       // e indexes the OUTPUT: coalesced writes
        Qt[e] = Qf[phi * CC + cc];
    }
}

template <int IM>
__global__ void __launch_bounds__(160) irfft32_mix3_kernel(const float2* __restrict__ Xf, const float2* __restrict__ Qt,
                                                           int B, int C, const float* __restrict__ bias, int group_sort,
                                                           float* __restrict__ y) {
    // RS = H (no row padding): phi = u * RS + v, so the mixing writes consecutive words, the column pass reads consecutive
    // words across lanes, and the row pass (lane = row pair, stride 2 RS float2 = 68 words) is 2-way instead of 4-way conflicted
    constexpr int n = 32, CH = 8, H = 17, RS = 17, RSI = n * RS + 1, F = n * H, THREADS = 160, OS = n + 4;
    extern __shared__ float4 smem4[];
    float2* T = reinterpret_cast<float2*>(smem4);                 // [IM][CH] images of [n][RS] (+1)
    const int groups_per_b = C / CH;
    const int bg = blockIdx.x / groups_per_b, g = blockIdx.x - bg * groups_per_b;
    const int b0 = bg * IM;
    auto chan = [&](int ci) {
        if (group_sort) return (ci < CH / 2) ? g * (CH / 2) + ci : C / 2 + g * (CH / 2) + (ci - CH / 2);
        return g * CH + ci;
    };
    // the IM input spectra (3 channels each) are staged once; every thread then reuses each Q value for all IM images
    float2* Xs = T + IM * CH * RSI + ((IM * CH * RSI) & 1);       // [IM][3][F], 16-byte aligned
    // asynchronous 16-byte copies (LDGSTS): all in flight at once, and they overlap the first batch of Q loads below
    for (int e = threadIdx.x; e < IM * 3 * F / 2; e += THREADS) {
        const int im = e / (3 * F / 2);
        if (b0 + im < B) __pipeline_memcpy_async(Xs + 2 * e, Xf + (size_t)b0 * 3 * F + 2 * e, 16);
        else *reinterpret_cast<float4*>(Xs + 2 * e) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __pipeline_commit();
    // step 1a: mix.  Yf[im][ci][phi] = sum_c Q[chan(ci)][c][phi] Xs[im][c][phi].  Lanes run along phi (coalesced Q reads, each
    // Q value read once per CTA); a thread does channels cp and cp + CH/2 of one phi, so every Xs value it reads from shared
    // memory feeds two outputs; the Q loads of UB iterations are issued together (they would otherwise serialise on L2 latency).
    {
        constexpr int UB = 5, ITERS = ((CH / 2) * F + THREADS - 1) / THREADS;      // 14 (phi, channel-pair) items per thread
        for (int it0 = 0; it0 < ITERS; it0 += UB) {
            float2 qa[UB][3], qb[UB][3];
#pragma unroll
            for (int j = 0; j < UB; ++j) {
                const int e = threadIdx.x + (it0 + j) * THREADS;
                if (it0 + j < ITERS && e < (CH / 2) * F) {
                    const int cp = e / F, phi = e - cp * F;
                    const float2* q = Qt + (size_t)chan(cp) * 3 * F + phi;
                    const float2* r = Qt + (size_t)chan(cp + CH / 2) * 3 * F + phi;
                    qa[j][0] = __ldg(q); qa[j][1] = __ldg(q + F); qa[j][2] = __ldg(q + 2 * F);
                    qb[j][0] = __ldg(r); qb[j][1] = __ldg(r + F); qb[j][2] = __ldg(r + 2 * F);
                }
            }
            if (it0 == 0) {                          // uniform: the staged spectra are needed from here on
                __pipeline_wait_prior(0);
                __syncthreads();
            }
#pragma unroll
            for (int j = 0; j < UB; ++j) {
                const int e = threadIdx.x + (it0 + j) * THREADS;
                if (it0 + j < ITERS && e < (CH / 2) * F) {
                    const int cp = e / F, phi = e - cp * F;
#pragma unroll
                    for (int im = 0; im < IM; ++im) {
                        const float2* xs = Xs + im * 3 * F + phi;
                        const float2 x0 = xs[0], x1 = xs[F], x2 = xs[2 * F];
                        float2 acc = make_float2(0.f, 0.f), bcc = make_float2(0.f, 0.f);
                        cfma(acc, qa[j][0], x0); cfma(acc, qa[j][1], x1); cfma(acc, qa[j][2], x2);
                        cfma(bcc, qb[j][0], x0); cfma(bcc, qb[j][1], x1); cfma(bcc, qb[j][2], x2);
                        T[(im * CH + cp) * RSI + phi] = acc;
                        T[(im * CH + cp + CH / 2) * RSI + phi] = bcc;
                    }
                }
            }
        }
    }
    __syncthreads();
    // step 1b: columns (inverse along u), in place.  task = (image, channel, v), v fastest
    for (int task = threadIdx.x; task < IM * CH * H; task += THREADS) {
        const int ic = task / H, v = task - ic * H;
        float2* col = T + ic * RSI + v;
        float2 z[n];
#pragma unroll
        for (int u = 0; u < n; ++u) z[u] = col[u * RS];
        fft_dif<n, true>(z);
#pragma unroll
        for (int a = 0; a < n; ++a) col[a * RS] = z[bitrev<n>(a)];
    }
    __syncthreads();
    // step 2: rows.  task = (image, channel pair (ci, ci + CH/2), row pair p): two real rows per complex FFT (rows 2p, 2p+1 packed
    // as re / im), both channels of a GroupSort pair in the same thread, so bias + (max, min) happen in registers.  The four
    // finished rows are staged in shared memory (over T / Xs, which are dead by then) and leave as coalesced float4 rows.
    const float scale = 1.0f / (float)(n * n);
    const int task = threadIdx.x;
    const bool has_task = task < IM * (CH / 2) * (n / 2);
    const int ic2 = task / (n / 2), p = task - ic2 * (n / 2), im = ic2 / (CH / 2), ci = ic2 - im * (CH / 2);
    float2 za[n], zb[n];
    if (has_task) {
        auto load_rows = [&](int c, float2 (&z)[n]) {
            const float2* ta = T + (im * CH + c) * RSI + 2 * p * RS;
            const float2* tb = ta + RS;
#pragma unroll
            for (int k = 0; k <= n / 2; ++k) {
                float2 A = ta[k], Bq = tb[k];
                if (k == 0 || k == n / 2) { A.y = 0.f; Bq.y = 0.f; }
                z[k] = make_float2(A.x - Bq.y, A.y + Bq.x);
                if (k > 0 && k < n / 2) z[n - k] = make_float2(A.x + Bq.y, -A.y + Bq.x);
            }
            fft_dif<n, true>(z);
        };
        load_rows(ci, za);
        load_rows(ci + CH / 2, zb);
    }
    __syncthreads();                                   // every read of T is done: reuse the buffer as the output stage
    float* O = reinterpret_cast<float*>(smem4);        // [IM * CH][n rows][OS] floats (OS = n + 4: conflict-free float4 rows)
    if (has_task) {
        const float bsa = bias ? bias[chan(ci)] : 0.f, bsb = bias ? bias[chan(ci + CH / 2)] : 0.f;
        float* oa = O + ((im * CH + ci) * n + 2 * p) * OS;
        float* ob = O + ((im * CH + ci + CH / 2) * n + 2 * p) * OS;
#pragma unroll
        for (int k = 0; k < n; k += 4) {
            float a0[4], a1[4], c0[4], c1[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 ra = za[bitrev<n>(k + q)], rb = zb[bitrev<n>(k + q)];
                const float u0 = ra.x * scale + bsa, u1 = ra.y * scale + bsa;       // channel a: row 2p, row 2p+1
                const float w0 = rb.x * scale + bsb, w1 = rb.y * scale + bsb;       // channel b
                a0[q] = group_sort ? fmaxf(u0, w0) : u0; c0[q] = group_sort ? fminf(u0, w0) : w0;
                a1[q] = group_sort ? fmaxf(u1, w1) : u1; c1[q] = group_sort ? fminf(u1, w1) : w1;
            }
            *reinterpret_cast<float4*>(oa + k) = make_float4(a0[0], a0[1], a0[2], a0[3]);
            *reinterpret_cast<float4*>(oa + OS + k) = make_float4(a1[0], a1[1], a1[2], a1[3]);
            *reinterpret_cast<float4*>(ob + k) = make_float4(c0[0], c0[1], c0[2], c0[3]);
            *reinterpret_cast<float4*>(ob + OS + k) = make_float4(c1[0], c1[1], c1[2], c1[3]);
        }
    }
    __syncthreads();
    // step 3: coalesced copy-out, one float4 per lane, a channel-image (4 KB) per 256 consecutive items
    for (int e = threadIdx.x; e < IM * CH * n * (n / 4); e += THREADS) {
        const int ic = e / (n * (n / 4)), rem = e - ic * (n * (n / 4)), row = rem / (n / 4), k4 = rem - row * (n / 4);
        const int im3 = ic / CH, c3 = ic - im3 * CH;
        if (b0 + im3 >= B) continue;
        const float4 v4 = *reinterpret_cast<const float4*>(O + (ic * n + row) * OS + k4 * 4);
        *reinterpret_cast<float4*>(y + ((size_t)(b0 + im3) * C + chan(c3)) * n * n + rem * 4) = v4;
    }
}
// Wide variant: 288 threads, so that the 272 column tasks and the 256 row tasks (ONE channel-image row pair each) are a single
// round, 18 warps per SM instead of 10 (ncu r01f: 160-thread CTAs leave 2.5 warps per scheduler, 'wait' + 'selected' are the top
// stall reasons, issue 52 %).  One FFT per thread keeps it under 112 registers; GroupSort moves to the copy-out.
template <int IM>
__global__ void __launch_bounds__(288, 2) irfft32_mix3_wide_kernel(const float2* __restrict__ Xf, const float2* __restrict__ Qt,
                                                           int B, int C, const float* __restrict__ bias, int group_sort,
                                                           float* __restrict__ y) {
    // RS = H (no row padding): phi = u * RS + v, so the mixing writes consecutive words, the column pass reads consecutive
    // words across lanes, and the row pass (lane = row pair, stride 2 RS float2 = 68 words) is 2-way instead of 4-way conflicted
    constexpr int n = 32, CH = 8, H = 17, RS = 17, RSI = n * RS + 1, F = n * H, THREADS = 288, OS = n + 4;
    extern __shared__ float4 smem4[];
    float2* T = reinterpret_cast<float2*>(smem4);                 // [IM][CH] images of [n][RS] (+1)
    const int groups_per_b = C / CH;
    const int bg = blockIdx.x / groups_per_b, g = blockIdx.x - bg * groups_per_b;
    const int b0 = bg * IM;
    auto chan = [&](int ci) {
        if (group_sort) return (ci < CH / 2) ? g * (CH / 2) + ci : C / 2 + g * (CH / 2) + (ci - CH / 2);
        return g * CH + ci;
    };
    // the IM input spectra (3 channels each) are staged once; every thread then reuses each Q value for all IM images
    float2* Xs = T + IM * CH * RSI + ((IM * CH * RSI) & 1);       // [IM][3][F], 16-byte aligned
    // asynchronous 16-byte copies (LDGSTS): all in flight at once, and they overlap the first batch of Q loads below
    for (int e = threadIdx.x; e < IM * 3 * F / 2; e += THREADS) {
        const int im = e / (3 * F / 2);
        if (b0 + im < B) __pipeline_memcpy_async(Xs + 2 * e, Xf + (size_t)b0 * 3 * F + 2 * e, 16);
        else *reinterpret_cast<float4*>(Xs + 2 * e) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __pipeline_commit();
    // step 1a: mix.  Yf[im][ci][phi] = sum_c Q[chan(ci)][c][phi] Xs[im][c][phi].  Lanes run along phi (coalesced Q reads, each
    // Q value read once per CTA); a thread does channels cp and cp + CH/2 of one phi, so every Xs value it reads from shared
    // memory feeds two outputs; the Q loads of UB iterations are issued together (they would otherwise serialise on L2 latency).
    {
        constexpr int UB = 4, ITERS = ((CH / 2) * F + THREADS - 1) / THREADS;      // 8 (phi, channel-pair) items per thread
        for (int it0 = 0; it0 < ITERS; it0 += UB) {
            float2 qa[UB][3], qb[UB][3];
#pragma unroll
            for (int j = 0; j < UB; ++j) {
                const int e = threadIdx.x + (it0 + j) * THREADS;
                if (it0 + j < ITERS && e < (CH / 2) * F) {
                    const int cp = e / F, phi = e - cp * F;
                    const float2* q = Qt + (size_t)chan(cp) * 3 * F + phi;
                    const float2* r = Qt + (size_t)chan(cp + CH / 2) * 3 * F + phi;
                    qa[j][0] = __ldg(q); qa[j][1] = __ldg(q + F); qa[j][2] = __ldg(q + 2 * F);
                    qb[j][0] = __ldg(r); qb[j][1] = __ldg(r + F); qb[j][2] = __ldg(r + 2 * F);
                }
            }
            if (it0 == 0) {                          // uniform: the staged spectra are needed from here on
                __pipeline_wait_prior(0);
                __syncthreads();
            }
#pragma unroll
            for (int j = 0; j < UB; ++j) {
                const int e = threadIdx.x + (it0 + j) * THREADS;
                if (it0 + j < ITERS && e < (CH / 2) * F) {
                    const int cp = e / F, phi = e - cp * F;
#pragma unroll
                    for (int im = 0; im < IM; ++im) {
                        const float2* xs = Xs + im * 3 * F + phi;
                        const float2 x0 = xs[0], x1 = xs[F], x2 = xs[2 * F];
                        float2 acc = make_float2(0.f, 0.f), bcc = make_float2(0.f, 0.f);
                        cfma(acc, qa[j][0], x0); cfma(acc, qa[j][1], x1); cfma(acc, qa[j][2], x2);
                        cfma(bcc, qb[j][0], x0); cfma(bcc, qb[j][1], x1); cfma(bcc, qb[j][2], x2);
                        T[(im * CH + cp) * RSI + phi] = acc;
                        T[(im * CH + cp + CH / 2) * RSI + phi] = bcc;
                    }
                }
            }
        }
    }
    __syncthreads();
    // step 1b: columns (inverse along u), in place.  task = (image, channel, v), v fastest
    for (int task = threadIdx.x; task < IM * CH * H; task += THREADS) {
        const int ic = task / H, v = task - ic * H;
        float2* col = T + ic * RSI + v;
        float2 z[n];
#pragma unroll
        for (int u = 0; u < n; ++u) z[u] = col[u * RS];
        fft_dif<n, true>(z);
#pragma unroll
        for (int a = 0; a < n; ++a) col[a * RS] = z[bitrev<n>(a)];
    }
    __syncthreads();
    // step 2: rows.  task = (image, channel, row pair p): two real rows per complex FFT (rows 2p, 2p+1 packed as re / im).  The two
    // finished rows (+ bias) are staged in shared memory (over T / Xs, which are dead by then).
    const float scale = 1.0f / (float)(n * n);
    const int task = threadIdx.x;
    const bool has_task = task < IM * CH * (n / 2);
    const int ic2 = task / (n / 2), p = task - ic2 * (n / 2), ci = ic2 % CH;
    float2 za[n];
    if (has_task) {
        const float2* ta = T + ic2 * RSI + 2 * p * RS;
        const float2* tb = ta + RS;
#pragma unroll
        for (int k = 0; k <= n / 2; ++k) {
            float2 A = ta[k], Bq = tb[k];
            if (k == 0 || k == n / 2) { A.y = 0.f; Bq.y = 0.f; }
            za[k] = make_float2(A.x - Bq.y, A.y + Bq.x);
            if (k > 0 && k < n / 2) za[n - k] = make_float2(A.x + Bq.y, -A.y + Bq.x);
        }
        fft_dif<n, true>(za);
    }
    __syncthreads();                                   // every read of T is done: reuse the buffer as the output stage
    float* O = reinterpret_cast<float*>(smem4);        // [IM * CH][n rows][OS] floats (OS = n + 4: conflict-free float4 rows)
    if (has_task) {
        const float bsa = bias ? bias[chan(ci)] : 0.f;
        float* oa = O + (ic2 * n + 2 * p) * OS;
#pragma unroll
        for (int k = 0; k < n; k += 4) {
            float a0[4], a1[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 ra = za[bitrev<n>(k + q)];
                a0[q] = ra.x * scale + bsa; a1[q] = ra.y * scale + bsa;             // row 2p, row 2p+1
            }
            *reinterpret_cast<float4*>(oa + k) = make_float4(a0[0], a0[1], a0[2], a0[3]);
            *reinterpret_cast<float4*>(oa + OS + k) = make_float4(a1[0], a1[1], a1[2], a1[3]);
        }
    }
    __syncthreads();
    // step 3: GroupSort of the (ci, ci + CH/2) pair (layers.py:127-131) + coalesced copy-out, one float4 of each partner per lane
    for (int e = threadIdx.x; e < IM * (CH / 2) * n * (n / 4); e += THREADS) {
        const int ip = e / (n * (n / 4)), rem = e - ip * (n * (n / 4)), row = rem / (n / 4), k4 = rem - row * (n / 4);
        const int im3 = ip / (CH / 2), c3 = ip - im3 * (CH / 2);
        if (b0 + im3 >= B) continue;
        const float4 u = *reinterpret_cast<const float4*>(O + ((im3 * CH + c3) * n + row) * OS + k4 * 4);
        const float4 w = *reinterpret_cast<const float4*>(O + ((im3 * CH + c3 + CH / 2) * n + row) * OS + k4 * 4);
        const float4 hi = group_sort ? make_float4(fmaxf(u.x, w.x), fmaxf(u.y, w.y), fmaxf(u.z, w.z), fmaxf(u.w, w.w)) : u;
        const float4 lo = group_sort ? make_float4(fminf(u.x, w.x), fminf(u.y, w.y), fminf(u.z, w.z), fminf(u.w, w.w)) : w;
        *reinterpret_cast<float4*>(y + ((size_t)(b0 + im3) * C + chan(c3)) * n * n + rem * 4) = hi;
        *reinterpret_cast<float4*>(y + ((size_t)(b0 + im3) * C + chan(c3 + CH / 2)) * n * n + rem * 4) = lo;
    }
}
// FIODE_MIX3_T=160 selects the 160-thread kernel (two FFTs per row task), default the 288-thread one
static bool mix3_wide() {
    static int v = [] { const char* e = getenv("FIODE_MIX3_T"); return e ? atoi(e) : 288; }();
    return v != 160;
}
template <int IM>
static int launch_irfft32_mix3_t(const float2* Xf, const float2* Qt, int B, int C, const float* bias, int gs, float* y,
                                 cudaStream_t st) {
    constexpr int F = 544;
    const size_t stage = (size_t)IM * 8 * (32 * 17 + 1) * sizeof(float2) + 16 + (size_t)IM * 3 * F * sizeof(float2);
    const size_t outb = (size_t)IM * 8 * 32 * 36 * sizeof(float);      // output stage reuses the same buffer
    const size_t smem = stage > outb ? stage : outb;
    cudaError_t e = cudaFuncSetAttribute(irfft32_mix3_kernel<IM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "irfft32_mix3 smem attr");
    if (mix3_wide()) {
        e = cudaFuncSetAttribute(irfft32_mix3_wide_kernel<IM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return cuda_fail(e, "irfft32_mix3 smem attr");
        irfft32_mix3_wide_kernel<IM><<<((B + IM - 1) / IM) * (C / 8), 288, smem, st>>>(Xf, Qt, B, C, bias, gs, y);
    } else {
        irfft32_mix3_kernel<IM><<<((B + IM - 1) / IM) * (C / 8), 160, smem, st>>>(Xf, Qt, B, C, bias, gs, y);
    }
    return 0;
}
// images per CTA: 2 reuses every Q value twice (2 CTAs / SM), 1 halves the footprint (4 CTAs / SM); FIODE_MIX3_IM overrides
static int mix3_images_per_cta() {
    static int v = [] { const char* e = getenv("FIODE_MIX3_IM"); const int x = e ? atoi(e) : 2; return x == 1 ? 1 : 2; }();
    return v;
}
int launch_transpose_q(const float2* Qf, int F, int CC, float2* Qt, cudaStream_t st) {
    transpose_q_kernel<<<(F * CC + 255) / 256, 256, 0, st>>>(Qf, F, CC, Qt);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
// Qt: F * cout * 3 float2, filled here from Qf; Qf == nullptr means Qt already holds the transposed spectrum
int launch_irfft32_mix3(const float2* Xf, const float2* Qf, float2* Qt, int B, int C, const float* bias, int gs, float* y,
                        cudaStream_t st) {
    constexpr int F = 544;
    if (C % 8) { set_error("irfft32_mix3: C=%d", C); return -1; }
    if (Qf) transpose_q_kernel<<<(F * C * 3 + 255) / 256, 256, 0, st>>>(Qf, F, C * 3, Qt);
    const int rc = mix3_images_per_cta() == 1 ? launch_irfft32_mix3_t<1>(Xf, Qt, B, C, bias, gs, y, st)
                                              : launch_irfft32_mix3_t<2>(Xf, Qt, B, C, bias, gs, y, st);
    if (rc) return rc;
    count_launch(Qf ? 2 : 1);
    return 0;
}

// ------------------------------------------------------------------------------------------------
// per-frequency orthogonal channel mixing: Yf[b][co][phi] = sum_ci Qf[phi][co][ci] * Xf[b][ci][phi]
// CTA = (4 consecutive frequencies, IB images); thread = 4 co x 4 b x 4 phi complex accumulators.
// ------------------------------------------------------------------------------------------------
template <int COUT, int CIN>
__global__ void __launch_bounds__((COUT / 4) * 16) mix_kernel(const float2* __restrict__ Xf,
                                                              const float2* __restrict__ Qf, int B, int F,
                                                              float2* __restrict__ Yf) {
    constexpr int IB = 64, FB = 4, THREADS = (COUT / 4) * 16;
    extern __shared__ float4 smem4[];
    float2* Qs = reinterpret_cast<float2*>(smem4);          // [FB][CIN][COUT]
    float2* Xs = Qs + FB * CIN * COUT;                      // [FB][CIN][IB]
    const int t = threadIdx.x;
    const int f0 = blockIdx.x * FB, b0 = blockIdx.y * IB;
    for (int e = t; e < FB * COUT * CIN; e += THREADS) {    // Qf[phi][co][ci] -> Qs[phi][ci][co]
        const int f = e / (COUT * CIN), r = e - f * (COUT * CIN), co = r / CIN, ci = r - co * CIN;
        Qs[(f * CIN + ci) * COUT + co] = Qf[(size_t)(f0 + f) * COUT * CIN + r];
    }
    for (int e = t; e < IB * CIN; e += THREADS) {           // 4 consecutive phi = 32 contiguous bytes
        const int bl = e / CIN, ci = e - bl * CIN;
        float4 v0 = make_float4(0, 0, 0, 0), v1 = v0;
        if (b0 + bl < B) {
            const float4* src = reinterpret_cast<const float4*>(Xf + ((size_t)(b0 + bl) * CIN + ci) * F + f0);
            v0 = src[0]; v1 = src[1];
        }
        Xs[(0 * CIN + ci) * IB + bl] = make_float2(v0.x, v0.y);
        Xs[(1 * CIN + ci) * IB + bl] = make_float2(v0.z, v0.w);
        Xs[(2 * CIN + ci) * IB + bl] = make_float2(v1.x, v1.y);
        Xs[(3 * CIN + ci) * IB + bl] = make_float2(v1.z, v1.w);
    }
    __syncthreads();
    const int tb = t & 15, tc = t >> 4;                     // b group (4 images), co group (4 channels)
    float2 acc[FB][4][4];
#pragma unroll
    for (int f = 0; f < FB; ++f)
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[f][i][j] = make_float2(0.f, 0.f);
#pragma unroll 1
    for (int ci = 0; ci < CIN; ++ci) {
#pragma unroll
        for (int f = 0; f < FB; ++f) {
            const float4* qp = reinterpret_cast<const float4*>(Qs + (f * CIN + ci) * COUT + tc * 4);
            const float4* xp = reinterpret_cast<const float4*>(Xs + (f * CIN + ci) * IB + tb * 4);
            const float4 q01 = qp[0], q23 = qp[1], x01 = xp[0], x23 = xp[1];
            const float2 q[4] = {{q01.x, q01.y}, {q01.z, q01.w}, {q23.x, q23.y}, {q23.z, q23.w}};
            const float2 xv[4] = {{x01.x, x01.y}, {x01.z, x01.w}, {x23.x, x23.y}, {x23.z, x23.w}};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) cfma(acc[f][i][j], q[i], xv[j]);
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int b = b0 + tb * 4 + j;
        if (b >= B) continue;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float4* dst = reinterpret_cast<float4*>(Yf + ((size_t)b * COUT + tc * 4 + i) * F + f0);
            dst[0] = make_float4(acc[0][i][j].x, acc[0][i][j].y, acc[1][i][j].x, acc[1][i][j].y);
            dst[1] = make_float4(acc[2][i][j].x, acc[2][i][j].y, acc[3][i][j].x, acc[3][i][j].y);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// stride-2 Cayley conv = space-to-depth + real orthogonal 1x1 conv + bias + GroupSort
//   y[b][o][w][h] = sum_{c,k1,k2} Q[o][4c + 2k1 + k2] * x[b][c][2w+k1][2h+k2] + bias[o]
// CTA = IMGS images, 256 threads, thread tile 4 o x 8 pixels.
// ------------------------------------------------------------------------------------------------
template <int COUT, int CIN, int n, int IMGS>   // CIN = channels before s2d; n = input spatial size
__global__ void __launch_bounds__(256) conv1x1_s2d_kernel(const float* __restrict__ x, const float* __restrict__ Q,
                                                          const float* __restrict__ bias, int B, int group_sort,
                                                          float* __restrict__ y) {
    constexpr int K = 4 * CIN, m = n / 2, P = m * m, PT = P * IMGS;
    static_assert((COUT / 4) * (PT / 8) == 256, "thread tiling");
    extern __shared__ float4 smem4[];
    float* xs = reinterpret_cast<float*>(smem4);   // [K][PT]  (reused for the pre-activation output [COUT][PT])
    float* Qs = xs + K * PT;                       // [K][COUT]
    const int t = threadIdx.x;
    const int b0 = blockIdx.x * IMGS;
    for (int e = t; e < COUT * K; e += 256) {      // Q[o][k] -> Qs[k][o]
        const int o = e / K, k = e - o * K;
        Qs[k * COUT + o] = Q[e];
    }
    for (int e = t; e < IMGS * CIN * n * n; e += 256) {   // coalesced read, s2d scatter (layers.py:45-48)
        const int im = e / (CIN * n * n), r = e - im * (CIN * n * n);
        const int c = r / (n * n), px = r - c * (n * n), a = px / n, bcol = px - a * n;
        const float v = (b0 + im < B) ? x[(size_t)(b0 + im) * CIN * n * n + r] : 0.f;
        const int k = 4 * c + 2 * (a & 1) + (bcol & 1);
        xs[k * PT + im * P + (a >> 1) * m + (bcol >> 1)] = v;
    }
    __syncthreads();
    const int tp = t % (PT / 8), to = t / (PT / 8);
    float acc[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
#pragma unroll 4
    for (int k = 0; k < K; ++k) {
        const float4 q = *reinterpret_cast<const float4*>(Qs + k * COUT + to * 4);
        const float4 x0 = *reinterpret_cast<const float4*>(xs + k * PT + tp * 4);
        const float4 x1 = *reinterpret_cast<const float4*>(xs + k * PT + PT / 2 + tp * 4);
        const float qv[4] = {q.x, q.y, q.z, q.w};
        const float xv[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(qv[i], xv[j], acc[i][j]);
    }
    __syncthreads();                               // all reads of xs done -> reuse as out[COUT][PT]
    float* outs = xs;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float bs = bias ? bias[to * 4 + i] : 0.f;
        float* dst = outs + (to * 4 + i) * PT;
        *reinterpret_cast<float4*>(dst + tp * 4) =
            make_float4(acc[i][0] + bs, acc[i][1] + bs, acc[i][2] + bs, acc[i][3] + bs);
        *reinterpret_cast<float4*>(dst + PT / 2 + tp * 4) =
            make_float4(acc[i][4] + bs, acc[i][5] + bs, acc[i][6] + bs, acc[i][7] + bs);
    }
    __syncthreads();
    if (group_sort) {
        for (int e = t; e < (COUT / 2) * PT; e += 256) {
            const int o = e / PT, pp = e - o * PT, im = pp / P, px = pp - im * P;
            if (b0 + im >= B) continue;
            const float u = outs[o * PT + pp], w = outs[(o + COUT / 2) * PT + pp];
            y[((size_t)(b0 + im) * COUT + o) * P + px] = fmaxf(u, w);
            y[((size_t)(b0 + im) * COUT + o + COUT / 2) * P + px] = fminf(u, w);
        }
    } else {
        for (int e = t; e < COUT * PT; e += 256) {
            const int o = e / PT, pp = e - o * PT, im = pp / P, px = pp - im * P;
            if (b0 + im >= B) continue;
            y[((size_t)(b0 + im) * COUT + o) * P + px] = outs[e];
        }
    }
}

__global__ void space_to_depth_kernel(const float* __restrict__ x, int B, int C, int n, float* __restrict__ y) {
    const size_t total = (size_t)B * C * n * n;
    const int m = n / 2;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int bcol = e % n, a = (e / n) % n, c = (e / ((size_t)n * n)) % C;
        const size_t b = e / ((size_t)C * n * n);
        y[((b * (4 * C) + 4 * c + 2 * (a & 1) + (bcol & 1)) * m + (a >> 1)) * m + (bcol >> 1)] = x[e];
    }
}

__global__ void group_sort_kernel(const float* __restrict__ x, int B, int C, int inner, float* __restrict__ y) {
    const size_t half = (size_t)(C / 2) * inner, total = (size_t)B * half;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t b = e / half, r = e - b * half;
        const float u = x[b * 2 * half + r], w = x[b * 2 * half + half + r];
        y[b * 2 * half + r] = fmaxf(u, w);
        y[b * 2 * half + half + r] = fminf(u, w);
    }
}

// host-side dispatch helpers used by backbone.cu too
int launch_rfft2(const float* x, int n_img, int C, int n, const float* mu, const float* sd, float2* Xf, cudaStream_t st,
                 bool freq_major, bool x_u8 = false, int xrow = 0) {
    if (xrow <= 0) xrow = n / 2 + 1;
    if (x_u8 && !(n == 32 && !freq_major)) { set_error("rfft2: 8-bit input only for the 32x32 image layer"); return -1; }
    if (x_u8) rfft2_kernel<32, false, true><<<(n_img + FftCfg<32>::CH - 1) / FftCfg<32>::CH, FftCfg<32>::THREADS, 0, st>>>(x, n_img, C, mu, sd, Xf, xrow);
    else if (n == 32 && !freq_major) rfft2_kernel<32, false><<<(n_img + FftCfg<32>::CH - 1) / FftCfg<32>::CH, FftCfg<32>::THREADS, 0, st>>>(x, n_img, C, mu, sd, Xf, xrow);
    else if (n == 16 && !freq_major) rfft2_kernel<16, false><<<(n_img + FftCfg<16>::CH - 1) / FftCfg<16>::CH, FftCfg<16>::THREADS, 0, st>>>(x, n_img, C, mu, sd, Xf, xrow);
    else if (n == 16 && C % FftCfg<16>::CH == 0) rfft2_kernel<16, true><<<(n_img + FftCfg<16>::CH - 1) / FftCfg<16>::CH, FftCfg<16>::THREADS, 0, st>>>(x, n_img, C, mu, sd, Xf, xrow);
    else { set_error("rfft2: unsupported n=%d (16 or 32), freq_major=%d", n, (int)freq_major); return -1; }
    count_launch(1);
    return 0;
}
int launch_irfft2(const float2* Yf, int B, int C, int n, const float* bias, int gs, float* y, cudaStream_t st,
                  int freq_major) {
    if (n == 32 && C % FftCfg<32>::CH == 0 && !freq_major) {
        using Cf = FftCfg<32>;
        const size_t smem = (size_t)Cf::CH * 32 * (Cf::RS * sizeof(float2) + 33 * sizeof(float)) + (Cf::CH + 2) * sizeof(float2);
        cudaError_t e = cudaFuncSetAttribute(irfft2_kernel<32, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return cuda_fail(e, "irfft2 smem attr");
        irfft2_kernel<32, 0><<<B * (C / Cf::CH), FftCfg<32>::THREADS, smem, st>>>(Yf, B, C, bias, gs, y);
    } else if (n == 16 && C % FftCfg<16>::CH == 0) {
        using Cf = FftCfg<16>;
        const size_t smem = (size_t)Cf::CH * 16 * (Cf::RS * sizeof(float2) + 17 * sizeof(float)) + (Cf::CH + 2) * sizeof(float2);
        if (freq_major == 2) irfft2_kernel<16, 2><<<B * (C / Cf::CH), FftCfg<16>::THREADS, smem, st>>>(Yf, B, C, bias, gs, y);
        else if (freq_major) irfft2_kernel<16, 1><<<B * (C / Cf::CH), FftCfg<16>::THREADS, smem, st>>>(Yf, B, C, bias, gs, y);
        else irfft2_kernel<16, 0><<<B * (C / Cf::CH), FftCfg<16>::THREADS, smem, st>>>(Yf, B, C, bias, gs, y);
    } else { set_error("irfft2: unsupported n=%d C=%d", n, C); return -1; }
    count_launch(1);
    return 0;
}
// Per-frequency complex mixing as a real GEMM operand: y = Q x with interleaved (re, im) rows / columns
//   W[phi][2 co + 0][2 ci + 0] =  Re Q   W[phi][2 co + 0][2 ci + 1] = -Im Q
//   W[phi][2 co + 1][2 ci + 0] =  Im Q   W[phi][2 co + 1][2 ci + 1] =  Re Q          (Q = Qf[phi][co][ci], mix_kernel)
__global__ void mix_real_weights_kernel(const float2* __restrict__ Qf, long long total, float* __restrict__ W, int cin) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const float2 q = Qf[e];
        const long long row = e / cin;            // phi * cout + co
        const int ci = (int)(e - row * cin);
        float* w0 = W + (row * 2) * (2 * cin) + 2 * ci;
        float* w1 = w0 + 2 * cin;
        w0[0] = q.x; w0[1] = -q.y;
        w1[0] = q.y; w1[1] = q.x;
    }
}
int launch_mix_real_weights(const float2* Qf, int F, int cout, int cin, float* W, cudaStream_t st) {
    const long long total = (long long)F * cout * cin;
    mix_real_weights_kernel<<<(unsigned)((total + 255) / 256 < 2048 ? (total + 255) / 256 : 2048), 256, 0, st>>>(Qf, total, W, cin);
    count_launch(1);
    return 0;
}

template <int COUT, int CIN> static int launch_mix_t(const float2* Xf, const float2* Qf, int B, int F, float2* Yf, cudaStream_t st) {
    const size_t smem = (size_t)4 * CIN * (COUT + 64) * sizeof(float2);
    cudaError_t e = cudaFuncSetAttribute(mix_kernel<COUT, CIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "mix smem attr");
    mix_kernel<COUT, CIN><<<dim3(F / 4, (B + 63) / 64), (COUT / 4) * 16, smem, st>>>(Xf, Qf, B, F, Yf);
    count_launch(1);
    return 0;
}
int launch_mix(const float2* Xf, const float2* Qf, int B, int F, int cout, int cin, float2* Yf, cudaStream_t st) {
    if (F % 4) { set_error("mix: F=%d not a multiple of 4", F); return -1; }
    if (cout == 32 && cin == 3) return launch_mix_t<32, 3>(Xf, Qf, B, F, Yf, st);
    if (cout == 64 && cin == 32) return launch_mix_t<64, 32>(Xf, Qf, B, F, Yf, st);
    if (cout == 32 && cin == 32) return launch_mix_t<32, 32>(Xf, Qf, B, F, Yf, st);
    set_error("mix: unsupported (cout=%d, cin=%d)", cout, cin);
    return -1;
}
int launch_conv1x1(const float* x, const float* Q, const float* bias, int B, int cin, int cout, int n, int gs, float* y,
                   cudaStream_t st) {
    if (cin == 32 && cout == 32 && n == 32) {
        auto k = conv1x1_s2d_kernel<32, 32, 32, 1>;
        const size_t smem = (size_t)(128 * 256 + 128 * 32) * sizeof(float);
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return cuda_fail(e, "conv1x1 smem attr");
        k<<<B, 256, smem, st>>>(x, Q, bias, B, gs, y);
    } else if (cin == 64 && cout == 64 && n == 16) {
        auto k = conv1x1_s2d_kernel<64, 64, 16, 2>;
        const size_t smem = (size_t)(256 * 128 + 256 * 64) * sizeof(float);
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return cuda_fail(e, "conv1x1 smem attr");
        k<<<(B + 1) / 2, 256, smem, st>>>(x, Q, bias, B, gs, y);
    } else { set_error("conv1x1_s2d: unsupported (cin=%d, cout=%d, n=%d)", cin, cout, n); return -1; }
    count_launch(1);
    return 0;
}

}  // namespace fiode

using namespace fiode;

extern "C" size_t fiode_cayley_conv_fwd_workspace_bytes(int B, int cin, int cout, int n) {
    const size_t F = (size_t)n * (n / 2 + 1);
    return (size_t)B * (cin + cout) * F * sizeof(float2) + 256;
}

extern "C" int fiode_cayley_conv_fwd(const float* x, const float* Qf, const float* bias, int B, int cin, int cout,
                                     int n, int group_sort, float* y, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!x || !Qf || !y || !ws) { set_error("fiode_cayley_conv_fwd: null pointer"); return -1; }
    if (ws_bytes < fiode_cayley_conv_fwd_workspace_bytes(B, cin, cout, n)) { set_error("fiode_cayley_conv_fwd: workspace too small"); return -1; }
    const int F = n * (n / 2 + 1);
    float2* Xf = reinterpret_cast<float2*>(ws);
    float2* Yf = Xf + (size_t)B * cin * F;
    int rc;
    if ((rc = launch_rfft2(x, B * cin, cin, n, nullptr, nullptr, Xf, st, false))) return rc;
    if ((rc = launch_mix(Xf, reinterpret_cast<const float2*>(Qf), B, F, cout, cin, Yf, st))) return rc;
    if ((rc = launch_irfft2(Yf, B, cout, n, bias, group_sort, y, st, false))) return rc;
    FIODE_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int fiode_conv1x1_s2d_fwd(const float* x, const float* Q, const float* bias, int B, int cin, int cout, int n,
                                     int group_sort, float* y, void* stream) {
    if (!x || !Q || !y) { set_error("fiode_conv1x1_s2d_fwd: null pointer"); return -1; }
    int rc = launch_conv1x1(x, Q, bias, B, cin, cout, n, group_sort, y, (cudaStream_t)stream);
    if (rc) return rc;
    FIODE_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int fiode_space_to_depth(const float* x, int B, int C, int n, float* y, void* stream) {
    if (!x || !y || (n & 1)) { set_error("fiode_space_to_depth: bad args"); return -1; }
    const size_t total = (size_t)B * C * n * n;
    space_to_depth_kernel<<<(unsigned)((total + 255) / 256 < 65535 ? (total + 255) / 256 : 65535), 256, 0, (cudaStream_t)stream>>>(x, B, C, n, y);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}

extern "C" int fiode_group_sort(const float* x, int B, int C, int inner, float* y, void* stream) {
    if (!x || !y || (C & 1)) { set_error("fiode_group_sort: bad args"); return -1; }
    const size_t total = (size_t)B * (C / 2) * inner;
    group_sort_kernel<<<(unsigned)((total + 255) / 256 < 65535 ? (total + 255) / 256 : 65535), 256, 0, (cudaStream_t)stream>>>(x, B, C, inner, y);
    FIODE_CUDA(cudaGetLastError());
    count_launch(1);
    return 0;
}
