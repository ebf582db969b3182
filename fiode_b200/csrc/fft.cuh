// fiode_b200/csrc/fft.cuh -- register-resident radix-2 FFTs (N = 8, 16, 32) for the Cayley
// convolutions (libs/ortho_conv/layers.py:78, 83 use torch.fft.rfft2 / irfft2 with n in {32,16}).
// One thread owns one whole transform; every index is a compile-time constant after unrolling,
// so the twiddles come straight out of the constant bank.
#pragma once
#include <cuda_runtime.h>

namespace fiode {

// exp(-2*pi*i*k/32), k = 0..15  (cos, -sin)
static __device__ __constant__ const float2 kTw32[16] = {
    {1.0000000000000000f, -0.0000000000000000f}, {0.9807852804032304f, -0.1950903220161283f},
    {0.9238795325112867f, -0.3826834323650898f}, {0.8314696123025452f, -0.5555702330196022f},
    {0.7071067811865476f, -0.7071067811865476f}, {0.5555702330196023f, -0.8314696123025452f},
    {0.3826834323650898f, -0.9238795325112867f}, {0.1950903220161283f, -0.9807852804032304f},
    {0.0000000000000000f, -1.0000000000000000f}, {-0.1950903220161282f, -0.9807852804032304f},
    {-0.3826834323650897f, -0.9238795325112867f}, {-0.5555702330196020f, -0.8314696123025455f},
    {-0.7071067811865475f, -0.7071067811865476f}, {-0.8314696123025453f, -0.5555702330196022f},
    {-0.9238795325112867f, -0.3826834323650899f}, {-0.9807852804032304f, -0.1950903220161286f}};

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cconj(float2 a) { return make_float2(a.x, -a.y); }
__device__ __forceinline__ void cfma(float2& acc, float2 a, float2 b) {
    acc.x = fmaf(a.x, b.x, acc.x); acc.x = fmaf(-a.y, b.y, acc.x);
    acc.y = fmaf(a.x, b.y, acc.y); acc.y = fmaf(a.y, b.x, acc.y);
}

template <int N> __host__ __device__ constexpr int bitrev(int i) {
    int r = 0;
    for (int b = 1; b < N; b <<= 1) { r = (r << 1) | (i & 1); i >>= 1; }
    return r;
}

// In-place decimation-in-frequency FFT; result lands in bit-reversed order: X[bitrev(i)] = x[i].
// INV = true uses the conjugate twiddles (unnormalised inverse).
template <int N, bool INV> __device__ __forceinline__ void fft_dif(float2 (&x)[N]) {
#pragma unroll
    for (int m = N / 2; m >= 1; m >>= 1) {
#pragma unroll
        for (int blk = 0; blk < N; blk += 2 * m) {
#pragma unroll
            for (int j = 0; j < m; ++j) {
                const float2 a = x[blk + j], b = x[blk + j + m];
                x[blk + j] = cadd(a, b);
                const float2 d = csub(a, b);
                const int tw = j * (16 / m);           // W_{2m}^j = W_32^{j*16/m}
                if (tw == 0) {
                    x[blk + j + m] = d;
                } else if (tw == 8) {                  // -i (forward) / +i (inverse)
                    x[blk + j + m] = INV ? make_float2(-d.y, d.x) : make_float2(d.y, -d.x);
                } else {
                    const float2 w = INV ? cconj(kTw32[tw]) : kTw32[tw];
                    x[blk + j + m] = cmul(d, w);
                }
            }
        }
    }
}

}  // namespace fiode
