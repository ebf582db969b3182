// fiode_b200/csrc/gemm_tc.cuh -- host-side interface of the tensor-core GEMMs (gemm_tc.cu)
#pragma once
#include "common.cuh"

namespace fiode {
size_t gemm_tc_packed_bytes(int N, int K);
bool gemm_tc_supported(int N, int K, int gs);
int gemm_tc_tile(int N, int gs);
int launch_pack_w(const float* W, int N, int K, int NT, int gs, void* Wp, cudaStream_t st, int nbatch = 1);
int launch_linear_tc_batched(const float* X, long long lda, long long sA, const void* Wp, long long M, int N, int K, int nbatch,
                             float* y, long long ldy, long long sO, cudaStream_t st);
int launch_linear_tc(const float* X, long long lda, const void* Wp, const float* bias, long long M, int N, int K, int gs,
                     float* y, long long ldy, cudaStream_t st);
int launch_conv1x1_tc(const float* x, const void* Wp, const float* bias, int B, int cin, int cout, int n, int gs, float* y,
                      cudaStream_t st);
int mlp_path();   // certify_tc.cu: 0 = tensor cores, 1 = FFMA
}  // namespace fiode
