// fiode_b200/csrc/rows.cuh -- declarations shared by the row-streaming RHS kernels (FFMA path rhs_rows.cu,
// tensor-core path certify_tc.cu).
#pragma once
#include "common.cuh"

namespace fiode {

constexpr int kIt = 32;       // max bisection candidates tracked (>= max_iter)

enum { kStore = 0, kCertify = 1 };

struct RowsArgs {
    fiode_dyn_weights w;
    // kStore
    const float* h; const float* s; long long N; long long s_rows; long long s_group;
    float* ft_out; unsigned* notconv;              // (N,10), 1 word
    const float* m1; const float* m2;              // optional dropout masks (N,128), already scaled by 1/(1-p)
    float* z1_out; float* z2_out;                  // optional activations (N,128) for the backward
    float drop_p; unsigned long long drop_seed;    // in-kernel Philox dropout when drop_p > 0 and m1 == m2 == NULL
    const unsigned long long* drop_step;           // device counter = Philox offset of this call (NULL: 0)
    // kCertify
    const uint8_t* gridz; const float* lut; int T; long long Ng; long long chunk; int n_chunks;
    const int* labels; int M; int P;               // images in this launch, CTAs per image
    float sa, sqrtn, sqrt2, dist, inv_T;           // fp32 roundings of the python-double constants
    unsigned* cmax;                                // (M, n_chunks, 2, kIt) ordered-uint maxima
    unsigned* cnot;                                // (M, n_chunks) not-converged OR words
};


// defined in rhs_rows.cu
// ftilde (N,10) [post scale_nominal] + OR of the not-converged words of every row; optional dropout masks / activations
int launch_rows_store(const fiode_dyn_weights* w, const float* h, const float* s, long long N, long long s_rows,
                      long long s_group, float* ft, unsigned* notconv, const float* m1, const float* m2, float* z1,
                      float* z2, cudaStream_t st, float drop_p = 0.f, unsigned long long drop_seed = 0,
                      const unsigned long long* drop_step = nullptr);
int launch_certify_ffma(const RowsArgs& a, cudaStream_t st);
int launch_certify_finish(const unsigned* cmax, const unsigned* cnot, const unsigned* status, int M, int n_chunks,
                          int max_iter, float kappa, float* out, int* iters, cudaStream_t st);
// defined in certify_tc.cu
int launch_certify_tc(const RowsArgs& a, cudaStream_t st);
int mlp_path();   // 0 = tensor cores (tcgen05, fp16x2 split), 1 = FFMA

}  // namespace fiode
