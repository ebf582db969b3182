/* fiode_b200.h -- C ABI of libfiode_b200.so (hand-written CUDA for sm_100a).
 *
 * The reference (yjhuangcd/FI-ODE) has NO FFI on this path: it is eager PyTorch calling
 * torchdiffeq.  The entry points below are therefore what a binding for the path *would* bind:
 * one per replaceable seam of SURVEY.md section 8(b).  Each cites the reference interface it replaces
 * (paths relative to the reference checkout).
 *
 * Conventions
 *   - every function returns 0 on success, <0 on error; fiode_last_error() gives the message
 *     (thread-local);
 *   - all tensor arguments are raw DEVICE pointers to contiguous fp32 (unless stated), owned by
 *     the caller; nothing is allocated behind the caller's back -- scratch comes from
 *     *_workspace_bytes() + a caller-provided workspace;
 *   - `stream` is a cudaStream_t passed as void*; all work is stream-ordered, no host sync;
 *   - no exceptions, no torch types.
 */
#ifndef FIODE_B200_H
#define FIODE_B200_H
#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define FIODE_API __attribute__((visibility("default")))
#else
#define FIODE_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

FIODE_API const char* fiode_last_error(void);
FIODE_API int fiode_version(void);
/* number of kernels this library has launched since load / since the last reset (bench.py's
 * gpu_launches claim is read from here) */
FIODE_API long long fiode_launch_count(void);
FIODE_API void fiode_reset_launch_count(void);
/* Process-wide A/B switch (default on; env FIODE_FORK=0 = off): the entry points that compute a layer's weight / bias gradient
 * next to its data gradient (fiode_linear_bwd, fiode_conv1x1_s2d_bwd, fiode_cayley_conv_layer_bwd, fiode_fi_loss_bwd,
 * fiode_cayley_linear_bwd*) run the two on the caller's stream and on an internal branch stream that is joined before the call
 * returns.  Results are identical either way (same kernels, same order per output).  No counterpart in the reference.     */
FIODE_API void fiode_set_side_streams(int on);

/* ---- Cayley transform (weight preparation, cached per weight version by the host) ----------
 * replaces libs/ortho_conv/layers.py:50-60 `cayley` as used by CayleyLinear.forward (:103-104)
 * and CayleyConv.forward (:82).                                                              */

/* Q = cayley(alpha * W / ||W||_F), W row-major (cout, cin) real.  Q row-major (cout, cin).
 * alpha is read from device memory (1 float).                                               */
FIODE_API size_t fiode_cayley_linear_workspace_bytes(int cout, int cin);
FIODE_API int fiode_cayley_linear(const float* W, const float* alpha, int cout, int cin, float* Q,
                        void* ws, size_t ws_bytes, void* stream);

/* lazily-created CayleyConv.alpha = || wfft ||_F over the half spectrum (layers.py:80-81).
 * w: (cout, cin, k, k); n = spatial size seen by the layer.  alpha_out: 1 float (device).     */
FIODE_API int fiode_conv_alpha_init(const float* w, int cout, int cin, int k, int n, float* alpha_out,
                          void* ws, size_t ws_bytes, void* stream);
/* Per-frequency orthogonal mixing matrices of a CayleyConv (layers.py:74-82):
 * Qf[phi] = cayley(alpha * wfft[phi] / ||wfft||), phi = u*(n/2+1)+v, stored as interleaved
 * complex64 (F, cout, cin).  For k == 1 the spectrum is frequency independent and real
 * (SURVEY App. C-1): then Qf is (cout, cin) real fp32.                                       */
FIODE_API size_t fiode_cayley_conv_workspace_bytes(int cout, int cin, int k, int n);
FIODE_API int fiode_cayley_conv(const float* w, const float* alpha, int cout, int cin, int k, int n,
                      float* Qf, void* ws, size_t ws_bytes, void* stream);

/* ---- backbone forward ------------------------------------------------------------------------
 * replaces models.py:23-35 Normalize + libs/ortho_conv/models_ortho.py:12-28 KWLarge_Concat
 * (4 x [CayleyConv, GroupSort], Flatten, 2 x [CayleyLinear, GroupSort], fc) in eval mode.     */
typedef struct fiode_backbone_weights {
    const float* mu;   /* (3)  */
    const float* std;  /* (3)  */
    const float* Qf0;  /* (544, 32, 3) complex64 interleaved : conv 3->32 @32x32, k=3      */
    const float* b0;   /* (32) */
    const float* Q2;   /* (32, 128) real : stride-2 conv as 1x1 on space-to-depth, @16x16  */
    const float* b2;   /* (32) */
    const float* Qf4;  /* (144, 64, 32) complex64 : conv 32->64 @16x16, k=3                */
    const float* b4;   /* (64) */
    const float* Q6;   /* (64, 256) real : stride-2 conv as 1x1, @8x8                      */
    const float* b6;   /* (64) */
    const float* Q9;   /* (512, 4096) */
    const float* b9;   /* (512) */
    const float* Q11;  /* (512, 512) */
    const float* b11;  /* (512) */
    const float* Qfc;  /* (10, 512) */
    const float* bfc;  /* (10) */
    const void* packed; /* NULL, or a block filled by fiode_backbone_pack for THESE weights: skips the per-call packing */
} fiode_backbone_weights;
FIODE_API size_t fiode_backbone_workspace_bytes(int B);
/* Kernel-ready copies of the prepared matrices (fp16x2-split tensor-core tiles, model.4's spectrum as real matrices,
 * model.0's transposed spectrum).  The reference caches Q per weight version in eval mode (layers.py:77-83,
 * 95-103: `self.Q` reused while not training); a caller that does the same keeps one block per weight version and
 * sets `packed`, so that a forward pass launches no weight preparation at all.                                   */
/* images per internal slab of fiode_backbone_fwd (each kernel of the chain is launched once per slab) */
FIODE_API int fiode_backbone_slab_images(void);
/* model.0 (libs/ortho_conv/layers.py:72-86 at 3 -> 32 channels): 2 (default) = mixing + the inverse transform as two DFT
 * GEMMs on tcgen05, warp-specialised (csrc/dft_tc.cu); 1 = its first, single-role version; 0 = mixing + inverse FFT on the
 * CUDA cores; same results to fp32 round-off */
FIODE_API void fiode_backbone_set_conv0_path(int path);
/* model.0 alone through one of the three paths (tests / tools/debug_conv0.py): x (B,3,32,32) -> a0 (B,32,32,32) after bias and
 * GroupSort; w->packed must be set; ws as for fiode_backbone_fwd, B <= fiode_backbone_slab_images() */
FIODE_API int fiode_debug_conv0(const fiode_backbone_weights* w, const float* x, int B, int path, float* a0_out, void* ws,
                                size_t ws_bytes, void* stream);
FIODE_API size_t fiode_backbone_packed_bytes(void);
FIODE_API int fiode_backbone_pack(const fiode_backbone_weights* w, void* packed, size_t packed_bytes, void* stream);
/* x: (B,3,32,32) in [0,1]; s_out: (B,10) */
FIODE_API int fiode_backbone_fwd(const fiode_backbone_weights* w, const float* x, int B, float* s_out,
                       void* ws, size_t ws_bytes, void* stream);
/* Same, on the dataset's own 8-bit pixels: x (B,3,32,32) uint8.  The first kernel applies torchvision's ToTensor
 * (float(p) / 255, dataset_loaders.py:76-78, the CIFAR-10 test transform) while loading, so the result is bit-identical to
 * fiode_backbone_fwd on x / 255 and the host -> device copy is a quarter of the float one.                        */
FIODE_API int fiode_backbone_fwd_u8(const fiode_backbone_weights* w, const unsigned char* x, int B, float* s_out,
                          void* ws, size_t ws_bytes, void* stream);

/* Per-stage device times of the LAST fiode_backbone_fwd (CUDA events on its stream; measurement aid for the
 * bench's roofline numbers, tensor-core path, batch <= one slab).  11 stages, see backbone.cu.  Synchronises. */
FIODE_API void fiode_backbone_enable_stage_timing(int on);
FIODE_API int fiode_backbone_stage_times(float* ms_out, int n);

/* single layers (the reference's nn.Module.forward seams; used by the per-layer parity tests) */
/* CayleyConv.forward, k=3 (layers.py:72-86): x (B,cin,n,n) -> y (B,cout,n,n), optional fused
 * GroupSort (layers.py:127-131).  n in {8,16,32}.                                             */
FIODE_API size_t fiode_cayley_conv_fwd_workspace_bytes(int B, int cin, int cout, int n);
FIODE_API int fiode_cayley_conv_fwd(const float* x, const float* Qf, const float* bias, int B, int cin, int cout,
                          int n, int group_sort, float* y, void* ws, size_t ws_bytes, void* stream);
/* F.linear(X, Q, b) (layers.py:105): X (M,K), Q (N,K), y (M,N); optional fused GroupSort.     */
FIODE_API size_t fiode_linear_workspace_bytes(int M, int N, int K);
FIODE_API int fiode_linear_fwd(const float* X, const float* Q, const float* bias, int M, int N, int K,
                     int group_sort, float* y, void* ws, size_t ws_bytes, void* stream);
/* fiode_linear_fwd without a workspace argument.  No counterpart in the reference: added
 * for callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The scratch
 * (split-K partial sums / re-tiled weights) is a built-in device array of fiode_linear_static_max_bytes() = 16 MiB that every
 * GPU loading the library holds until exit; the call allocates nothing (no cudaMalloc / cudaFree, no allocating library call)
 * and fails with a clear fiode_last_error() when the product it takes needs more scratch than that (never more than
 * fiode_linear_workspace_bytes(M, N, K); e.g. M = 512, N = 512, K = 4096 is 256 bytes over).  All calls on a device
 * share the array (it keeps the last call's contents; nothing in it needs clearing): calls that overlap on the GPU race.
 * Same results as fiode_linear_fwd, bit for bit.  Make the first call on each device outside stream capture (it resolves the
 * array's device address once per device).                                                                                  */
FIODE_API size_t fiode_linear_static_max_bytes(void);
FIODE_API int fiode_linear_fwd_static(const float* X, const float* Q, const float* bias, int M, int N, int K,
                            int group_sort, float* y, void* stream);
/* stride-2 CayleyConv (StridedConv pre-hook layers.py:24-48 + CayleyConv.forward with k=1, whose
 * spectrum is frequency independent and real): space-to-depth + real orthogonal
 * 1x1 conv + bias (+GroupSort).  x (B,cin,n,n) -> y (B,cout,n/2,n/2); Q (cout, 4*cin).          */
FIODE_API int fiode_conv1x1_s2d_fwd(const float* x, const float* Q, const float* bias, int B, int cin, int cout,
                          int n, int group_sort, float* y, void* stream);
/* einops 'b c (w 2) (h 2) -> b (c 4) w h' (layers.py:45-48) */
FIODE_API int fiode_space_to_depth(const float* x, int B, int C, int n, float* y, void* stream);
FIODE_API int fiode_group_sort(const float* x, int B, int C, int inner, float* y, void* stream);

/* ---- training: adjoints of the backbone layers and of the Cayley transform ---------------------
 * The reference differentiates these through autograd (loss.backward(), pl_modules.py:504-506).  Each
 * entry point is the hand-written adjoint of the forward named beside it; optional outputs may be NULL.
 * Batch sums are ordered (no float atomics): gradients are bit-reproducible.                         */
/* Normalize.forward (models.py:23-26): y = (x - mu[c]) / std[c], x (B,C,inner).                      */
FIODE_API int fiode_normalize(const float* x, int B, int C, int inner, const float* mu, const float* std, float* y,
                    void* stream);
/* adjoint of GroupSort (layers.py:127-131): x_pre is the layer's input; ties split evenly like torch.max/min. */
FIODE_API int fiode_group_sort_bwd(const float* x_pre, const float* gy, int B, int C, int inner, float* gx, void* stream);
/* gb[c] = sum over batch and inner of gy (B,C,inner): adjoint of the `+ bias` of every layer.          */
FIODE_API int fiode_bias_grad(const float* gy, int B, int C, int inner, float* gb, void* stream);
/* adjoint of F.linear(X, Q, b) (layers.py:105): gX = gY Q (M,K), gQ = gY^T X (N,K), gb = column sums.  */
FIODE_API size_t fiode_linear_bwd_workspace_bytes(int M, int N, int K);
FIODE_API int fiode_linear_bwd(const float* X, const float* Q, const float* gY, int M, int N, int K, float* gX, float* gQ,
                     float* gb, void* ws, size_t ws_bytes, void* stream);
/* adjoint of fiode_conv1x1_s2d_fwd: gx (B,cin,n,n), gQ (cout,4cin), gb (cout).                          */
FIODE_API size_t fiode_conv1x1_s2d_bwd_workspace_bytes(int B, int cin, int cout, int n);
FIODE_API int fiode_conv1x1_s2d_bwd(const float* x, const float* Q, const float* gy, int B, int cin, int cout, int n,
                          float* gx, float* gQ, float* gb, void* ws, size_t ws_bytes, void* stream);
/* adjoint of fiode_cayley_conv_fwd (CayleyConv.forward layers.py:78-85, k = 3): gx = irfft2(Q^H rfft2(gy)),
 * gQf[phi] = c_phi / n^2 sum_b rfft2(gy)[b,:,phi] rfft2(x)[b,:,phi]^H  (F,cout,cin) complex64, gb (cout).   */
FIODE_API size_t fiode_cayley_conv_layer_bwd_workspace_bytes(int B, int cin, int cout, int n);
FIODE_API int fiode_cayley_conv_layer_bwd(const float* x, const float* Qf, const float* gy, int B, int cin, int cout, int n,
                                float* gx, float* gQf, float* gb, void* ws, size_t ws_bytes, void* stream);
/* adjoint of fiode_cayley_linear (cayley(alpha W / ||W||), layers.py:50-60, 103-104): gQ (cout,cin) ->
 * gW (cout,cin), galpha (1).  Forward intermediates are recomputed.                                     */
FIODE_API size_t fiode_cayley_linear_bwd_workspace_bytes(int cout, int cin);
FIODE_API int fiode_cayley_linear_bwd(const float* W, const float* alpha, int cout, int cin, const float* gQ, float* gW,
                            float* galpha, void* ws, size_t ws_bytes, void* stream);
/* Same pair with (I + A)^-1 handed from the forward call to its adjoint: Minv (n, n), n = min(cout, cin).  The reference's
 * autograd keeps the forward's intermediates for backward (layers.py:50-60 under loss.backward()); without it the adjoint
 * inverts again (n = 512: ~0.5 ms per layer per step).                                                              */
FIODE_API int fiode_cayley_linear_save(const float* W, const float* alpha, int cout, int cin, float* Q, float* Minv,
                             void* ws, size_t ws_bytes, void* stream);
FIODE_API int fiode_cayley_linear_bwd_saved(const float* W, const float* alpha, int cout, int cin, const float* gQ,
                                  const float* Minv, float* gW, float* galpha, void* ws, size_t ws_bytes, void* stream);
/* Training: the forward transform (layers.py:103-104, evaluated in CayleyLinear.forward under autograd) that keeps what its
 * adjoint needs in ONE caller buffer of fiode_cayley_linear_ctx_floats(cout, cin) floats (16-byte aligned) -- the scale pair,
 * (I + A)^-1 and its transpose, the scaled working matrix and V^T, as autograd keeps a forward's intermediates; small layers
 * keep (I + A)^-1 only.  fiode_cayley_linear_bwd_ctx then starts at the incoming gradient.  Same Q as fiode_cayley_linear and
 * same (gW, galpha) as fiode_cayley_linear_bwd, bit for bit.                                                                */
FIODE_API size_t fiode_cayley_linear_ctx_floats(int cout, int cin);
FIODE_API int fiode_cayley_linear_fwd_ctx(const float* W, const float* alpha, int cout, int cin, float* Q, float* ctx,
                                void* ws, size_t ws_bytes, void* stream);
FIODE_API int fiode_cayley_linear_bwd_ctx(const float* W, const float* alpha, int cout, int cin, const float* gQ,
                                const float* ctx, float* gW, float* galpha, void* ws, size_t ws_bytes, void* stream);
/* Several small layers (fiode_cayley_small_ok: min(cout,cin) <= 64 and max*min <= 16384 -- the dynamics' 128x10 / 10x128
 * layers, fc 10x512, the stride-2 convs' 1x1 matrices) in ONE launch, one CTA per layer: the Cayley transforms of a training
 * step's layers are independent (libs/ortho_conv/layers.py:103-104 / :74-82 evaluate them layer by layer).  The lists are HOST
 * arrays of device pointers; mult[i] = 1 for a CayleyLinear, sqrt(n (n/2+1)) for a k = 1 CayleyConv (layers.py:80-81), NULL =
 * all 1.  Minv[i] (n x n, n = min(cout,cin)) is written by fwd when non-NULL and read by bwd when non-NULL (else recomputed).
 * Results are bit-identical with fiode_cayley_linear(_save) / fiode_cayley_linear_bwd(_saved) of the same layer.        */
FIODE_API int fiode_cayley_small_ok(int cout, int cin);
FIODE_API int fiode_cayley_small_batch_fwd(int count, const float* const* W, const float* const* alpha, const int* cout,
                                 const int* cin, const float* mult, float* const* Q, float* const* Minv, void* stream);
FIODE_API int fiode_cayley_small_batch_bwd(int count, const float* const* W, const float* const* alpha, const int* cout,
                                 const int* cin, const float* mult, const float* const* gQ, const float* const* Minv,
                                 float* const* gW, float* const* galpha, void* stream);
/* adjoint of fiode_cayley_conv (layers.py:74-82): gQf as produced by the layer adjoints above
 * (k = 3: (F,cout,cin) complex64; k = 1: real (cout,cin)) -> gw (cout,cin,k,k), galpha (1).              */
FIODE_API size_t fiode_cayley_conv_bwd_workspace_bytes(int cout, int cin, int k, int n);
FIODE_API int fiode_cayley_conv_bwd(const float* w, const float* alpha, int cout, int cin, int k, int n, const float* gQf,
                          float* gw, float* galpha, void* ws, size_t ws_bytes, void* stream);

/* ---- ODE right-hand side -----------------------------------------------------------------------
 * replaces dynamics/classification.py:96-126 (_h_dot_raw, eval_dot, eval_dot_light) with
 * barrier_projection/barrier_projection.py:217-269 (FastBarrierProjectionNoUpper.forward).
 * Q* are the Cayley-transformed matrices (CayleyLinear.Q), row-major (out, in).              */
typedef struct fiode_dyn_weights {
    const float* Q1;  /* hidden_to_mlp.Q (128,10) */
    const float* b1;  /* (128) */
    const float* Qx;  /* U_x.Q (128,10) */
    const float* bx;  /* (128) */
    const float* Q2;  /* mlp_to_mlp.Q (128,128) */
    const float* b2;  /* (128) */
    const float* Q3;  /* mlp_to_hidden.Q (10,128) */
    const float* b3;  /* (10) */
    float alpha_1, alpha_2, sigma_1;
    int scale_nominal;
    int max_iter; /* 30 (classification.py:63) */
    float tol;    /* 1e-4 */
} fiode_dyn_weights;

/* f = eval_dot_light(h, s): h (N,10); s (s_rows,10) with s_rows in {1, N} (row r uses
 * s[r % s_rows] when s_rows == 1, else s[r / s_group] -- s_group rows of h share one s row,
 * which is how compute_loss expands static_state, pl_modules.py:395-400).
 * The projection's stop iteration is global over the N rows of the call, like the reference.
 * Optional outputs (may be NULL): ftilde (N,10) pre-projection nominal, mu (N), n_iter (1 int). */
FIODE_API size_t fiode_rhs_workspace_bytes(long long N);
FIODE_API int fiode_rhs_fwd(const fiode_dyn_weights* w, const float* h, const float* s, long long N,
                  long long s_rows, long long s_group, float* f, float* ftilde, float* mu,
                  int* n_iter, void* ws, size_t ws_bytes, void* stream);

/* autograd.Function seam: FastBarrierProjectionNoUpper(max_iter, tol).apply(lower, nominal)
 * (barrier_projection.py:217-313).  forward -> v, mu; backward -> d_lower, d_nominal.          */
FIODE_API int fiode_barrier_project_fwd(const float* lower, const float* upper /* NULL = no upper */,
                              const float* nominal, long long N, int max_iter, float tol, float* v,
                              float* mu, int* n_iter, void* ws, size_t ws_bytes, void* stream);
FIODE_API int fiode_barrier_project_bwd(const float* grad_v, const float* v, const float* mu, const float* nominal,
                              long long N, float* d_lower, float* d_nominal, void* stream);
/* backward of the upper-bounded projection (barrier_projection.py:164-214: FastBarrierProjection.backward) from the saved
 * forward state (v, mu, nominal) -> d_lower, d_upper, d_nominal, each (N,10) */
FIODE_API int fiode_barrier_project_bwd_upper(const float* grad_v, const float* v, const float* mu, const float* nominal,
                                              long long N, float* d_lower, float* d_upper, float* d_nominal, void* stream);

/* ---- Lipschitz grid certification ----------------------------------------------------------------
 * replaces the per-image body of robustness/certify_lipschitz.py:104-143 (h_vdot :37-42, Lf_eta /
 * perturb :122-124, tie mask :126-128, running maxima :131-134) for M images at once.
 * gridz: (Ng,10) uint8 numerators z of the canonical label-0 grid eta = z/T (eval_utils.py:31-61);
 * the per-label column swap (eval_utils.py:64-69) is applied on load.  lut: (T+1) floats,
 * lut[z] = float32(z / T) (what grid.float() holds).  s: (M,10) static features; labels: (M).
 * `batches` reproduces the reference's chunking (:100-102, :116-119) because the projection's stop
 * iteration is global per chunk.  out: (M,2) = {max(h+perturb+kappa), max(h+kappa)};
 * iters (may be NULL): (M, n_chunks) bisection iterations the reference would have run.        */
FIODE_API size_t fiode_certify_workspace_bytes(int M, int n_chunks);
/* Which pipe evaluates the three dense layers of the dynamics MLP in the row-streaming kernels
 * (certification, FI loss, RHS over many rows): 0 = tcgen05 tensor cores with the fp16x2 split
 * (FP32 emulation: x = hi + lo, three MMAs, fp32 accumulate; default), 1 = FP32 FFMA.  The
 * environment variable FIODE_MLP_PATH=ffma selects 1 at load.  If an activation leaves the fp16x2
 * range (|x| >= 1023) the tensor-core kernels write NaN results instead of a verdict.            */
FIODE_API void fiode_set_mlp_path(int path);
FIODE_API int fiode_get_mlp_path(void);
FIODE_API int fiode_certify_grid(const fiode_dyn_weights* w, const uint8_t* gridz, long long Ng, int T,
                       const float* lut, const float* s, const int* labels, int M, int batches,
                       float kappa, float* out, int* iters, void* ws, size_t ws_bytes, void* stream);

/* ---- forward-invariance training workload ------------------------------------------------------------
 * Samplers (sampling/sampler.py): pure functions of pre-drawn Exp(1) variates e (the host draws them
 * with the framework's RNG, like torch.distributions.Exponential.sample at sampler.py:36,116,142).
 *   kind 0 UniformSimplexSampling  (:24-38)   e (n,10), one set shared by every image (:209-210)
 *   kind 1 CorrectConeSampling     (:104-128) e (B,n,10), y (B)
 *   kind 2 DecisionBoundarySampling(:130-153) e (B,n,9),  y (B)
 * writes rows [s0, s0+n) of every image's block of the composite sample tensor h (B,S,10)
 * (CompositeSampler.forward :195-216: cat on the sample axis, flatten to (B*S,10)).            */
FIODE_API int fiode_sampler_fill(int kind, const float* e, const int* y, int B, int n, int S, int s0, float* h,
                       void* stream);
/* lya_cands.py:72-94 DecisionBoundary.forward: h (N,10), y (N) -> V (N) = 1 + max_{j != y} p_j - p_y (log_mode: its log);
 * p = h (on_simplex) or softmax(h) */
FIODE_API int fiode_decision_boundary(const float* h, const int* y, long long N, int on_simplex, int log_mode, float* V,
                                      void* stream);

/* LyapunovLearning.compute_loss, order 1, act 'relu' (pl_modules.py:390-484) with the DecisionBoundary
 * candidate (lya_cands.py:79-94): h (B*S,10) samples, s (B,10) static features (row r uses s[r/S],
 * pl_modules.py:395-400), y (B) labels.  m1/m2: optional dropout masks (B*S,128), already scaled by
 * 1/(1-p) (classification.py:98,100; NULL = eval).  Saved for the backward: f, ftilde, mu, z1, z2,
 * V, Vdot.  out4 (device) = {loss, effective_batch_size, mean_active_constraints, bisection iterations}. */
FIODE_API size_t fiode_fi_loss_workspace_bytes(long long N, int B);
FIODE_API int fiode_fi_loss_fwd(const fiode_dyn_weights* w, const float* h, const float* s, const int* y, int B, int S,
                      float kappa, const float* m1, const float* m2, float* f, float* ftilde, float* mu,
                      float* z1, float* z2, float* V, float* Vdot, float* out4, void* ws, size_t ws_bytes,
                      void* stream);
/* Backward of the above (what loss.backward() computes through BarrierProjectionFn.backward
 * barrier_projection.py:272-311 and F.linear): gradients w.r.t. the four Cayley matrices Q* (same
 * layouts as fiode_dyn_weights), the biases (db1 is the gradient of hidden_to_mlp.bias AND of
 * U_x.bias), and the static features ds (B,10).  grad_loss = d(objective)/d(loss).              */
FIODE_API int fiode_fi_loss_bwd(const fiode_dyn_weights* w, const float* h, const float* s, const int* y, int B, int S,
                      float kappa, float grad_loss, const float* m1, const float* m2, const float* f,
                      const float* ftilde, const float* mu, const float* z1, const float* z2, const float* V,
                      const float* Vdot, float* dQ1, float* dQx, float* dQ2, float* dQ3, float* db1, float* db2,
                      float* db3, float* ds, void* ws, size_t ws_bytes, void* stream);

/* Options for the _opt variants of the two calls above.  kappa_dev: device scalar read instead of the by-value kappa, so that the
 * kappa warm-up (pl_modules.py:447-450, kappa_length = 30 000 steps in cifar_train.yaml) survives CUDA-graph replay.
 * dropout_p > 0 (with m1 == m2 == NULL): the nn.Dropout of classification.py:96-102 is drawn INSIDE the kernel
 * (Philox4x32-10 keyed by seed, offset *step_dev), masks never touch HBM; the forward increments *step_dev so every call /
 * graph replay draws fresh masks; the backward needs only dropout_p (a kept unit's mask value is 1 / (1 - p)).        */
typedef struct fiode_fi_options {
    const float* kappa_dev;
    float dropout_p;
    unsigned long long seed;
    unsigned long long* step_dev;
} fiode_fi_options;
FIODE_API int fiode_fi_loss_fwd_opt(const fiode_dyn_weights* w, const float* h, const float* s, const int* y, int B, int S,
                          float kappa, const float* m1, const float* m2, const fiode_fi_options* opt, float* f,
                          float* ftilde, float* mu, float* z1, float* z2, float* V, float* Vdot, float* out4, void* ws,
                          size_t ws_bytes, void* stream);
FIODE_API int fiode_fi_loss_bwd_opt(const fiode_dyn_weights* w, const float* h, const float* s, const int* y, int B, int S,
                          float kappa, float grad_loss, const float* m1, const float* m2, const fiode_fi_options* opt,
                          const float* f, const float* ftilde, const float* mu, const float* z1, const float* z2,
                          const float* V, const float* Vdot, float* dQ1, float* dQx, float* dQ2, float* dQ3, float* db1,
                          float* db2, float* db3, float* ds, void* ws, size_t ws_bytes, void* stream);
/* The (N,128) mask (0 or 1/(1-p)) the forward draws for layer 0 (after hidden_to_mlp + U_x) / 1 (after mlp_to_mlp) at the
 * CURRENT value of *step_dev: replaying a step with explicit masks (parity tests).  Synchronises the stream.        */
FIODE_API int fiode_dropout_mask(float dropout_p, unsigned long long seed, const unsigned long long* step_dev, int layer,
                       long long N, float* mask, void* stream);

/* ---- decision-boundary grid -----------------------------------------------------------------------
 * replaces robustness/eval_utils.py:31-61 sample_decision_boundary and :72-89
 * count_samples_decision_boundary: the set { z in Z^n_{>=0} : sum z = T, z_0 = max_{i>=1} z_i } in a
 * canonical order (by z_0, then lexicographic), as uint8 numerators (eta = z / T).  Rows
 * [first, first + count) are written to out (count, n) -- the grid can be generated in shards.   */
FIODE_API long long fiode_grid_count(int n, int T);
FIODE_API size_t fiode_grid_workspace_bytes(int n, int T);
FIODE_API int fiode_grid_enumerate(int n, int T, long long first, long long count, uint8_t* out, void* ws,
                         size_t ws_bytes, void* stream);

/* ---- Segway NN-controller forward invariance (control/) ----------------------------------------------
 * weights of libs/core/core/controllers/nn_controller.py:6-34 NNController(3 -> 32 -> 32 -> 1) as stored
 * by nn.Linear (out, in), and P of control/control_utils.py:215-234 LyaQuadratic (3,3) row-major.     */
typedef struct fiode_segway_weights {
    const float* W0;  /* model.0.weight (32,3)  */
    const float* b0;  /* model.0.bias   (32)    */
    const float* W1;  /* model.2.weight (32,32) */
    const float* b1;  /* model.2.bias   (32)    */
    const float* W2;  /* model.4.weight (1,32)  */
    const float* b2;  /* model.4.bias   (1)     */
    const float* P;   /* lya.P.weight   (3,3)   */
} fiode_segway_weights;
/* control/certify_segway.py:55-59: grid_uniform_3d (grid_sampler.py:109-117, np.meshgrid 'xy' order over
 * the axes d0 (phi), d1 (v), d2 (phi_dot), given as float32 arrays on the device), reject_sampling
 * (:91-98) to the shell level_lb <= ||P x||^2 <= level_ub, and val = SegwaySingleBarrierModel(eta)
 * (control_utils.py:97-106) over grid indices [first, first+count).  Kept points are written in grid
 * order to eta (M,3) / val (M); stats2 (device) = {M, max val}.  eta/val must hold `capacity` >= M rows. */
FIODE_API size_t fiode_segway_grid_workspace_bytes(void);
FIODE_API int fiode_segway_grid(const fiode_segway_weights* w, const float* d0, int n0, const float* d1, int n1,
                      const float* d2, int n2, float level_lb, float level_ub, long long first, long long count,
                      long long capacity, float* eta, float* val, float* stats2, void* ws, size_t ws_bytes,
                      void* stream);
/* libs/core/core/dynamics/system_dynamics.py:54-91 simulate (processed=True): zero-order hold of
 * u = ctrl(x) over each [ts[j], ts[j+1]], one dopri5 solve (torchdiffeq semantics, batch-wide RMS norm
 * over the B <= 32 trajectories) per interval.  x0 (B,3); ts (N) float32; xs (B,N,3); us (B,N-1,1);
 * stats4 (device) = {status, accepted steps, rejected steps, RHS evaluations}.                       */
FIODE_API int fiode_segway_simulate(const fiode_segway_weights* w, const float* x0, int B, const float* ts, int N,
                          double rtol, double atol, float* xs, float* us, int* stats4, void* stream);

/* ---- Segway adversarial training (control/train_segway.py:119-158) ----------------------------------------------------
 * The barrier model at arbitrary points and its gradients: val_r = sum(P^T (P f(x_r, u(x_r))) * x_r)
 * (control_utils.py:97-111 SegwaySingleBarrierModel.forward / forward_adv).
 *   fiode_segway_hdot      x (N,3) -> val (N)
 *   fiode_segway_hdot_bwd  gradients of sum_r gout_r val_r (gout NULL = all ones): grad_x (N,3) and / or grad_params
 *                          (fiode_segway_param_count() = 1226 floats: W0 96 | b0 32 | W1 1024 | b1 32 | W2 32 | b2 1 | P 9),
 *                          deterministic two-stage reduction through ws.  P's gradient is the reference graph's: through the inner
 *                          self.P(f) only, g (P x) f^T -- LyaQuadratic.h_dot refreshes P_t through .data (control_utils.py:230)
 *   fiode_segway_pgd_step  advertorch perturb_iterative (ord = inf), one iteration, in place on x_adv:
 *                          delta = clamp(x_adv - x0 + eps_iter sign(grad), -eps, eps); x_adv = clamp(x0 + delta, clip_min, clip_max) */
FIODE_API int fiode_segway_param_count(void);
FIODE_API size_t fiode_segway_hdot_bwd_workspace_bytes(long long N);
FIODE_API int fiode_segway_hdot(const fiode_segway_weights* w, const float* x, long long N, float* val, void* stream);
FIODE_API int fiode_segway_hdot_bwd(const fiode_segway_weights* w, const float* x, long long N, const float* gout, float* grad_x,
                                    float* grad_params, void* ws, size_t ws_bytes, void* stream);
FIODE_API int fiode_segway_pgd_step(float* x_adv, const float* x0, const float* grad, long long n, float eps, float eps_iter,
                                    float clip_min, float clip_max, void* stream);

/* ---- device-resident adaptive dopri5 -----------------------------------------------------------
 * replaces torchdiffeq==0.2.2 odeint(func, y0, t, rtol, atol, method='dopri5') as called at
 * models.py:236-241 with func = IVP.h_dot -> dyn_fun.ode_forward (classification.py:128-132).
 * One cooperative persistent kernel: 7 FSAL stages, batch-global RMS error norm, fp64 controller,
 * accept/reject and dense output all on the device -- zero host syncs.                         */
typedef struct fiode_solve_stats {
    int n_accept, n_reject, nfe;
    int status;        /* 0 ok; 1 'underflow in dt'; 2 'non-finite values in state'; 3 'max_num_steps exceeded' */
    double t_final;    /* right edge of the last accepted step */
    double last_dt;
    int n_logged;      /* attempts recorded below (first FIODE_LOG_STEPS) */
    int pad_;
    double dts[64];
    float ratios[64];
} fiode_solve_stats;
#define FIODE_LOG_STEPS 64

FIODE_API size_t fiode_dopri5_workspace_bytes(int B);
/* h0 (B,10); s (s_rows,10), s_rows in {1,B}; t: nt fp64 output times on the DEVICE (increasing);
 * out (nt,B,10); stats: device struct; rows_per_cta 0 = auto.                                 */
FIODE_API int fiode_dopri5_simplex(const fiode_dyn_weights* w, const float* h0, const float* s, int s_rows, int B,
                         const double* t, int nt, double rtol, double atol, long long max_num_steps,
                         float* out, fiode_solve_stats* stats, void* ws, size_t ws_bytes,
                         int rows_per_cta, void* stream);

/* The same solve with a caller-owned device log of EVERY step attempt (fiode_solve_stats keeps the first
 * FIODE_LOG_STEPS only): log_dts[i] = dt of attempt i as the fp64 controller held it, log_ratios[i] = its fp32 error
 * ratio (accepted iff <= 1).  Trained FI-ODE dynamics take hundreds of attempts; the differentiable solve
 * (train_ode pl_modules.py:490-500, PGD :189-199) replays the accepted ones.  Attempts beyond log_capacity are dropped;
 * the attempt count is n_accept + n_reject, so the caller re-runs with a larger log if it overflowed.          */
FIODE_API int fiode_dopri5_simplex_logged(const fiode_dyn_weights* w, const float* h0, const float* s, int s_rows, int B,
                                const double* t, int nt, double rtol, double atol, long long max_num_steps,
                                float* out, fiode_solve_stats* stats, void* ws, size_t ws_bytes, int rows_per_cta,
                                double* log_dts, float* log_ratios, long long log_capacity, void* stream);

/* G independent solves of B rows each (row g*B+r belongs to solve g) in one launch: one thread-block
 * cluster per solve, so every solve keeps ITS OWN batch-wide RMS norm and projection stop rule -- the
 * result of solve g equals fiode_dopri5_simplex on rows [g*B, (g+1)*B).  This is how whole batches
 * shard across SMs (and, one level up, across GPUs).  h0, s (1 | G*B rows), out (nt, G*B, 10),
 * stats: G structs.  cluster_size 0 = auto (8 CTAs x 32 rows for B = 256, 16 CTAs for B > 512); 1, 2, 4, 8 or 16 (16 is a
 * non-portable cluster size: granted on B200); B <= 64 * cluster_size, i.e. up to 1024 rows per solve.   */
FIODE_API int fiode_dopri5_simplex_batched(const fiode_dyn_weights* w, const float* h0, const float* s, int s_rows,
                                 int G, int B, const double* t, int nt, double rtol, double atol,
                                 long long max_num_steps, float* out, fiode_solve_stats* stats,
                                 int cluster_size, void* stream);

/* ---- differentiable solve: vector-Jacobian products of the right-hand side -------------------------
 * The reference back-propagates through torchdiffeq's dopri5 (use_adjoint=False, pl_modules.py:303; train_ode :490-500,
 * PGD :189-199).  fiode_b200/odeint.py replays the accepted steps the device solver logged and walks them backwards
 * with these calls (one stage each); the weight gradients of all stages are formed at the end from the stacked deltas.  */
FIODE_API size_t fiode_rhs_train_workspace_bytes(long long N);
FIODE_API int fiode_rhs_fwd_train(const fiode_dyn_weights* w, const float* h, const float* s, long long N, float* f,
                        float* ftilde, float* mu, float* z1, float* z2, void* ws, size_t ws_bytes, void* stream);
FIODE_API int fiode_dyn_transposes(const fiode_dyn_weights* w, float* Q1T, float* QxT, float* Q2T, float* Q3T, void* stream);
FIODE_API int fiode_rhs_bwd_data(const fiode_dyn_weights* w, const float* h, long long N, const float* g_f, const float* f,
                       const float* ftilde, const float* mu, const float* z1, const float* z2, const float* Q1T,
                       const float* QxT, const float* Q2T, const float* Q3T, float* dft, float* d2, float* d1,
                       float* g_h, float* g_s, void* ws, size_t ws_bytes, void* stream);
FIODE_API size_t fiode_rhs_bwd_params_workspace_bytes(long long M);
FIODE_API int fiode_rhs_bwd_params(const float* h, const float* s, const float* z1, const float* z2, const float* dft,
                         const float* d2, const float* d1, long long M, float* dQ1, float* dQx, float* dQ2,
                         float* dQ3, float* db1, float* db2, float* db3, void* ws, size_t ws_bytes, void* stream);
/* out = sum_i c[i] x[i], n_terms <= 8 (the Runge-Kutta combinations of the replay and of its adjoint). */
FIODE_API int fiode_lincomb(float* out, long long n, int n_terms, const float* const* x, const float* c, void* stream);

/* ---- interval step after the CROWN bound (robustness/certify_crown.py:139-150) ------------------
 * CROWN itself (auto_LiRPA) stays on the reference; lb <= ftilde <= ub (N,10) are its outputs over [eta-eps, eta+eps].
 * fiode_ibp_sigmoid   : CrownOrthoClassDynProjectSimplexLips.ibp_sigmoid (dynamics/classification.py:175-181)
 * fiode_ibp_cbf_qp    : ....ibp_cbf_qp(h, eps, lb, ub, upper=False) (:208-242): 10 N projection problems, one batch-global
 *                       stop iteration per call (lower / upper), coordinate i of problem (n, i) kept.  iters2 (2 ints,
 *                       optional) returns the iterations each call ran.
 * fiode_perturbed_vdot: certify_crown.py:29-34; label (1 or N ints), runner_up (N,10) uint8 mask; vmax optional.          */
FIODE_API size_t fiode_ibp_workspace_bytes(void);
FIODE_API int fiode_ibp_sigmoid(const float* lb, const float* ub, const float* h, long long N, float eps, float alpha_1,
                      float alpha_2, float sigma_1, float* out_lb, float* out_ub, void* stream);
FIODE_API int fiode_ibp_cbf_qp(const float* h, const float* lb, const float* ub, long long N, float eps, float alpha_1,
                     float sigma_1, int max_iter, float tol, float* f_lb, float* f_ub, int* iters2, void* ws,
                     size_t ws_bytes, void* stream);
/* certify_crown.py:146-148: mask[n][j] = (j != label) && eta[n][j] >= max_j eta[n][j] - margin.                          */
FIODE_API int fiode_runner_up_mask(const float* eta, long long N, int label, float margin, uint8_t* mask, void* stream);
FIODE_API int fiode_perturbed_vdot(const float* f_lb, const float* f_ub, const int* label, int label_rows,
                         const uint8_t* runner_up, long long N, float* vdot, float* vmax, void* ws, size_t ws_bytes,
                         void* stream);
/* Solves of B rows that one fiode_dopri5_simplex_batched launch keeps co-resident (one thread-block cluster per
 * solve; cudaOccupancyMaxActiveClusters).  More solves than this run in waves of a full solve latency each:
 * size G to a multiple of it.  cluster_size 0 = the launcher's default.  < 0 on error.                       */
FIODE_API int fiode_dopri5_max_concurrent_solves(int B, int cluster_size);
#ifdef __cplusplus
}
#endif
#endif /* FIODE_B200_H */
