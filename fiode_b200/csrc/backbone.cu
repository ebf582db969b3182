// fiode_b200/csrc/backbone.cu -- forward of the static-feature backbone s = param_map(x).
//
// Replaces models.py:17-35 (Normalize + make_ortho_KWLarge_Concat) and
// libs/ortho_conv/models_ortho.py:12-28 (KWLarge_Concat: 4 x [CayleyConv, GroupSort], Flatten,
// 2 x [CayleyLinear, GroupSort], fc) in eval mode, with every Cayley matrix prepared once per
// weight version (cayley.cu).  The two stride-2 layers are real 1x1 convolutions on the
// space-to-depth tensor and run without any FFT.
//
// The batch is processed in slabs so that the per-layer activations of a slab stay L2-resident
// (126 MB) between the producing and the consuming kernel.
#include "common.cuh"
#include "gemm_tc.cuh"
#include <cstdlib>

namespace fiode {
int launch_rfft2(const float* x, int n_img, int C, int n, const float* mu, const float* sd, float2* Xf, cudaStream_t st,
                 bool freq_major = false, bool x_u8 = false, int xrow = 0);
int launch_irfft2(const float2* Yf, int B, int C, int n, const float* bias, int gs, float* y, cudaStream_t st,
                  int freq_major = 0);
int launch_irfft32_mix3(const float2* Xf, const float2* Qf, float2* Qt, int B, int C, const float* bias, int gs, float* y,
                        cudaStream_t st);
size_t dft32_consts_bytes();
int dft32_tc2_xrow();
int launch_dft32_consts(void* out, cudaStream_t st);
int launch_irfft32_mix3_tc(const float2* Xf, const float2* Qt, const void* consts, int B, const float* bias, int gs, float* y,
                           cudaStream_t st);
int launch_irfft32_mix3_tc2(const float2* Xf, const float2* Qt, const void* consts, int B, const float* bias, int gs, float* y,
                            cudaStream_t st);
int launch_mix_real_weights(const float2* Qf, int F, int cout, int cin, float* W, cudaStream_t st);
int launch_transpose_q(const float2* Qf, int F, int CC, float2* Qt, cudaStream_t st);
int launch_mix(const float2* Xf, const float2* Qf, int B, int F, int cout, int cin, float2* Yf, cudaStream_t st);
int launch_conv1x1(const float* x, const float* Q, const float* bias, int B, int cin, int cout, int n, int gs, float* y,
                   cudaStream_t st);
int gemm_nt(const float* A, int lda, const float* Bm, int ldb, int M, int N, int K, const float* bias, float alpha,
            int group_sort, float* y, int ldy, float* ws, size_t ws_bytes, cudaStream_t st);
size_t gemm_nt_workspace(int M, int N, int K);

// images per slab: ~0.55 MB of activations per image, so nothing but the smallest slabs would be L2-resident, and those
// leave the SMs idle between short launches.  Measured on the 8192-image step (profiles/r01e_slab_sweep.txt): slab 512
// 4.38 ms, 1024 3.36, 2048 2.80, 4096 2.54, 8192 2.44 -> one slab of up to 8192 images (4.5 GB of workspace).
// FIODE_SLAB overrides.
static int slab_size() {
    static int v = 0;
    if (!v) { const char* e = getenv("FIODE_SLAB"); v = e ? atoi(e) : 8192; if (v < 1) v = 8192; }
    return v;
}
#define kSlab slab_size()

// ---- optional per-stage device timing (bench.py's live roofline numbers): CUDA events on the launching stream ----
constexpr int kStages = 11;
static bool g_time_stages = false;
static cudaEvent_t g_ev[kStages + 1];
static bool g_ev_ready = false;
static float g_stage_ms[kStages];
static void mark(int i, cudaStream_t st) {
    if (!g_time_stages) return;
    if (!g_ev_ready) { for (auto& e : g_ev) cudaEventCreate(&e); g_ev_ready = true; }
    cudaEventRecord(g_ev[i], st);
}

struct SlabLayout {
    size_t xf0, yf0, a0, a2, xf4, yf4, a4, a6, l9, l11, gws, packed, total, gws_bytes;
};
// Kernel-ready copies of the prepared weights (one block, filled by pack_weights): fp16x2-split MMA tiles of the four
// real matrices, model.4's mixing as 144 real (128 x 64) matrices (fp32 staging + tiles), model.0's transposed spectrum.
struct PackedLayout { size_t p2, p6, p9, p11, w4, p4, qt0, dft0, total; };
static size_t al(size_t x) { return (x + 255) & ~(size_t)255; }
static PackedLayout packed_layout() {
    PackedLayout L; size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o += al(bytes); return r; };
    L.p2 = take(gemm_tc_packed_bytes(32, 128)); L.p6 = take(gemm_tc_packed_bytes(64, 256));
    L.p9 = take(gemm_tc_packed_bytes(512, 4096)); L.p11 = take(gemm_tc_packed_bytes(512, 512));
    L.w4 = take((size_t)144 * 128 * 64 * 4); L.p4 = take((size_t)144 * gemm_tc_packed_bytes(128, 64));
    L.qt0 = take((size_t)544 * 32 * 3 * 8);
    L.dft0 = take(dft32_consts_bytes());         // fp16 hi / lo DFT matrices of the tensor-core inverse transform (dft_tc.cu)
    L.total = o;
    return L;
}
static SlabLayout slab_layout(int S) {
    SlabLayout L; size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o += al(bytes); return r; };
    L.xf0 = take((size_t)S * 3 * 32 * 20 * 8);          // rows of 17, or padded to 20 for the tcgen05 DFT kernel
    L.yf0 = take((size_t)S * 32 * 544 * 8);
    L.a0 = take((size_t)S * 32 * 1024 * 4);
    L.a2 = take((size_t)S * 32 * 256 * 4);
    L.xf4 = take((size_t)S * 32 * 144 * 8);
    L.yf4 = take((size_t)S * 64 * 144 * 8);
    L.a4 = take((size_t)S * 64 * 256 * 4);
    L.a6 = take((size_t)S * 4096 * 4);
    L.l9 = take((size_t)S * 512 * 4);
    L.l11 = take((size_t)S * 512 * 4);
    L.gws_bytes = gemm_nt_workspace(S, 512, 4096);
    const size_t g2 = gemm_nt_workspace(S, 512, 512), g3 = gemm_nt_workspace(S, 10, 512);
    if (g2 > L.gws_bytes) L.gws_bytes = g2;
    if (g3 > L.gws_bytes) L.gws_bytes = g3;
    L.gws = take(L.gws_bytes);
    L.packed = take(packed_layout().total);     // used when the caller passes no prepared block (w->packed == NULL)
    L.total = o;
    return L;
}
}  // namespace fiode

using namespace fiode;

static int pack_weights(const fiode_backbone_weights* w, char* pk, cudaStream_t st) {
    const PackedLayout P = packed_layout();
    int rc;
    if ((rc = launch_pack_w(w->Q2, 32, 128, gemm_tc_tile(32, 1), 1, pk + P.p2, st))) return rc;
    if ((rc = launch_pack_w(w->Q6, 64, 256, gemm_tc_tile(64, 1), 1, pk + P.p6, st))) return rc;
    if ((rc = launch_pack_w(w->Q9, 512, 4096, gemm_tc_tile(512, 1), 1, pk + P.p9, st))) return rc;
    if ((rc = launch_pack_w(w->Q11, 512, 512, gemm_tc_tile(512, 1), 1, pk + P.p11, st))) return rc;
    if ((rc = launch_mix_real_weights((const float2*)w->Qf4, 144, 64, 32, (float*)(pk + P.w4), st))) return rc;
    if ((rc = launch_pack_w((const float*)(pk + P.w4), 128, 64, 128, 0, pk + P.p4, st, 144))) return rc;
    if ((rc = launch_transpose_q((const float2*)w->Qf0, 544, 32 * 3, (float2*)(pk + P.qt0), st))) return rc;
    return launch_dft32_consts(pk + P.dft0, st);
}
// model.0's inverse transform: 2 (default) = two DFT GEMMs on tcgen05, warp-specialised (dft_tc.cu, irfft32_mix3_tc2_kernel);
// 1 = the first tcgen05 version (one serial pipeline per warpgroup); 0 = mixing + inverse FFT on the CUDA cores.
// FIODE_CONV0 = tc2 | tc | fft, or fiode_backbone_set_conv0_path.  Per 8192 images: 0.441 / 0.584 / 0.592 ms.
static int g_conv0_path = -1;
static int conv0_path() {
    if (g_conv0_path < 0) {
        const char* e = getenv("FIODE_CONV0");
        g_conv0_path = !e ? 2 : (e[0] == 'f' ? 0 : ((e[0] == 't' && e[1] == 'c' && e[2] == '\0') ? 1 : 2));
    }
    return g_conv0_path;
}

extern "C" int fiode_backbone_slab_images(void) { return kSlab; }
extern "C" void fiode_backbone_set_conv0_path(int path) { g_conv0_path = (path == 0 || path == 1) ? path : 2; }
extern "C" size_t fiode_backbone_packed_bytes(void) { return packed_layout().total; }
extern "C" int fiode_backbone_pack(const fiode_backbone_weights* w, void* packed, size_t packed_bytes, void* stream) {
    if (!w || !packed) { set_error("fiode_backbone_pack: null pointer"); return -1; }
    if (packed_bytes < packed_layout().total) { set_error("fiode_backbone_pack: block too small (%zu < %zu)", packed_bytes, packed_layout().total); return -1; }
    if (mlp_path() != 0) return 0;     // FFMA debug path (FIODE_MLP): reads the fp32 matrices directly
    return pack_weights(w, reinterpret_cast<char*>(packed), (cudaStream_t)stream);
}

extern "C" size_t fiode_backbone_workspace_bytes(int B) {
    const int S = B < kSlab ? B : kSlab;
    return slab_layout(S).total;
}

static int backbone_fwd(const fiode_backbone_weights* w, const float* x, bool x_u8, int B, float* s_out, void* ws,
                        size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!w || !x || !s_out || !ws) { set_error("fiode_backbone_fwd: null pointer"); return -1; }
    if (B <= 0) { set_error("fiode_backbone_fwd: B=%d", B); return -1; }
    const int S = B < kSlab ? B : kSlab;
    const SlabLayout L = slab_layout(S);
    if (ws_bytes < L.total) { set_error("fiode_backbone_fwd: workspace too small (%zu < %zu)", ws_bytes, L.total); return -1; }
    char* base = reinterpret_cast<char*>(ws);
    float2* xf0 = (float2*)(base + L.xf0); float2* yf0 = (float2*)(base + L.yf0);
    float* a0 = (float*)(base + L.a0); float* a2 = (float*)(base + L.a2);
    float2* xf4 = (float2*)(base + L.xf4); float2* yf4 = (float2*)(base + L.yf4);
    float* a4 = (float*)(base + L.a4); float* a6 = (float*)(base + L.a6);
    float* l9 = (float*)(base + L.l9); float* l11 = (float*)(base + L.l11);
    float* gws = (float*)(base + L.gws);
    int rc;
    const bool tcp = mlp_path() == 0;
    const PackedLayout P = packed_layout();
    char* pk = w->packed ? reinterpret_cast<char*>(const_cast<void*>(w->packed)) : base + L.packed;
    mark(0, st);
    if (tcp && !w->packed && (rc = pack_weights(w, pk, st))) return rc;     // per call unless the caller keeps a block
    mark(1, st);
    for (int b0 = 0; b0 < B; b0 += S) {
        const int nb = (B - b0) < S ? (B - b0) : S;
        const float* xs = x_u8 ? reinterpret_cast<const float*>(reinterpret_cast<const unsigned char*>(x) + (size_t)b0 * 3 * 1024)
                               : x + (size_t)b0 * 3 * 1024;
        // model.0: CayleyConv(3, 32, 3) @32x32 + GroupSort (ToTensor of 8-bit pixels and Normalize fused into the FFT load)
        if ((rc = launch_rfft2(xs, nb * 3, 3, 32, w->mu, w->std, xf0, st, false, x_u8, tcp && conv0_path() == 2 ? dft32_tc2_xrow() : 0))) return rc;
        mark(2, st);
        if (tcp) {   // mixing fused into the inverse FFT: the 32-channel spectrum never reaches HBM
            if (conv0_path() == 2) {
                if ((rc = launch_irfft32_mix3_tc2(xf0, (const float2*)(pk + P.qt0), pk + P.dft0, nb, w->b0, 1, a0, st))) return rc;
            } else if (conv0_path() == 1) {
                if ((rc = launch_irfft32_mix3_tc(xf0, (const float2*)(pk + P.qt0), pk + P.dft0, nb, w->b0, 1, a0, st))) return rc;
            } else if ((rc = launch_irfft32_mix3(xf0, nullptr, (float2*)(pk + P.qt0), nb, 32, w->b0, 1, a0, st))) return rc;
        } else {
            if ((rc = launch_mix(xf0, (const float2*)w->Qf0, nb, 544, 32, 3, yf0, st))) return rc;
            if ((rc = launch_irfft2(yf0, nb, 32, 32, w->b0, 1, a0, st))) return rc;
        }
        mark(3, st);
        // model.2: CayleyConv(32, 32, 3, stride=2) == orthogonal 1x1 on space-to-depth + GroupSort
        if ((rc = tcp ? launch_conv1x1_tc(a0, pk + P.p2, w->b2, nb, 32, 32, 32, 1, a2, st)
                      : launch_conv1x1(a0, w->Q2, w->b2, nb, 32, 32, 32, 1, a2, st))) return rc;
        mark(4, st);
        // model.4: CayleyConv(32, 64, 3) @16x16 + GroupSort
        if (tcp) {   // frequency-major spectra: the mixing of frequency phi is the row-major GEMM (nb x 64) . (128 x 64)^T
            if ((rc = launch_rfft2(a2, nb * 32, 32, 16, nullptr, nullptr, xf4, st, true))) return rc;
            mark(5, st);
            // mixed spectrum as [b][phi][co] (row stride 144 x 128 floats, one frequency = 128 floats further): scattered 512-byte
            // row writes for the GEMM, one contiguous 74 KB block per image for the inverse FFT.  FIODE_YF4=1: [phi][b][co].
            static const bool ifm = [] { const char* e = getenv("FIODE_YF4"); return !(e && atoi(e) == 1); }();
            if ((rc = launch_linear_tc_batched((const float*)xf4, 64, (long long)nb * 64, pk + P.p4, nb, 128, 64, 144,
                                               (float*)yf4, ifm ? 144 * 128 : 128, ifm ? 128 : (long long)nb * 128, st))) return rc;
            mark(6, st);
            if ((rc = launch_irfft2(yf4, nb, 64, 16, w->b4, 1, a4, st, ifm ? 2 : 1))) return rc;
        } else {
            if ((rc = launch_rfft2(a2, nb * 32, 32, 16, nullptr, nullptr, xf4, st))) return rc;
            if ((rc = launch_mix(xf4, (const float2*)w->Qf4, nb, 144, 64, 32, yf4, st))) return rc;
            if ((rc = launch_irfft2(yf4, nb, 64, 16, w->b4, 1, a4, st))) return rc;
        }
        mark(7, st);
        // model.6: CayleyConv(64, 64, 3, stride=2) + GroupSort -> (nb, 64, 8, 8) == Flatten 4096
        if ((rc = tcp ? launch_conv1x1_tc(a4, pk + P.p6, w->b6, nb, 64, 64, 16, 1, a6, st)
                      : launch_conv1x1(a4, w->Q6, w->b6, nb, 64, 64, 16, 1, a6, st))) return rc;
        mark(8, st);
        // model.9 / model.11 / fc
        if (tcp) {
            if ((rc = launch_linear_tc(a6, 4096, pk + P.p9, w->b9, nb, 512, 4096, 1, l9, 512, st))) return rc;
            mark(9, st);
            if ((rc = launch_linear_tc(l9, 512, pk + P.p11, w->b11, nb, 512, 512, 1, l11, 512, st))) return rc;
        } else {
            if ((rc = gemm_nt(a6, 4096, w->Q9, 4096, nb, 512, 4096, w->b9, 1.f, 1, l9, 512, gws, L.gws_bytes, st))) return rc;
            if ((rc = gemm_nt(l9, 512, w->Q11, 512, nb, 512, 512, w->b11, 1.f, 1, l11, 512, gws, L.gws_bytes, st))) return rc;
        }
        mark(10, st);
        if ((rc = gemm_nt(l11, 512, w->Qfc, 512, nb, 10, 512, w->bfc, 1.f, 0, s_out + (size_t)b0 * 10, 10, gws,
                          L.gws_bytes, st))) return rc;
        mark(11, st);
    }
    FIODE_CUDA(cudaGetLastError());
    return 0;
}

// model.0 alone (debug / tests): x (B,3,32,32) -> a0 (B,32,32,32) through conv0 path 0 | 1 | 2; ws as for fiode_backbone_fwd, B <= slab
extern "C" int fiode_debug_conv0(const fiode_backbone_weights* w, const float* x, int B, int path, float* a0_out, void* ws,
                                 size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const SlabLayout L = slab_layout(B);
    if (!w || !w->packed || ws_bytes < L.total || B > kSlab) { set_error("fiode_debug_conv0: bad arguments"); return -1; }
    char* base = reinterpret_cast<char*>(ws);
    float2* xf0 = (float2*)(base + L.xf0);
    const PackedLayout P = packed_layout();
    char* pk = reinterpret_cast<char*>(const_cast<void*>(w->packed));
    int rc;
    if ((rc = launch_rfft2(x, B * 3, 3, 32, w->mu, w->std, xf0, st, false, false, path == 2 ? dft32_tc2_xrow() : 0))) return rc;
    if (path == 2) return launch_irfft32_mix3_tc2(xf0, (const float2*)(pk + P.qt0), pk + P.dft0, B, w->b0, 1, a0_out, st);
    if (path == 1) return launch_irfft32_mix3_tc(xf0, (const float2*)(pk + P.qt0), pk + P.dft0, B, w->b0, 1, a0_out, st);
    return launch_irfft32_mix3(xf0, nullptr, (float2*)(pk + P.qt0), B, 32, w->b0, 1, a0_out, st);
}

extern "C" int fiode_backbone_fwd(const fiode_backbone_weights* w, const float* x, int B, float* s_out, void* ws,
                                  size_t ws_bytes, void* stream) {
    return backbone_fwd(w, x, false, B, s_out, ws, ws_bytes, stream);
}
extern "C" int fiode_backbone_fwd_u8(const fiode_backbone_weights* w, const unsigned char* x, int B, float* s_out, void* ws,
                                     size_t ws_bytes, void* stream) {
    return backbone_fwd(w, reinterpret_cast<const float*>(x), true, B, s_out, ws, ws_bytes, stream);
}

/* Stage timing for bench.py (tensor-core path, one slab): enable, run fiode_backbone_fwd once, then read.  Stages:
 * 0 weight packing | 1 rfft 32x32 (3 ch) | 2 mix + irfft 32x32 (32 ch) | 3 conv 128->32 @16 (tcgen05) | 4 rfft 16x16 (32 ch)
 * 5 mixing 144 x (64 -> 128) (tcgen05) | 6 irfft 16x16 (64 ch) | 7 conv 256->64 @8 (tcgen05) | 8 linear 4096->512 (tcgen05)
 * 9 linear 512->512 (tcgen05) | 10 fc 512->10.  Synchronises the device.                                            */
extern "C" void fiode_backbone_enable_stage_timing(int on) { fiode::g_time_stages = on != 0; }
extern "C" int fiode_backbone_stage_times(float* ms_out, int n) {
    if (!fiode::g_ev_ready) { set_error("fiode_backbone_stage_times: no timed run yet"); return -1; }
    if (cudaDeviceSynchronize() != cudaSuccess) { set_error("fiode_backbone_stage_times: sync failed"); return -1; }
    for (int i = 0; i < fiode::kStages && i < n; ++i)
        if (cudaEventElapsedTime(&ms_out[i], fiode::g_ev[i], fiode::g_ev[i + 1]) != cudaSuccess) {
            set_error("fiode_backbone_stage_times: events incomplete (tensor-core path, batch <= slab only)"); return -1; }
    return 0;
}
