// tools/probe/tc_probe.cu -- bring-up probe for tcgen05 (A in TMEM fp16, B in smem K-major no-swizzle).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tc_probe tc_probe.cu ; run on a B200.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cstdint>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols));
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.u32 %0, 1, 0, P1;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) { while (!mbar_try(bar, parity)) {} }
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void mma_ts_f16(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_ss_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t addr, const uint32_t* v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                 "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t addr, uint32_t* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(addr) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// K-major no-swizzle smem descriptor: core matrix = 8 rows x 16 B (128 B contiguous);
// LBO = byte distance between core matrices adjacent in K, SBO = between core matrices adjacent in M/N.
__host__ __device__ inline uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;   // version = 1 (Blackwell)
    return d;                 // layout_type = 0 (no swizzle), base_offset 0
}
__host__ __device__ inline uint32_t make_idesc_f16(int M, int N) {
    return (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// B smem layout: [kchunk (16 B = 8 halves)][n][8 halves]  => LBO = N*16, SBO = 128
template <int N, int K>
__global__ void __launch_bounds__(128, 1) probe_kernel(const __half* A, const __half* B, float* D, int reps, long long* cycles, int swapAB) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(8) uint64_t bar;
    const int t = threadIdx.x, warp = t >> 5;
    __half* Bs = reinterpret_cast<__half*>(smem);
    if (warp == 0) tmem_alloc(&tmem_base_s, 512);
    if (t == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    for (int e = t; e < N * K / 8; e += 128) {
        const int kc = e / N, n = e % N;
        *reinterpret_cast<uint4*>(Bs + ((size_t)kc * N + n) * 8) = *reinterpret_cast<const uint4*>(B + (size_t)n * K + kc * 8);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tmem = tmem_base_s;
    const uint32_t d_col = 0, a_col = 256;
    // A row t -> TMEM columns a_col .. a_col + K/2
    const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
    for (int c = 0; c < K / 2; c += 16) {
        uint32_t v[16];
        for (int i = 0; i < 16; ++i) v[i] = *reinterpret_cast<const uint32_t*>(A + (size_t)t * K + (c + i) * 2);
        tmem_st16(lane_addr + a_col + c, v);
    }
    tmem_wait_st();
    fence_before();
    __syncthreads();
    long long t0 = 0, t1 = 0;
    const uint32_t idesc = make_idesc_f16(128, N);
    if (t == 0) {
        fence_after();
        t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            for (int ks = 0; ks < K / 16; ++ks) {
                const uint64_t bd = make_desc(smem_u32(Bs) + ks * 2 * N * 16, N * 16, 128);
                mma_ts_f16(tmem + d_col, tmem + a_col + ks * 8, bd, idesc, ks > 0 ? 1u : 0u);
            }
        }
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    if (t == 0) { t1 = clock64(); cycles[0] = t1 - t0; }
    fence_after();
    for (int c = 0; c < N; c += 16) {
        uint32_t v[16];
        tmem_ld16(lane_addr + d_col + c, v);
        tmem_wait_ld();
        for (int i = 0; i < 16; ++i) D[(size_t)t * N + c + i] = __uint_as_float(v[i]);
    }
    // TMEM ld/st throughput: all 4 warps, 128 columns each way, many reps
    __syncthreads();
    {
        uint32_t v[16];
        for (int i = 0; i < 16; ++i) v[i] = t + i;
        long long s0 = clock64();
        for (int r = 0; r < 64; ++r)
            for (int c = 0; c < 128; c += 16) tmem_st16(lane_addr + a_col + c, v);
        tmem_wait_st();
        __syncthreads();
        long long s1 = clock64();
        uint32_t acc = 0;
        for (int r = 0; r < 64; ++r)
            for (int c = 0; c < 128; c += 16) { tmem_ld16(lane_addr + a_col + c, v); tmem_wait_ld(); acc += v[0] + v[15]; }
        __syncthreads();
        long long s2 = clock64();
        if (t == 0) { cycles[1] = s1 - s0; cycles[2] = s2 - s1; cycles[3] = acc; }
    }
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int N, int K>
int run(int reps) {
    const int M = 128;
    std::vector<__half> A(M * K), B(N * K);
    std::vector<float> Af(M * K), Bf(N * K);
    srand(1);
    for (int i = 0; i < M * K; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; A[i] = __float2half(v); Af[i] = __half2float(A[i]); }
    for (int i = 0; i < N * K; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; B[i] = __float2half(v); Bf[i] = __half2float(B[i]); }
    __half *dA, *dB; float* dD; long long* dC;
    CK(cudaMalloc(&dA, A.size() * 2)); CK(cudaMalloc(&dB, B.size() * 2)); CK(cudaMalloc(&dD, M * N * 4)); CK(cudaMalloc(&dC, 64));
    CK(cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemset(dD, 0, M * N * 4));
    const size_t smem = (size_t)N * K * 2 + 1024;
    CK(cudaFuncSetAttribute(probe_kernel<N, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    probe_kernel<N, K><<<1, 128, smem>>>(dA, dB, dD, reps, dC, 0);
    CK(cudaDeviceSynchronize());
    std::vector<float> D(M * N); long long cyc[4];
    CK(cudaMemcpy(D.data(), dD, M * N * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(cyc, dC, 32, cudaMemcpyDeviceToHost));
    double maxerr = 0; int bad = 0;
    for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) {
        double ref = 0; for (int k = 0; k < K; ++k) ref += (double)Af[m * K + k] * Bf[n * K + k];
        double e = fabs(ref - D[m * N + n]); if (e > maxerr) maxerr = e; if (e > 1e-3) ++bad;
    }
    printf("N=%d K=%d reps=%d: max err %.3e bad %d ; mma cycles %lld (%.1f per instr) ; st 64x128 cols: %lld cyc ; ld 64x128 cols: %lld cyc\n",
           N, K, reps, maxerr, bad, cyc[0], (double)cyc[0] / (reps * (K / 16)), cyc[1], cyc[2]);
    if (bad) printf("  D[0][0..3] = %f %f %f %f ; D[1][0]=%f D[32][0]=%f\n", D[0], D[1], D[2], D[3], D[N], D[32 * N]);
    return bad;
}

int main() {
    int bad = 0;
    bad += run<128, 128>(1);
    bad += run<128, 128>(64);
    bad += run<16, 128>(1);
    bad += run<16, 128>(64);
    bad += run<128, 16>(1);
    bad += run<128, 16>(64);
    printf(bad ? "PROBE FAILED\n" : "PROBE OK\n");
    return bad ? 1 : 0;
}
