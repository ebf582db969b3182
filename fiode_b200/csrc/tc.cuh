// fiode_b200/csrc/tc.cuh -- thin inline-PTX layer over the sm_100a tensor-core path: tcgen05.mma with the
// A operand in tensor memory (TMEM) and B in shared memory, TMEM alloc / ld / st, mbarriers.
//
// Number format used by every tensor-core kernel here ("fp16x2 split", an FP32 emulation):
//   x = hi + lo,  hi = fp16_rn(x),  lo = fp16_rn(x - hi)       (x - hi is exact in fp32)
//   A.B ~= A_hi.B_hi + A_lo.B_hi + A_hi.B_lo                   (three kind::f16 MMAs, fp32 accumulate)
// Dropped term lo.lo is O(2^-22) relative, the same order as one fp32 rounding.  Weights are pre-scaled
// by 2^6 (exact) so that their lo halves stay out of the fp16 subnormal range; the epilogue multiplies the
// accumulator by 2^-6 (exact).  |values| must stay below 65504 / 64: kernels flag non-finite results.
#pragma once
#include <cuda_fp16.h>
#include <cstdint>

namespace fiode {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- TMEM allocation (one warp, .sync.aligned) ---------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy smem writes -> visible to the async proxy (tensor core operand reads)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- mbarrier ----------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {   // may suspend for a bounded time
    uint32_t ok;
    asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.u32 %0, 1, 0, P1;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {  // non-blocking
    uint32_t ok;
    asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.u32 %0, 1, 0, P1;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) { while (!mbar_try_wait(bar, parity)) {} }

// ---- MMA ---------------------------------------------------------------------------------------------
// all previously issued MMAs of this thread complete -> one arrival on `bar`
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[tmem] . B[smem]^T ; M = 128 rows (TMEM lanes), fp16 operands, fp32 accumulate
__device__ __forceinline__ void mma_ts_f16(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_ss_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// Eight MMAs over consecutive 16-wide k slabs of one operand pair, as ONE asm block: the A address advances
// by 8 TMEM columns and the B descriptor by `b_step` (descriptor units of 16 bytes) per slab.  Written as a
// dependent add chain on purpose: left to the compiler, all descriptors are precomputed and spilled.
#define FIODE_MMA_STEP_ "add.s64 bd, bd, %4;\n\tadd.u32 ta, ta, 8;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [ta], bd, %3, q;\n\t"
#define FIODE_MMA_HEAD_ \
        "{\n\t.reg .pred p, q;\n\t.reg .b64 bd;\n\t.reg .b32 ta;\n\t" \
        "setp.ne.b32 p, %5, 0;\n\tsetp.eq.b32 q, 0, 0;\n\tmov.b64 bd, %2;\n\tmov.b32 ta, %1;\n\t" \
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [ta], bd, %3, p;\n\t"
__device__ __forceinline__ void mma_ts_f16_x8(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint64_t b_step, uint32_t idesc,
                                              uint32_t accumulate_first) {
    asm volatile(FIODE_MMA_HEAD_
        FIODE_MMA_STEP_ FIODE_MMA_STEP_ FIODE_MMA_STEP_ FIODE_MMA_STEP_ FIODE_MMA_STEP_ FIODE_MMA_STEP_ FIODE_MMA_STEP_
        "}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "l"(b_step), "r"(accumulate_first) : "memory");
}
// four consecutive k slabs (half of a 128-deep operand)
__device__ __forceinline__ void mma_ts_f16_x4(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint64_t b_step, uint32_t idesc,
                                              uint32_t accumulate_first) {
    asm volatile(FIODE_MMA_HEAD_
        FIODE_MMA_STEP_ FIODE_MMA_STEP_ FIODE_MMA_STEP_
        "}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "l"(b_step), "r"(accumulate_first) : "memory");
}
#undef FIODE_MMA_HEAD_
#undef FIODE_MMA_STEP_

// ---- issue from a CONVERGED warp: every lane executes the surrounding code, one elected lane issues.  With
// warp-uniform operands ptxas keeps addresses / descriptors in uniform registers (no per-MMA R2UR waterfall).
__device__ __forceinline__ void mma_ts_f16_elect(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p, e;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|e, 0xffffffff;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_commit_elect(uint64_t* bar) {
    asm volatile("{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
                 "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}

// Shared-memory matrix descriptor, K-major, no swizzle.  The operand is stored as 8-row x 16-byte core
// matrices (128 contiguous bytes each): lbo = byte distance between core matrices adjacent in K,
// sbo = between core matrices adjacent along M/N.  (Verified on B200 by tools/probe/tc_probe.cu.)
__host__ __device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo >> 4) & 0x3fffu) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3fffu) << 32) | ((uint64_t)1 << 46);
}
// instruction descriptor: fp16 x fp16 -> fp32, both operands K-major, M x N tile
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N) {
    return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---- TMEM <-> registers: shape 32x32b -- the warp's 32 lanes <-> TMEM lanes 32*(warp%4) + lane,
//      register i <-> column base + i -------------------------------------------------------------------
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void tmem_st8(uint32_t addr, const uint32_t* v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"r"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
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
__device__ __forceinline__ void tmem_ld32(uint32_t addr, uint32_t* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(addr) : "memory");
}

// ---- fp16x2 split of a pair of fp32 values: hi/lo packed half2 words (element 0 in the low half) ------
__device__ __forceinline__ void split2(float x0, float x1, uint32_t& hi, uint32_t& lo) {
    const __half2 h = __floats2half2_rn(x0, x1);
    const float2 f = __half22float2(h);
    const __half2 l = __floats2half2_rn(x0 - f.x, x1 - f.y);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}
__device__ __forceinline__ void split1(float x, unsigned short& hi, unsigned short& lo) {
    const __half h = __float2half_rn(x);
    const __half l = __float2half_rn(x - __half2float(h));
    hi = __half_as_ushort(h);
    lo = __half_as_ushort(l);
}

}  // namespace tc
}  // namespace fiode
