// tcgen05 / tensor-memory / mbarrier helpers shared by the tensor-core kernels (gn.cu: gn_tc_kernel, fbp.cu: fbp_filter_tc_kernel).
// sm_100a only.  All of them are thin wrappers over one PTX instruction (plus the elect / predicate set-up it needs).
#pragma once
#include <stdint.h>

namespace dexct {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Wait for the phase of the given parity to complete.  A wait that lasts seconds can only mean a protocol error: trap
// instead of hanging the device.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok = 0, spins = 0;
    for (;;) {
        // try_wait suspends the warp until the phase completes or a hardware time limit passes: no issue slots burnt
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) return;
        if (++spins > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&v)[8])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(v[0]), "r"(v[1]),
                 "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
                 "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
                 "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16])
{
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int q = 0; q < 16; ++q) v[q] = __uint_as_float(r[q]);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32])
{
    uint32_t r[32];
#pragma unroll
    for (int h = 0; h < 2; ++h)
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                     : "=r"(r[16 * h + 0]), "=r"(r[16 * h + 1]), "=r"(r[16 * h + 2]), "=r"(r[16 * h + 3]), "=r"(r[16 * h + 4]),
                       "=r"(r[16 * h + 5]), "=r"(r[16 * h + 6]), "=r"(r[16 * h + 7]), "=r"(r[16 * h + 8]), "=r"(r[16 * h + 9]),
                       "=r"(r[16 * h + 10]), "=r"(r[16 * h + 11]), "=r"(r[16 * h + 12]), "=r"(r[16 * h + 13]), "=r"(r[16 * h + 14]),
                       "=r"(r[16 * h + 15])
                     : "r"(taddr + 16 * h) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int q = 0; q < 32; ++q) v[q] = __uint_as_float(r[q]);
}
// D[tmem] (+)= A[tmem] x B[smem descriptor], M128 K8, TF32 in, FP32 accumulate
// Called by ALL lanes of a converged warp with warp-uniform arguments; one elected lane issues.  (Issued from inside an
// `if (lane == 0)` region the operands reach the uniform registers the instruction needs through a per-lane waterfall --
// ELECT / R2UR / BRA.U.ANY around every MMA -- and the issuing warp, not the tensor pipe, bounds the kernel: profiles/r02j.)
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{ .reg .pred p, q; elect.sync _|q, 0xffffffff; setp.ne.b32 p, %4, 0;\n"
                 "  @q tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p; }" ::"r"(d_tmem),
                 "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar)
{
    asm volatile("{ .reg .pred q; elect.sync _|q, 0xffffffff;\n"
                 "  @q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0]; }" ::"r"(bar) : "memory");
}

}  // namespace dexct
