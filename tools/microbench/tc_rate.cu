// Microbenchmark: cost of one tcgen05.mma (M128, K = 32 bytes) on one SM of sm_100a as a function of N, of the operand
// kind (tf32 / bf16) and of where A comes from (tensor memory or shared memory).  One CTA of 128 threads; thread 0 issues
// REP back-to-back MMAs on the same operands, commits, and everybody waits on the mbarrier; clock64 around it.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tc_rate tc_rate.cu ; run: ./tc_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <bool TF32, bool A_TMEM>
__global__ void k(long long *cyc, int N, int rep, int nacc, int nissue, int wstride)
{
    extern __shared__ __align__(128) unsigned char smem[];     // zeros: A [128 x 32 B] then B [256 x 32 B], core-matrix layout
    __shared__ __align__(8) unsigned long long bar;
    __shared__ uint32_t tmem_sm;
    for (int i = threadIdx.x; i < (128 * 32 + 256 * 32) / 4; i += blockDim.x) ((uint32_t *)smem)[i] = 0;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(nissue) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_sm)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_sm;
    // zero the A columns in tensor memory (columns 256..263)
    {
        const uint32_t a = tmem + ((uint32_t)(((threadIdx.x >> 5) & 3) * 32) << 16) + 256;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(a), "r"(0u) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t idesc = (1u << 4) | ((TF32 ? 2u : 1u) << 7) | ((TF32 ? 2u : 1u) << 10) | ((uint32_t)(N >> 3) << 17) | (8u << 24);
    // K-major, no swizzle: core matrix 8 rows x 16 B; LBO = next 16 B along K, SBO = next 8 rows
    const uint64_t dA = ((uint64_t)((smem_u32(smem) & 0x3ffff) >> 4)) | ((uint64_t)(128 * 16 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
    const uint64_t dB = ((uint64_t)((smem_u32(smem + 128 * 32) & 0x3ffff) >> 4)) | ((uint64_t)(256 * 16 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
    long long t0 = clock64();
    if ((threadIdx.x & 31) == 0 && (int)(threadIdx.x >> 5) % wstride == 0 && (int)(threadIdx.x >> 5) / wstride < nissue) {       // nissue warps issue at the same time (warps 0, wstride, 2 wstride, ...: wstride = 4 puts them on one scheduler)
        for (int r = 0; r < rep; ++r) {
            const uint32_t dcol = tmem + (uint32_t)((r + (threadIdx.x >> 5)) & (nacc - 1)) * 64;    // nacc (a power of two) > 1: successive MMAs feed different accumulators; no integer division in the issue loop (r02d measured the division, not the MMA)
            if (A_TMEM) {
                if (TF32) asm volatile("{ .reg .pred p; setp.ne.b32 p, %4, 0; tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p; }" ::"r"(dcol), "r"(tmem + 256), "l"(dB), "r"(idesc), "r"(1u) : "memory");
                else asm volatile("{ .reg .pred p; setp.ne.b32 p, %4, 0; tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p; }" ::"r"(dcol), "r"(tmem + 256), "l"(dB), "r"(idesc), "r"(1u) : "memory");
            } else {
                if (TF32) asm volatile("{ .reg .pred p; setp.ne.b32 p, %4, 0; tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p; }" ::"r"(dcol), "l"(dA), "l"(dB), "r"(idesc), "r"(1u) : "memory");
                else asm volatile("{ .reg .pred p; setp.ne.b32 p, %4, 0; tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p; }" ::"r"(dcol), "l"(dA), "l"(dB), "r"(idesc), "r"(1u) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    long long t_issue = clock64();
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(smem_u32(&bar)) : "memory");
    long long t1 = clock64();
    if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t_issue - t0; }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

template <bool TF32, bool A_TMEM>
static void run(const char *name, int nacc, int nissue = 1, int wstride = 1)
{
    long long *d, h[2];
    cudaMalloc(&d, 16);
    const int smem = 128 * 32 + 256 * 32;
    for (int N : {16, 32, 64, 128, 256}) {
        if (nacc > 1 && N > 64) continue;
        long long c[2][2];
        int reps[2] = {64, 576};
        for (int q = 0; q < 2; ++q) {
            k<TF32, A_TMEM><<<1, wstride == 1 ? 128 : 32 * wstride * nissue, smem>>>(d, N, reps[q], nacc, nissue, wstride);
            k<TF32, A_TMEM><<<1, wstride == 1 ? 128 : 32 * wstride * nissue, smem>>>(d, N, reps[q], nacc, nissue, wstride);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("%s N=%d: %s\n", name, N, cudaGetErrorString(e)); return; }
            cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
            c[q][0] = h[0]; c[q][1] = h[1];
        }
        printf("%-22s %d accumulator(s) %d issuer(s) warp stride %d N=%3d: %7.1f cycles per MMA (SM aggregate), %6.1f per issue of one thread; %lld cycles for 64\n", name, nacc, nissue, wstride, N,
               (double)(c[1][0] - c[0][0]) / (512.0 * nissue), (double)(c[1][1] - c[0][1]) / 512.0, c[0][0]);
    }
    cudaFree(d);
}

int main()
{
    for (int nacc : {1, 2, 4}) {
        run<true, true>("tf32 K=8  A in TMEM", nacc);
        run<false, true>("bf16 K=16 A in TMEM", nacc);
    }
    for (int ni : {2, 4}) { run<true, true>("tf32 K=8  A in TMEM", 4, ni); run<true, true>("tf32 K=8  A in TMEM", 1, ni); }
    for (int ni : {2, 4}) run<true, true>("tf32 K=8  A in TMEM", 1, ni, 4);
    run<true, false>("tf32 K=8  A in smem", 1);
    run<false, false>("bf16 K=16 A in smem", 1);
    return 0;
}
