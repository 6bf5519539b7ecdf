// Microbenchmark: issue rate of FFMA / FFMA2 / MUFU.EX2 on one SM (sm_100a), GN-kernel-like operand pattern.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fma_pipe fma_pipe.cu ; run: ./fma_pipe
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
#define REP 256

__constant__ float ctab[64 * 16];

template <int MODE>
__global__ void k(float *out, long long *cyc, float seed)
{
    __shared__ __align__(16) float smem[64 * 16];
    for (int i = threadIdx.x; i < 64 * 16; i += blockDim.x) smem[i] = 1e-3f * i;
    // 12 "table" values, 1 broadcast multiplier, 12 (or 24) accumulators
    float t[12], acc[24];
    for (int i = 0; i < 12; ++i) t[i] = seed + i + threadIdx.x * 1e-3f;
    for (int i = 0; i < 24; ++i) acc[i] = 0.f;
    float att = seed * 0.5f, att2 = seed * 0.25f;
    u64 tp[6], ap[12];
    for (int i = 0; i < 6; ++i) asm("mov.b64 %0, {%1,%2};" : "=l"(tp[i]) : "f"(t[2 * i]), "f"(t[2 * i + 1]));
    for (int i = 0; i < 12; ++i) ap[i] = 0ull;
    __syncthreads();
    long long c0 = clock64();
#pragma unroll 1
    for (int r = 0; r < REP; ++r) {
        if (MODE == 0) {            // 24 scalar FFMA: acc[i] = t[i%12] * att + acc[i]
#pragma unroll
            for (int i = 0; i < 12; ++i) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc[i]) : "f"(t[i]), "f"(att));
#pragma unroll
            for (int i = 0; i < 12; ++i) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc[12 + i]) : "f"(t[i]), "f"(att2));
        } else if (MODE == 1) {     // 12 FFMA2 with a broadcast scalar (pack of the same value)
            u64 a1, a2;
            asm("mov.b64 %0, {%1,%1};" : "=l"(a1) : "f"(att));
            asm("mov.b64 %0, {%1,%1};" : "=l"(a2) : "f"(att2));
#pragma unroll
            for (int i = 0; i < 6; ++i) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(ap[i]) : "l"(tp[i]), "l"(a1));
#pragma unroll
            for (int i = 0; i < 6; ++i) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(ap[6 + i]) : "l"(tp[i]), "l"(a2));
        } else if (MODE == 2) {     // 24 scalar FFMA, two-register form: acc = acc * att + acc (few distinct registers)
#pragma unroll
            for (int i = 0; i < 24; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %0;" : "+f"(acc[i]) : "f"(att));
        } else if (MODE == 3) {     // 8 MUFU.EX2
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(acc[i]));
        } else if (MODE == 4) {     // 12 FFMA2, fully packed operands (no broadcast)
#pragma unroll
            for (int i = 0; i < 6; ++i) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(ap[i]) : "l"(tp[i]), "l"(tp[(i + 1) % 6]));
#pragma unroll
            for (int i = 0; i < 6; ++i) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(ap[6 + i]) : "l"(tp[i]), "l"(tp[(i + 2) % 6]));
        } else if (MODE == 5) {     // GN mix: 4 scalar + 12 FFMA2 + 2 MUFU
            float x0 = fmaf(att, t[0], att2 * t[1]), x1 = fmaf(att2, t[2], att * t[3]);
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x0));
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x1));
            u64 a1, a2;
            asm("mov.b64 %0, {%1,%1};" : "=l"(a1) : "f"(x0));
            asm("mov.b64 %0, {%1,%1};" : "=l"(a2) : "f"(x1));
#pragma unroll
            for (int i = 0; i < 6; ++i) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(ap[i]) : "l"(tp[i]), "l"(a1));
#pragma unroll
            for (int i = 0; i < 6; ++i) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(ap[6 + i]) : "l"(tp[i]), "l"(a2));
        } else if (MODE == 7 || MODE == 8) {     // GN mix packed with the table row read from shared (7) / constant (8) memory
            const ulonglong2 *row = MODE == 7 ? (const ulonglong2 *)(smem + (r & 63) * 16) : (const ulonglong2 *)(ctab + (r & 63) * 16);
            const ulonglong2 r0 = row[0], r1 = row[1], r2 = row[2], r3 = row[3];
            float t0, t1;
            asm("mov.b64 {%0,%1}, %2;" : "=f"(t0), "=f"(t1) : "l"(r0.x));
            float x0 = fmaf(att, t0, att2 * t1), x1 = fmaf(att2, t0, att * t1);
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x0));
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x1));
            u64 a1, a2;
            asm("mov.b64 %0, {%1,%1};" : "=l"(a1) : "f"(x0));
            asm("mov.b64 %0, {%1,%1};" : "=l"(a2) : "f"(x1));
            const u64 tq[6] = {r0.y, r1.x, r1.y, r2.x, r2.y, r3.x};
#pragma unroll
            for (int i = 0; i < 6; ++i) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(ap[i]) : "l"(tq[i]), "l"(a1));
#pragma unroll
            for (int i = 0; i < 6; ++i) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(ap[6 + i]) : "l"(tq[i]), "l"(a2));
        } else if (MODE == 6) {     // GN mix scalar: 4 + 24 FFMA + 2 MUFU
            float x0 = fmaf(att, t[0], att2 * t[1]), x1 = fmaf(att2, t[2], att * t[3]);
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x0));
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x1));
#pragma unroll
            for (int i = 0; i < 12; ++i) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc[i]) : "f"(t[i]), "f"(x0));
#pragma unroll
            for (int i = 0; i < 12; ++i) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc[12 + i]) : "f"(t[i]), "f"(x1));
        }
    }
    long long c1 = clock64();
    float s = 0.f;
    for (int i = 0; i < 24; ++i) s += acc[i];
    for (int i = 0; i < 12; ++i) { float x, y; asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(ap[i])); s += x + y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if ((threadIdx.x & 31) == 0) { cyc[2 * (threadIdx.x >> 5)] = c0; cyc[2 * (threadIdx.x >> 5) + 1] = c1; }
}

template <int MODE> void run(const char *name, int n_instr)
{
    float *out; long long *cyc;
    cudaMalloc(&out, 4 * 1024 * 148 * 4); cudaMalloc(&cyc, 8 * 148 * 4);
    for (int warps = 4; warps <= 32; warps *= 2) {                 // warps per SM (one CTA on one SM)
        k<MODE><<<1, warps * 32>>>(out, cyc, 1.0f);
        k<MODE><<<1, warps * 32>>>(out, cyc, 1.0f);
        long long cc[64]; cudaMemcpy(cc, cyc, 8 * 2 * warps, cudaMemcpyDeviceToHost);
        long long lo = cc[0], hi = cc[1];
        for (int w = 0; w < warps; ++w) { if (cc[2 * w] < lo) lo = cc[2 * w]; if (cc[2 * w + 1] > hi) hi = cc[2 * w + 1]; }
        long long c = hi - lo;
        double per_smsp = (double)c / ((double)REP * n_instr * (warps / 4.0));
        printf("%-34s warps/SM %2d  cycles %8lld  cycles per warp-instr per SMSP %.3f\n", name, warps, c, per_smsp);
    }
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    run<0>("FFMA 3-reg (t*att+acc) x24", 24);
    run<2>("FFMA acc*att+acc x24", 24);
    run<1>("FFMA2 bcast x12", 12);
    run<4>("FFMA2 packed x12", 12);
    run<3>("MUFU.EX2 x8", 8);
    run<5>("GN mix packed (4+12 FFMA2+2 MUFU)", 18);
    run<6>("GN mix scalar (4+24 FFMA+2 MUFU)", 30);
    run<7>("GN mix packed + 4 LDS.128", 22);
    run<8>("GN mix packed + const table", 22);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return 0;
}
