// Microbenchmark: the GN inner loop (per energy row: PX exponentials + 6 PX FFMA2 on a 16-float table row) with the row
// fetched from shared memory (LDS.128 broadcast), the constant bank (uniform loads), or half from each.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gn_mix gn_mix.cu ; run: ./gn_mix
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
#define REP 512

__constant__ ulonglong2 ctab[64 * 4];

__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c)
{
    u64 r;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

// SRC: 0 = registers (no loads), 1 = shared, 2 = constant, 3 = rows 0,1 shared + rows 2,3 constant, 4 = 0 const + 1..3 shared,
//      5 = constant at a loop-invariant address (uniform-register operands, no loads in the loop),
//      6 = constant through per-thread LDC (vector registers)
template <int PX, int SRC>
__global__ void k(float *out, long long *cyc, float seed)
{
    __shared__ __align__(16) ulonglong2 smem[64 * 4];
    for (int i = threadIdx.x; i < 64 * 4; i += blockDim.x) {
        float a = 1e-3f * i, b = 2e-3f * i;
        u64 p;
        asm("mov.b64 %0, {%1,%2};" : "=l"(p) : "f"(a), "f"(b));
        smem[i] = make_ulonglong2(p, p);
    }
    float a1[PX], a2[PX];
    u64 acc[PX][6];
    for (int p = 0; p < PX; ++p) {
        a1[p] = seed * (p + 1) * 1e-3f + threadIdx.x * 1e-6f;
        a2[p] = seed * (p + 2) * 1e-3f;
        for (int q = 0; q < 6; ++q) acc[p][q] = 0ull;
    }
    ulonglong2 rr[4];
    for (int q = 0; q < 4; ++q) rr[q] = smem[q];
    __syncthreads();
    long long c0 = clock64();
#pragma unroll 2
    for (int r = 0; r < REP; ++r) {
        const int e = SRC == 5 ? (int)seed : SRC == 6 ? ((r & 63) + (threadIdx.x >> 20)) : (r & 63);
        ulonglong2 row[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const bool from_s = SRC == 1 || (SRC == 3 && q < 2) || (SRC == 4 && q > 0);
            const bool from_c = (SRC >= 10 && q < SRC - 10) || SRC == 2 || SRC == 5 || SRC == 6 || (SRC == 3 && q >= 2) || (SRC == 4 && q == 0);
            row[q] = from_s ? smem[e * 4 + q] : from_c ? ctab[e * 4 + q] : rr[q];
        }
        float t0, t1;
        asm("mov.b64 {%0,%1}, %2;" : "=f"(t0), "=f"(t1) : "l"(row[0].x));
#pragma unroll
        for (int p = 0; p < PX; ++p) {
            float x = fmaf(a2[p], t1, a1[p] * t0);
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x));
            u64 x2;
            asm("mov.b64 %0, {%1,%1};" : "=l"(x2) : "f"(x));
            acc[p][0] = ffma2(row[0].y, x2, acc[p][0]);
            acc[p][1] = ffma2(row[1].x, x2, acc[p][1]);
            acc[p][2] = ffma2(row[1].y, x2, acc[p][2]);
            acc[p][3] = ffma2(row[2].x, x2, acc[p][3]);
            acc[p][4] = ffma2(row[2].y, x2, acc[p][4]);
            acc[p][5] = ffma2(row[3].x, x2, acc[p][5]);
        }
    }
    long long c1 = clock64();
    float s = 0.f;
    for (int p = 0; p < PX; ++p)
        for (int q = 0; q < 6; ++q) { float x, y; asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(acc[p][q])); s += x + y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if ((threadIdx.x & 31) == 0) { cyc[2 * (threadIdx.x >> 5)] = c0; cyc[2 * (threadIdx.x >> 5) + 1] = c1; }
}

// Software-pipelined variant: the exponentials of row r+1 (exponent pair from its own small constant table) are issued
// before the FFMA2 of row r, whose moment pairs come from the constant bank at the time of use.
__constant__ u64 cexp[64];
template <int PX, int SRC>
__global__ void kp(float *out, long long *cyc, float seed)
{
    float a1[PX], a2[PX], att[PX];
    u64 acc[PX][6];
    for (int p = 0; p < PX; ++p) {
        a1[p] = seed * (p + 1) * 1e-3f + threadIdx.x * 1e-6f;
        a2[p] = seed * (p + 2) * 1e-3f;
        att[p] = seed;
        for (int q = 0; q < 6; ++q) acc[p][q] = 0ull;
    }
    long long c0 = clock64();
#pragma unroll 2
    for (int r = 0; r < REP; ++r) {
        const int e = r & 63, en = (r + 1) & 63;
        float t0, t1;
        asm("mov.b64 {%0,%1}, %2;" : "=f"(t0), "=f"(t1) : "l"(cexp[en]));
        float nxt[PX];
#pragma unroll
        for (int p = 0; p < PX; ++p) {
            nxt[p] = fmaf(a2[p], t1, a1[p] * t0);
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(nxt[p]));
        }
        const ulonglong2 r0 = ctab[e * 4], r1 = ctab[e * 4 + 1], r2 = ctab[e * 4 + 2];
#pragma unroll
        for (int p = 0; p < PX; ++p) {
            u64 x2;
            asm("mov.b64 %0, {%1,%1};" : "=l"(x2) : "f"(att[p]));
            acc[p][0] = ffma2(r0.x, x2, acc[p][0]);
            acc[p][1] = ffma2(r0.y, x2, acc[p][1]);
            acc[p][2] = ffma2(r1.x, x2, acc[p][2]);
            acc[p][3] = ffma2(r1.y, x2, acc[p][3]);
            acc[p][4] = ffma2(r2.x, x2, acc[p][4]);
            acc[p][5] = ffma2(r2.y, x2, acc[p][5]);
        }
#pragma unroll
        for (int p = 0; p < PX; ++p) att[p] = nxt[p];
    }
    long long c1 = clock64();
    float s = 0.f;
    for (int p = 0; p < PX; ++p)
        for (int q = 0; q < 6; ++q) { float x, y; asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(acc[p][q])); s += x + y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if ((threadIdx.x & 31) == 0) { cyc[2 * (threadIdx.x >> 5)] = c0; cyc[2 * (threadIdx.x >> 5) + 1] = c1; }
}

template <int PX, int SRC> void run(const char *name)
{
    float *out; long long *cyc;
    cudaMalloc(&out, 4 * 1024 * 4); cudaMalloc(&cyc, 8 * 64 * 2);
    for (int warps = 8; warps <= 32; warps *= 2) {
        if (warps * 32 * (PX * 16 + 24) > 65536) continue;           // register file
        if (SRC == 20) { kp<PX, SRC><<<1, warps * 32>>>(out, cyc, 1.0f); kp<PX, SRC><<<1, warps * 32>>>(out, cyc, 1.0f); }
        else { k<PX, SRC><<<1, warps * 32>>>(out, cyc, 1.0f); k<PX, SRC><<<1, warps * 32>>>(out, cyc, 1.0f); }
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s: launch failed\n", name); cudaGetLastError(); continue; }
        long long cc[128]; cudaMemcpy(cc, cyc, 8 * 2 * warps, cudaMemcpyDeviceToHost);
        long long lo = cc[0], hi = cc[1];
        for (int w = 0; w < warps; ++w) { if (cc[2 * w] < lo) lo = cc[2 * w]; if (cc[2 * w + 1] > hi) hi = cc[2 * w + 1]; }
        const double per_unit = (double)(hi - lo) / ((double)REP * PX * (warps / 4.0));
        printf("%-40s warps/SM %2d  cycles per (pixel, energy) unit per SMSP %.2f   (FMA floor 14)\n", name, warps, per_unit);
    }
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    run<2, 0>("PX=2 row in registers");
    run<2, 1>("PX=2 row from shared (4 LDS.128)");
    run<2, 2>("PX=2 row from constant");
    run<2, 3>("PX=2 rows 0,1 shared + 2,3 constant");
    run<2, 4>("PX=2 row 0 constant + 1..3 shared");
    run<2, 5>("PX=2 row in uniform registers");
    run<2, 6>("PX=2 row from constant via LDC");
    run<4, 5>("PX=4 row in uniform registers");
    run<4, 6>("PX=4 row from constant via LDC");
    run<2, 11>("PX=2 1 of 4 quarter-rows from constant");
    run<2, 12>("PX=2 2 of 4 quarter-rows from constant");
    run<2, 13>("PX=2 3 of 4 quarter-rows from constant");
    run<4, 11>("PX=4 1 of 4 quarter-rows from constant");
    run<4, 12>("PX=4 2 of 4 quarter-rows from constant");
    run<4, 13>("PX=4 3 of 4 quarter-rows from constant");
    run<1, 0>("PX=1 row in registers");
    run<1, 2>("PX=1 row from constant");
    run<1, 20>("PX=1 pipelined, constant");
    run<2, 20>("PX=2 pipelined, constant");
    run<3, 20>("PX=3 pipelined, constant");
    run<4, 20>("PX=4 pipelined, constant");
    run<3, 2>("PX=3 row from constant");
    run<3, 3>("PX=3 rows 0,1 shared + 2,3 constant");
    run<4, 0>("PX=4 row in registers");
    run<4, 1>("PX=4 row from shared");
    run<4, 2>("PX=4 row from constant");
    run<4, 3>("PX=4 rows 0,1 shared + 2,3 constant");
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
