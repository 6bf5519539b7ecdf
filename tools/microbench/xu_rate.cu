// Microbenchmark: MUFU.EX2 throughput per SM on sm_100a, alone and with the other instructions of a gn_tc_kernel k-step
// (per exponential: half an FMUL2, half an FFMA2, one LOP3, half an FADD2), as a function of resident warps per scheduler.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o xu_rate xu_rate.cu ; run: ./xu_rate
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float x, float y) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(x), "f"(y)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float &x, float &y) { asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(v)); }
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

template <int MODE>
__global__ void k(float *out, int reps, float a1, float a2)
{
    float acc = 0.f;
    const f32x2 a1p = pack2(a1, a1), a2p = pack2(a2, a2);
    float t = threadIdx.x * 1e-6f;
    for (int r = 0; r < reps; ++r) {
        float e[8];
        if (MODE == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) e[j] = ex2(t + j);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc += e[j];          // 8 FADD per 8 MUFU
        } else {
            unsigned h[8];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                f32x2 m, xx;
                asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(m) : "l"(a1p), "l"(pack2(t + j, t - j)));
                asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(xx) : "l"(a2p), "l"(pack2(t, t + 1.f)), "l"(m));
                float x0, x1;
                unpack2(xx, x0, x1);
                e[2 * j] = ex2(x0); e[2 * j + 1] = ex2(x1);
                h[2 * j] = __float_as_uint(e[2 * j]) & 0xffffe000u;
                h[2 * j + 1] = __float_as_uint(e[2 * j + 1]) & 0xffffe000u;
                f32x2 l;
                asm volatile("sub.rn.f32x2 %0, %1, %2;" : "=l"(l) : "l"(pack2(e[2 * j], e[2 * j + 1])), "l"(pack2(__uint_as_float(h[2 * j]), __uint_as_float(h[2 * j + 1]))));
                float l0, l1;
                unpack2(l, l0, l1);
                acc += l0 + l1;
            }
        }
        t += 1e-3f;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int MODE>
static void run(const char *name, int warps_per_smsp)
{
    int dev = 0, sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    const int threads = 128 * warps_per_smsp, reps = 20000;
    float *out;
    cudaMalloc(&out, sizeof(float) * sms * threads);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 3; ++w) k<MODE><<<sms, threads>>>(out, reps, 1.0f, 0.5f);
    cudaEventRecord(e0);
    for (int w = 0; w < 5; ++w) k<MODE><<<sms, threads>>>(out, reps, 1.0f, 0.5f);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double ops = 5.0 * sms * threads * (double)reps * 8.0;
    printf("%-34s %2d warps per scheduler: %6.2f ex2 per nominal clock and SM (%.3f ms per launch, nominal %.3f GHz)\n", name, warps_per_smsp,
           ops / (ms * 1e-3) / sms / (khz * 1e3), ms / 5, khz * 1e-6);
    cudaFree(out);
}

int main()
{
    for (int w : {1, 2, 4, 8}) run<0>("MUFU.EX2 + FADD", w);
    for (int w : {1, 2, 4, 8}) run<1>("k-step mix (no memory, no tmem)", w);
    return 0;
}
