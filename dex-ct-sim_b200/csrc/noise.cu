// Quantum noise on detected counts and virtual monoenergetic images from basis images.
//
// The reference scales its spectra to photons per detector pixel per view (/root/reference/main.py:68,
// spec.rescale_counts(A_iso * dose / N_proj)) and studies image noise against dose (plots.py:631-717), so the simulated
// counts are Poisson variates; the sampler itself sits in the absent xtomosim submodule.  Here: a counter-based generator
// (Philox4x32-10, Salmon et al. 2011) addressed by the GLOBAL element index, so a sinogram does not depend on how its
// views are sharded over GPUs, and an exact Poisson sampler (inversion below lambda = 10, Hoermann's PTRS transformed
// rejection above).  oracle/noise_oracle.py restates both operation for operation; the arithmetic below is written with
// single-rounding intrinsics so that no FMA contraction separates the two.
#include "common.cuh"

namespace dexct {
namespace {

struct U4 { uint32_t x, y, z, w; };

__device__ __forceinline__ U4 philox4x32_10(U4 c, uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = U4{hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0};
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c;
}

// uniform on (0, 1): (x + 1/2) 2^-32, exact in double
__device__ __forceinline__ double u01(uint32_t x) { return __dmul_rn(__dadd_rn((double)x, 0.5), 2.3283064365386963e-10); }

__device__ __forceinline__ uint32_t word(const U4 &v, int i) { return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w; }

// One Poisson(lam) variate for element `idx` of random stream `stream_id` under key (k0, k1).
__device__ double poisson_sample(double lam, uint64_t idx, uint32_t k0, uint32_t k1, uint32_t stream_id)
{
    if (!(lam > 0.0)) return lam != lam ? lam : 0.0;          // NaN stays NaN, non-positive means gives 0
    const uint32_t c0 = (uint32_t)idx, c1 = (uint32_t)(idx >> 32);
    if (lam < 10.0) {
        // inversion by sequential search from one uniform
        const double u = u01(philox4x32_10(U4{c0, c1, 0u, stream_id}, k0, k1).x);
        double p = exp(-lam), F = p;
        int k = 0;
        while (u > F && k < 200) {
            ++k;
            p = __dmul_rn(p, __ddiv_rn(lam, (double)k));
            F = __dadd_rn(F, p);
        }
        return (double)k;
    }
    // PTRS (W. Hoermann, "The transformed rejection method for generating Poisson random variables", 1993)
    const double slam = __dsqrt_rn(lam), loglam = log(lam);
    const double b = __dadd_rn(0.931, __dmul_rn(2.53, slam));
    const double a = __dadd_rn(-0.059, __dmul_rn(0.02483, b));
    const double invalpha = __dadd_rn(1.1239, __ddiv_rn(1.1328, __dsub_rn(b, 3.4)));
    const double vr = __dsub_rn(0.9277, __ddiv_rn(3.6224, __dsub_rn(b, 2.0)));
    U4 blk{};
    for (int t = 0; t < 64; ++t) {
        if ((t & 1) == 0) blk = philox4x32_10(U4{c0, c1, (uint32_t)(t >> 1), stream_id}, k0, k1);
        const double U = __dsub_rn(u01(word(blk, 2 * (t & 1))), 0.5);
        const double V = u01(word(blk, 2 * (t & 1) + 1));
        const double us = __dsub_rn(0.5, fabs(U));
        const double k = floor(__dadd_rn(__dadd_rn(__dmul_rn(__dadd_rn(__ddiv_rn(__dmul_rn(2.0, a), us), b), U), lam), 0.43));
        if (us >= 0.07 && V <= vr) return k;
        if (k < 0.0 || (us < 0.013 && V > us)) continue;
        const double lhs = __dsub_rn(__dadd_rn(log(V), log(invalpha)), log(__dadd_rn(__ddiv_rn(a, __dmul_rn(us, us)), b)));
        const double rhs = __dsub_rn(__dadd_rn(-lam, __dmul_rn(k, loglam)), lgamma(__dadd_rn(k, 1.0)));
        if (lhs <= rhs) return k;
    }
    return floor(__dadd_rn(lam, 0.5));                       // 64 rejections in a row: probability < 1e-60
}

template <typename R>
__global__ void __launch_bounds__(256) poisson_kernel(R *__restrict__ counts, R *__restrict__ logs, long long n,
                                                      unsigned long long offset, uint32_t k0, uint32_t k1,
                                                      uint32_t stream_id, double air, double floor_counts)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double k = poisson_sample((double)counts[i], offset + (unsigned long long)i, k0, k1, stream_id);
        counts[i] = (R)k;
        if (logs) logs[i] = (R)(-log(__ddiv_rn(fmax(k, floor_counts), air)));
    }
}

template <typename R>
__global__ void __launch_bounds__(256) vmi_kernel(const R *__restrict__ m1, const R *__restrict__ m2, long long n,
                                                  R u1, R u2, R uw, int hu, float *__restrict__ out)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        // plots.py:141-143: vmi = u1 M1 + u2 M2 (two products, one sum), then 1000 (vmi - uw) / uw
        R v = Rn<R>::add(Rn<R>::mul(u1, m1[i]), Rn<R>::mul(u2, m2[i]));
        if (hu) v = Rn<R>::mul((R)1000, Rn<R>::sub(v, uw)) / uw;
        out[i] = (float)v;
    }
}

}  // namespace
}  // namespace dexct

using namespace dexct;

extern "C" int dexct_poisson_noise(void *counts, void *logs, int64_t n, uint64_t seed, uint32_t stream_id, uint64_t offset,
                                   double air_counts, double floor_counts, int32_t dtype, void *stream)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(n >= 0, "n >= 0 required");
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    if (n == 0) return DEXCT_OK;
    DEXCT_REQUIRE(counts, "counts must be non-null");
    DEXCT_REQUIRE(!logs || (air_counts > 0 && floor_counts > 0), "air_counts > 0 and floor_counts > 0 required with logs");
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = (unsigned)min((long long)sm_count() * 8, (long long)((n + 255) / 256));
    const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    if (dtype == DEXCT_F32)
        poisson_kernel<float><<<grid, 256, 0, st>>>((float *)counts, (float *)logs, n, offset, k0, k1, stream_id, air_counts, floor_counts);
    else
        poisson_kernel<double><<<grid, 256, 0, st>>>((double *)counts, (double *)logs, n, offset, k0, k1, stream_id, air_counts, floor_counts);
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}

extern "C" int dexct_vmi(const void *m1, const void *m2, int64_t n, double u1, double u2, double u_water, int32_t hu,
                         float *out, int32_t dtype, void *stream)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(n >= 0, "n >= 0 required");
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    DEXCT_REQUIRE(!hu || u_water > 0, "u_water > 0 required for Hounsfield output");
    if (n == 0) return DEXCT_OK;
    DEXCT_REQUIRE(m1 && m2 && out, "m1, m2 and out must be non-null");
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = (unsigned)min((long long)sm_count() * 8, (long long)((n + 255) / 256));
    if (dtype == DEXCT_F32)
        vmi_kernel<float><<<grid, 256, 0, st>>>((const float *)m1, (const float *)m2, n, (float)u1, (float)u2, (float)u_water, hu, out);
    else
        vmi_kernel<double><<<grid, 256, 0, st>>>((const double *)m1, (const double *)m2, n, u1, u2, u_water, hu, out);
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}
