// Gauss-Newton two-basis material decomposition, one independent 2x2 Newton problem per sinogram pixel.
// Replaces optimize_sino_cpu (/root/reference/matdecomp.py:87-127) and the masking of get_basismat_sinos
// (matdecomp.py:195-205).
//
// Per pixel and iteration (matdecomp.py:116-125):
//   att_e   = exp(clip(-(a1 mu1_e + a2 mu2_e), +-700))
//   nu_m    = sum_e i0_me att_e             grad_mp = -sum_e i0_me mu_pe att_e        hess_mpq = sum_e i0_me mu_pe mu_qe att_e
//   r_m     = g_m/nu_m - 1                  dF_p    = -sum_m r_m grad_mp
//   H_pq    = -sum_m ( r_m hess_mpq - g_m/nu_m^2 grad_mp grad_mq )                    a -= H^-1 dF
// The twelve energy sums are twelve FMAs against a per-energy table row kept in shared memory:
//   row_e = { -mu1_e*s, -mu2_e*s, i0_1e {1, mu1, mu2, mu1^2, mu1 mu2, mu2^2}, i0_2e {...} }   (s = log2(e) in float)
// Every thread of a warp reads the same row (broadcast LDS.128), so the kernel is bound by the FP32 FMA pipe:
// 14 FMA + 1 MUFU.EX2 per pixel, energy and iteration.  PX pixels per thread share each row load.
#include "common.cuh"
#include <mutex>
#include "tc05.cuh"
#include <algorithm>
#include <stdlib.h>

namespace dexct {

constexpr int GN_ROW = 16;
constexpr float GN_A_FLOOR = -20.0f;      // single precision hands pixels whose iterates go below this to the float64 pass

// Packed single-precision pairs (FFMA2 on sm_100a: two FMAs per issue slot).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float x, float y)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(x), "f"(y));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float &x, float &y) { asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(v)); }
template <typename T> __device__ __forceinline__ void unpack2(f32x2, T &, T &) {}      // never called: double path
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}   // floats per table row (14 used), 64 B so a row is four LDS.128

template <typename R> struct GnMath;
template <> struct GnMath<float> {
    static constexpr float scale = 1.4426950408889634f;                 // exponent kept in the log2 domain
    static __device__ __forceinline__ float att(float x)                // x = -(a.mu) * log2(e)
    {
        // float: the reference's clip(-700, 700) is the identity here -- anything beyond +-128 (log2 domain) already
        // saturates to 0 / inf in single precision, exactly as the clipped value would.
        float y;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
        return y;
    }
    static __device__ __forceinline__ float fma(float a, float b, float c) { return fmaf(a, b, c); }
};
template <> struct GnMath<double> {
    static constexpr double scale = 1.0;
    static __device__ __forceinline__ double att(double x) { return exp(fmin(fmax(x, -700.0), 700.0)); }   // :116
    static __device__ __forceinline__ double fma(double a, double b, double c) { return ::fma(a, b, c); }
};

// max over g[0][0..P) -> *out (scratch).  Non-negative counts: compare as ordered ints after a float reinterpretation
// is avoided on purpose; a two-level block reduction plus one atomic per block on an integer view of the value.
template <typename R>
__global__ void gn_max_kernel(const R *__restrict__ g, long long P, R *out)
{
    R m = -INFINITY;
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += (long long)gridDim.x * blockDim.x)
        m = fmax(m, g[p]);
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    __shared__ R wm[32];
    if ((threadIdx.x & 31) == 0) wm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = threadIdx.x < (blockDim.x >> 5) ? wm[threadIdx.x] : -INFINITY;
        for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (threadIdx.x == 0) {
            // atomic max on a monotone integer key of the floating value (works for negatives too)
            if (sizeof(R) == 4) {
                int k = __float_as_int((float)m);
                k = k >= 0 ? k : k ^ 0x7fffffff;
                atomicMax((int *)out, k);
            } else {
                long long k = __double_as_longlong((double)m);
                k = k >= 0 ? k : k ^ 0x7fffffffffffffffLL;
                atomicMax((long long *)out, k);
            }
        }
    }
}

template <typename R> __device__ __forceinline__ R decode_max(const void *scratch);
template <> __device__ __forceinline__ float decode_max<float>(const void *s)
{
    int k = *(const int *)s;
    k = k >= 0 ? k : k ^ 0x7fffffff;
    return __int_as_float(k);
}
template <> __device__ __forceinline__ double decode_max<double>(const void *s)
{
    long long k = *(const long long *)s;
    k = k >= 0 ? k : k ^ 0x7fffffffffffffffLL;
    return __longlong_as_double(k);
}

// Single precision keeps the table in the constant bank: every thread of a warp reads the same row, and uniform loads
// (LDCU into uniform registers, consumed directly by FFMA2) cost no shared-memory bandwidth -- measured with
// tools/microbench/gn_mix.cu (profiles/r01f_gn_mix_microbench.txt): a row fed from registers runs at 12-13 cycles per
// (pixel, energy) and SMSP, from the constant bank at 17, from shared memory (four broadcast LDS.128) at 22.
// Two tables, so that the exponentials of the NEXT bin can be issued ahead of the FFMA2 of the current one (software
// pipeline, the microbenchmark's "pipelined" rows): the exponent pair of a bin, and its six moment pairs.
constexpr int GN_CTAB_ROWS = 512;                             // nE + 2 zero rows must fit: nE <= 510
constexpr int GN_CROW = 12;                                   // floats per moment row: 6 (measurement 1, measurement 2) pairs
__constant__ ulonglong2 gn_ctab[GN_CTAB_ROWS * (GN_CROW / 4)];
__constant__ f32x2 gn_cexp[GN_CTAB_ROWS];
// Unified form (loop variants 1..3): one 64-byte row per bin = { exponent pair, six moment pairs, pad }, so that the loop
// carries ONE uniform byte offset (add, compare, branch) instead of two scaled indices (eight uniform-ALU instructions per
// trip in variant 0: profiles/r02a_gn_sass_loop.txt).  gn_crow overlays nothing: variant 0 keeps its own two tables.
constexpr int GN_UROW = 16;                                   // floats per unified row
__constant__ ulonglong2 gn_crow[(GN_CTAB_ROWS + 4) * (GN_UROW / 4)];

// Staging image of both tables: [nE + 2][12] moment rows, then [nE + 2][2] exponent pairs; the two rows past nE are zero.
__global__ void gn_table_kernel(const float *__restrict__ i0, const float *__restrict__ mus, int nE, float *__restrict__ tab)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nE + 2) return;
    float *row = tab + e * GN_CROW;
    float *ex = tab + (nE + 2) * GN_CROW + 2 * e;
    if (e >= nE) {
#pragma unroll
        for (int q = 0; q < GN_CROW; ++q) row[q] = 0.f;
        ex[0] = ex[1] = 0.f;
        return;
    }
    const float m1 = mus[e], m2 = mus[nE + e], s1 = i0[e], s2 = i0[nE + e];
    ex[0] = -m1 * GnMath<float>::scale; ex[1] = -m2 * GnMath<float>::scale;
    const float f[6] = {1.f, m1, m2, m1 * m1, m1 * m2, m2 * m2};
#pragma unroll
    for (int q = 0; q < 6; ++q) { row[2 * q] = s1 * f[q]; row[2 * q + 1] = s2 * f[q]; }
}

// Unified rows: [nE + 4][16] = { -mu1 s, -mu2 s, six (measurement 1, measurement 2) moment pairs, 0, 0 }; four zero rows past nE.
__global__ void gn_table_unified_kernel(const float *__restrict__ i0, const float *__restrict__ mus, int nE, float *__restrict__ tab)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nE + 4) return;
    float *row = tab + e * GN_UROW;
#pragma unroll
    for (int q = 0; q < GN_UROW; ++q) row[q] = 0.f;
    if (e >= nE) return;
    const float m1 = mus[e], m2 = mus[nE + e], s1 = i0[e], s2 = i0[nE + e];
    row[0] = -m1 * GnMath<float>::scale; row[1] = -m2 * GnMath<float>::scale;
    const float f[6] = {1.f, m1, m2, m1 * m1, m1 * m2, m2 * m2};
#pragma unroll
    for (int q = 0; q < 6; ++q) { row[2 + 2 * q] = s1 * f[q]; row[3 + 2 * q] = s2 * f[q]; }
}

// Further copies of the result in peer memory (view-sharded scans, peer.cu): the epilogue stores every pixel to each of
// them as well, so the basis sinograms are assembled on all ranks by the time the kernel ends -- no gather afterwards.
struct GnMirrors { void *p[DEXCT_MAX_PEERS]; int n; };
// A finished buffer (the count/log shard the projection just wrote) that rides along: the threads of the decomposition
// kernel copy it into the peers' copies a slice per Newton iteration, so the transfer is spread evenly over all SMs and
// over the kernel's run time instead of competing for a few SMs as a separate copy kernel would (peer.cu).
struct GnRider { const uint4 *src; long long n16; void *dst[DEXCT_MAX_PEERS]; int n; };

__device__ __forceinline__ void gn_rider_slice(const GnRider &rd, int k, int nk)
{
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x + nthreads * k; i < rd.n16; i += nthreads * nk) {
        const uint4 v = rd.src[i];
        for (int d = 0; d < rd.n; ++d) ((uint4 *)rd.dst[d])[i] = v;
    }
}


// One Newton step of a pixel from its twelve energy sums A[0..5] (measurement 1), A[6..11] (measurement 2):
// {nu, sum i0 mu1 att, sum i0 mu2 att, sum i0 mu1^2 att, sum i0 mu1 mu2 att, sum i0 mu2^2 att}  (matdecomp.py:118-125).
template <typename R>
__device__ __forceinline__ void newton_update(const R *A, R g1, R g2, R &a1, R &a2, bool &sane)
{
    typedef GnMath<R> M;
    const R nu1 = A[0], nu2 = A[6];
    if (sizeof(R) == 4) {
        // single precision cannot follow the reference where its exponent clip (matdecomp.py:116) is what keeps
        // the sums finite: any sum beyond these bounds (inf and NaN included) sends the pixel to the float64 pass
        R big = fmax(fabs(A[3]), fabs(A[9]));
        big = fmax(big, fmax(fmax(fabs(A[4]), fabs(A[10])), fmax(fabs(A[5]), fabs(A[11]))));
        sane = sane && (nu1 < (R)1e18) && (nu2 < (R)1e18) && (big < (R)1e30) && (nu1 > (R)0) && (nu2 > (R)0);
    }
    const R r1 = g1 / nu1 - (R)1, r2 = g2 / nu2 - (R)1;            // g/nu - 1
    const R s1 = g1 / (nu1 * nu1), s2 = g2 / (nu2 * nu2);          // g/nu^2
    // grad_mp = -A[1+p]; the two minus signs of dF cancel, grad grad^T is sign-free
    const R d1 = r1 * A[1] + r2 * A[7];
    const R d2 = r1 * A[2] + r2 * A[8];
    const R h11 = -(r1 * A[3] + r2 * A[9]) + (s1 * A[1] * A[1] + s2 * A[7] * A[7]);
    const R h12 = -(r1 * A[4] + r2 * A[10]) + (s1 * A[1] * A[2] + s2 * A[7] * A[8]);
    const R h22 = -(r1 * A[5] + r2 * A[11]) + (s1 * A[2] * A[2] + s2 * A[8] * A[8]);
    // 2x2 solve; the determinant is a difference of like-sized products -> compensate the first product
    const R w = h12 * h12;
    const R err = M::fma(-h12, h12, w);
    const R det = M::fma(h11, h22, -w) + err;
    const R inv = (R)1 / det;
    a1 -= (h22 * d1 - h12 * d2) * inv;
    a2 -= (h11 * d2 - h12 * d1) * inv;
    // An indefinite Hessian means the iterate has left the basin of the minimum: from there the Newton map is chaotic and
    // single precision cannot retrace the reference's float64 trajectory; neither can it where a basis thickness runs far
    // below zero (photon-starved pixels: the exponent a.mu is then a difference of huge terms).  GN_A_FLOOR [g/cm2] is far
    // outside anything physical.
    if (sizeof(R) == 4) sane = sane && det > (R)0 && h11 > (R)0 && fmin(a1, a2) > (R)GN_A_FLOOR;
}

// The same step for the tensor-core kernel, where the instructions of the update are no longer hidden behind an FMA-bound
// loop (they were 14 % of the kernel: profiles/r02n): three approximate reciprocals (MUFU.RCP, 1 ulp) instead of five IEEE
// divisions.  g/nu - 1 is formed with one FMA, so the rounding of the quotient does not enter the residual twice.
__device__ __forceinline__ float rcp_fast(float x)
{
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void newton_update_fast(const float *A, float g1, float g2, float &a1, float &a2, bool &sane)
{
    const float nu1 = A[0], nu2 = A[6];
    float big = fmaxf(fabsf(A[3]), fabsf(A[9]));
    big = fmaxf(big, fmaxf(fmaxf(fabsf(A[4]), fabsf(A[10])), fmaxf(fabsf(A[5]), fabsf(A[11]))));
    sane = sane && (nu1 < 1e18f) && (nu2 < 1e18f) && (big < 1e30f) && (nu1 > 0.f) && (nu2 > 0.f);
    const float i1 = rcp_fast(nu1), i2 = rcp_fast(nu2);
    const float q1 = g1 * i1, q2 = g2 * i2;
    const float r1 = fmaf(g1, i1, -1.f), r2 = fmaf(g2, i2, -1.f);   // g/nu - 1
    const float s1 = q1 * i1, s2 = q2 * i2;                        // g/nu^2
    const float d1 = r1 * A[1] + r2 * A[7];
    const float d2 = r1 * A[2] + r2 * A[8];
    const float h11 = -(r1 * A[3] + r2 * A[9]) + (s1 * A[1] * A[1] + s2 * A[7] * A[7]);
    const float h12 = -(r1 * A[4] + r2 * A[10]) + (s1 * A[1] * A[2] + s2 * A[7] * A[8]);
    const float h22 = -(r1 * A[5] + r2 * A[11]) + (s1 * A[2] * A[2] + s2 * A[8] * A[8]);
    const float w = h12 * h12;
    const float err = fmaf(-h12, h12, w);
    const float det = fmaf(h11, h22, -w) + err;
    const float inv = rcp_fast(det);
    a1 -= (h22 * d1 - h12 * d2) * inv;
    a2 -= (h11 * d2 - h12 * d1) * inv;
    sane = sane && det > 0.f && h11 > 0.f && fminf(a1, a2) > GN_A_FLOOR;
}

template <typename R, int PX, bool CTAB, int VAR>
__global__ void __launch_bounds__(256, (PX == 4 && CTAB) ? 3 : 1) gn_kernel(const R *__restrict__ g, const R *__restrict__ i0, const R *__restrict__ mus,
                                                 int nE, long long P, int n_iters, R mask_thresh, void *scratch,
                                                 R *__restrict__ a_out, const GnMirrors mir, const GnRider rider)
{
    typedef GnMath<R> M;
    constexpr bool PACKED = sizeof(R) == 4;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    R *tab = (R *)smem_raw;                                   // [nE][GN_ROW]
    for (int e = threadIdx.x; !CTAB && e < nE; e += blockDim.x) {
        const R m1 = mus[e], m2 = mus[nE + e], s1 = i0[e], s2 = i0[nE + e];
        R *row = tab + e * GN_ROW;
        row[0] = -m1 * M::scale; row[1] = -m2 * M::scale;
        const R f[6] = {(R)1, m1, m2, m1 * m1, m1 * m2, m2 * m2};
        if (PACKED) {                     // (measurement 1, measurement 2) pairs: one FFMA2 per moment
#pragma unroll
            for (int q = 0; q < 6; ++q) { row[2 + 2 * q] = s1 * f[q]; row[3 + 2 * q] = s2 * f[q]; }
        } else {
#pragma unroll
            for (int q = 0; q < 6; ++q) { row[2 + q] = s1 * f[q]; row[8 + q] = s2 * f[q]; }
        }
        row[14] = (R)0; row[15] = (R)0;
    }
    if (!CTAB) {
        for (int t = threadIdx.x; t < 2 * GN_ROW; t += blockDim.x) tab[nE * GN_ROW + t] = (R)0;  // two zero rows (pipelined loop)
        __syncthreads();
    }

    const R thr = mask_thresh > (R)0 ? mask_thresh * decode_max<R>(scratch) : (R)INFINITY;

    // CTA-stride over groups of blockDim * PX pixels; pixel p_k = base + k * blockDim keeps every global access coalesced.
    // The loop bounds are uniform across the CTA on purpose: inside a loop whose trip count depends on the thread the
    // compiler may not use the uniform datapath, and the table reads fall back from LDCU to per-thread LDC (2x slower).
    const long long stride = blockDim.x;
    const long long n_groups = (P + stride * PX - 1) / (stride * PX);
    // the rider's copy is cut into one slice per (group of this CTA, Newton iteration), so that it trickles out over the
    // whole run of the kernel whatever the number of pixels; CTAs that own one group fewer send their last slices at the end
    const int rider_per_grp = n_iters > 0 ? n_iters : 1;
    const int rider_nk = rider_per_grp * (int)((n_groups + gridDim.x - 1) / gridDim.x);
    int rider_k = 0;
    const bool ride = rider.n != 0;                           // CTA-uniform
    for (long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x) {
        const long long base = grp * stride * PX + threadIdx.x;
        R a1[PX], a2[PX], g1[PX], g2[PX];
        bool sane[PX];
#pragma unroll
        for (int k = 0; k < PX; ++k) {
            const long long p = base + k * stride;
            sane[k] = true;
            a1[k] = a2[k] = (R)1e-6;                                         // matdecomp.py:98
            g1[k] = p < P ? g[p] : (R)1;
            g2[k] = p < P ? g[P + p] : (R)1;
        }
        for (int it = 0; it < n_iters; ++it) {
            if (ride) gn_rider_slice(rider, rider_k++, rider_nk);
            R acc[PX][12];
            if constexpr (PACKED) {
                // FFMA2 (sm_100a): accumulator pair q = (measurement 1, measurement 2) of moment q; att is broadcast
                f32x2 acc2[PX][6];
#pragma unroll
                for (int k = 0; k < PX; ++k)
#pragma unroll
                    for (int q = 0; q < 6; ++q) acc2[k][q] = 0ull;
                // Software pipeline over energy pairs: the exponentials of the NEXT bin are issued before the twelve
                // FFMA2 of the current one, so MUFU and LDS latency hide behind the FMA pipe.  The table carries two
                // zero rows past nE (i0 = 0: no contribution), which keeps the loop free of tails and bounds checks.
                const ulonglong2 *rows = (const ulonglong2 *)tab;
                ulonglong2 rA[4], rB[4];
                float attA[PX], attB[PX];
                auto load_row = [&](int e, ulonglong2(&r)[4]) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) r[q] = rows[e * (GN_ROW / 4) + q];
                };
                auto exps = [&](const ulonglong2(&r)[4], float(&att)[PX]) {
                    float t0, t1;
                    unpack2(r[0].x, t0, t1);
#pragma unroll
                    for (int k = 0; k < PX; ++k) att[k] = M::att(fmaf(a2[k], t1, a1[k] * t0));
                };
                auto accum = [&](const ulonglong2(&r)[4], const float(&att)[PX]) {
#pragma unroll
                    for (int k = 0; k < PX; ++k) {
                        const f32x2 att2 = pack2(att[k], att[k]);
                        acc2[k][0] = ffma2(r[0].y, att2, acc2[k][0]);
                        acc2[k][1] = ffma2(r[1].x, att2, acc2[k][1]);
                        acc2[k][2] = ffma2(r[1].y, att2, acc2[k][2]);
                        acc2[k][3] = ffma2(r[2].x, att2, acc2[k][3]);
                        acc2[k][4] = ffma2(r[2].y, att2, acc2[k][4]);
                        acc2[k][5] = ffma2(r[3].x, att2, acc2[k][5]);
                    }
                };
                if constexpr (CTAB && VAR >= 1) {
                    // unified 64-byte rows behind one uniform byte offset; pipelined like variant 0
                    float att[PX], nxt[PX];
                    auto exps_u = [&](f32x2 tt, float(&o)[PX]) {
                        float t0, t1;
                        unpack2(tt, t0, t1);
#pragma unroll
                        for (int k = 0; k < PX; ++k) o[k] = M::att(fmaf(a2[k], t1, a1[k] * t0));
                    };
                    auto accum_u = [&](const ulonglong2 *r, const float(&w)[PX]) {
                        const f32x2 m0 = r[0].y, m5 = r[3].x;
                        const ulonglong2 r1 = r[1], r2 = r[2];
#pragma unroll
                        for (int k = 0; k < PX; ++k) {
                            const f32x2 att2 = pack2(w[k], w[k]);
                            acc2[k][0] = ffma2(m0, att2, acc2[k][0]);
                            acc2[k][1] = ffma2(r1.x, att2, acc2[k][1]);
                            acc2[k][2] = ffma2(r1.y, att2, acc2[k][2]);
                            acc2[k][3] = ffma2(r2.x, att2, acc2[k][3]);
                            acc2[k][4] = ffma2(r2.y, att2, acc2[k][4]);
                            acc2[k][5] = ffma2(m5, att2, acc2[k][5]);
                        }
                    };
                    const ulonglong2 *r = gn_crow;
                    const ulonglong2 *const rend = gn_crow + 4 * nE;
                    exps_u(r[0].x, att);
                    if constexpr (VAR == 1) {
#pragma unroll 1
                        for (; r < rend; r += 8) {
                            exps_u(r[4].x, nxt);
                            accum_u(r, att);
                            exps_u(r[8].x, att);
                            accum_u(r + 4, nxt);
                        }
                    } else {
#pragma unroll 1
                        for (; r < rend; r += 16) {
                            exps_u(r[4].x, nxt);
                            accum_u(r, att);
                            exps_u(r[8].x, att);
                            accum_u(r + 4, nxt);
                            exps_u(r[12].x, nxt);
                            accum_u(r + 8, att);
                            exps_u(r[16].x, att);
                            accum_u(r + 12, nxt);
                        }
                    }
                } else if constexpr (CTAB) {
                    // pipelined over bins: exponentials of bin e + 1 first, then the FFMA2 of bin e on its moment row
                    float att[PX], nxt[PX];
                    {
                        float t0, t1;
                        unpack2(gn_cexp[0], t0, t1);
#pragma unroll
                        for (int k = 0; k < PX; ++k) att[k] = M::att(fmaf(a2[k], t1, a1[k] * t0));
                    }
                    auto exps_c = [&](int e, float(&o)[PX]) {
                        float t0, t1;
                        unpack2(gn_cexp[e], t0, t1);
#pragma unroll
                        for (int k = 0; k < PX; ++k) o[k] = M::att(fmaf(a2[k], t1, a1[k] * t0));
                    };
                    auto accum_c = [&](int e, const float(&w)[PX]) {
                        const ulonglong2 r0 = gn_ctab[e * 3], r1 = gn_ctab[e * 3 + 1], r2 = gn_ctab[e * 3 + 2];
#pragma unroll
                        for (int k = 0; k < PX; ++k) {
                            const f32x2 att2 = pack2(w[k], w[k]);
                            acc2[k][0] = ffma2(r0.x, att2, acc2[k][0]);
                            acc2[k][1] = ffma2(r0.y, att2, acc2[k][1]);
                            acc2[k][2] = ffma2(r1.x, att2, acc2[k][2]);
                            acc2[k][3] = ffma2(r1.y, att2, acc2[k][3]);
                            acc2[k][4] = ffma2(r2.x, att2, acc2[k][4]);
                            acc2[k][5] = ffma2(r2.y, att2, acc2[k][5]);
                        }
                    };
                    // two bins per trip (the zero rows past nE make an odd nE harmless); a rolled loop with a uniform
                    // counter keeps the table reads on the uniform datapath (LDCU), which an unrolled one loses to LDC
#pragma unroll 1
                    for (int e = 0; e < nE; e += 2) {
                        exps_c(e + 1, nxt);
                        accum_c(e, att);
                        exps_c(e + 2, att);
                        accum_c(e + 1, nxt);
                    }
                } else {
                    load_row(0, rA);
                    exps(rA, attA);
#pragma unroll 1
                    for (int e = 0; e < nE; e += 2) {
                        load_row(e + 1, rB);
                        exps(rB, attB);
                        accum(rA, attA);
                        load_row(e + 2, rA);
                        exps(rA, attA);
                        accum(rB, attB);
                    }
                }
#pragma unroll
                for (int k = 0; k < PX; ++k)
#pragma unroll
                    for (int q = 0; q < 6; ++q) unpack2(acc2[k][q], acc[k][q], acc[k][6 + q]);
            } else {
#pragma unroll
            for (int k = 0; k < PX; ++k)
#pragma unroll
                for (int q = 0; q < 12; ++q) acc[k][q] = (R)0;
#pragma unroll 2
            for (int e = 0; e < nE; ++e) {
                const R *row = tab + e * GN_ROW;
                R t[14];
#pragma unroll
                for (int q = 0; q < 14; ++q) t[q] = row[q];
#pragma unroll
                for (int k = 0; k < PX; ++k) {
                    const R att = M::att(M::fma(a2[k], t[1], a1[k] * t[0]));
#pragma unroll
                    for (int q = 0; q < 12; ++q) acc[k][q] = M::fma(t[2 + q], att, acc[k][q]);
                }
            }
            }
#pragma unroll
            for (int k = 0; k < PX; ++k) newton_update<R>(acc[k], g1[k], g2[k], a1[k], a2[k], sane[k]);
        }
#pragma unroll
        for (int k = 0; k < PX; ++k) {
            const long long p = base + k * stride;
            if (p < P) {
                const bool masked = g1[k] >= thr;                             // matdecomp.py:196,204-205
                // a_out [P][2]: two adjacent values per pixel
                if (sizeof(R) == 4) {
                    float2 o = make_float2(masked ? 0.f : (float)a1[k], masked ? 0.f : (float)a2[k]);
                    ((float2 *)a_out)[p] = o;
                    for (int d = 0; d < mir.n; ++d) ((float2 *)mir.p[d])[p] = o;
                    const bool finite = fabs(a1[k]) < (R)INFINITY && fabs(a2[k]) < (R)INFINITY;
                    if (!masked && !(sane[k] && finite)) {                    // rare: hand the pixel to gn_fixup_kernel
                        unsigned *fl = (unsigned *)scratch;
                        fl[4 + atomicAdd(fl + 2, 1u)] = (unsigned)p;
                    }
                } else {
                    double2 o = make_double2(masked ? 0.0 : (double)a1[k], masked ? 0.0 : (double)a2[k]);
                    ((double2 *)a_out)[p] = o;
                    for (int d = 0; d < mir.n; ++d) ((double2 *)mir.p[d])[p] = o;
                }
            }
        }
    }
    if (ride)
        for (; rider_k < rider_nk; ++rider_k) gn_rider_slice(rider, rider_k, rider_nk);
}

// ---------------------------------------------------------------------------------------------------------------------
// Tensor-core form of the single-precision decomposition (sm_100a: tcgen05.mma + tensor memory).
//
// The twelve energy sums of a pixel are one row of the product  att[pixel][bin] x row[bin][moment]  -- a contraction over
// the energy bins with N = 12 (padded to 16) columns -- and they are 24 of the 28 floating-point operations per (pixel,
// bin).  Here they leave the FP32 pipe: a CTA owns a tile of 128 pixels (thread t <-> pixel t <-> lane t of tensor memory);
// per chunk of bins each thread evaluates its exponentials, splits every value into a TF32 head and a tail
// (att = hi + lo exactly: hi = att with the low 13 mantissa bits cleared) and stores both into tensor memory as the A operand
// (tcgen05.st); one thread of a fifth warp then issues, per eight bins, two M128 x K8 TF32 MMAs
//     D[0..31] += hi x [Bhi | Blo]  (N = 32),     D[0..15] += lo x Bhi  (N = 16)      (Bhi + Blo = the moment rows, split once)
// -- the "3xTF32" scheme: products of 11-bit significands are exact in the FP32 accumulator, and the one missing term
// lo x Blo is below 2^-22 of the sum.  An MMA costs ~45 cycles whatever its N up to 64 (tools/microbench/tc_rate.cu,
// profiles/r02c_tc_rate.txt), so the two products with hi share one instruction; the thread adds the two halves of its
// accumulator row when it reads it back.  B is read from shared memory through a K-major no-swizzle matrix descriptor, D
// (128 lanes x 32 columns of FP32) accumulates in tensor memory over all bins of a Newton iteration.  The threads read their
// row of D back (tcgen05.ld), do the 2x2 Newton step in registers as before, and go round again.  A-operand chunks are double
// buffered (mbarriers: full[b] by the compute warps, free[b] and done by tcgen05.commit), four CTAs share an SM (4 x 128
// columns of tensor memory), so exponentials, MMAs and Newton updates of different tiles overlap.
// What is left per (pixel, bin) on the ordinary pipes: FMUL2/FFMA2 halves for the exponent, one MUFU.EX2, one LOP3, half an
// FADD2 -- the kernel is bound by the MUFU unit (16 ex2 per clock and SM) instead of the FMA pipe.
constexpr int GT_KS = 8;                    // bins per MMA: 32 bytes of TF32 along K
#ifndef GT_CHUNK_KS_VARIANT                 // (tuning builds: make variant VARFLAGS="-DGT_CHUNK_KS_VARIANT=2 -DGT_NBUF_VARIANT=3")
#define GT_CHUNK_KS_VARIANT 3
#endif
#ifndef GT_NBUF_VARIANT
#define GT_NBUF_VARIANT 2
#endif
constexpr int GT_CHUNK_KS = GT_CHUNK_KS_VARIANT;   // k-steps per A-operand chunk
constexpr int GT_KCOLS = 2 * GT_KS;         // tensor-memory columns of one k-step of A: 8 heads, then 8 tails
constexpr int GT_CH = GT_KCOLS * GT_CHUNK_KS;   // columns per chunk
constexpr int GT_NBUF = GT_NBUF_VARIANT;
constexpr int GT_COLS = 128;                // tensor-memory columns per CTA: 2 x 48 (A operand chunks) + 32 (accumulator)
constexpr int GT_DCOL = GT_NBUF * GT_CH;
constexpr int GT_THREADS = 160;             // four compute warps (one per 32 lanes of tensor memory) + the MMA warp
constexpr int GT_MAX_KS = 56;               // 56 KiB of B operand + 3.5 KiB of exponent pairs fit the 64 KiB staging slot
// instruction descriptors: F32 += TF32 x TF32, K-major A and B, M = 128, N = 32 (hi x [Bhi | Blo]) and N = 16 (lo x Bhi)
constexpr uint32_t GT_IDESC32 = (1u << 4) | (2u << 7) | (2u << 10) | ((32u >> 3) << 17) | ((128u >> 4) << 24);
constexpr uint32_t GT_IDESC16 = (1u << 4) | (2u << 7) | (2u << 10) | ((16u >> 3) << 17) | ((128u >> 4) << 24);

// B operand image: per k-step one 32-row matrix [Bhi (16 moment rows) ; Blo (16 rows)] x 8 bins in the canonical K-major
// no-swizzle layout [K half (4 bins)][row group of 8][row][4 bins] -- core matrices of 8 rows x 16 bytes; LBO (next K half)
// = 512 B, SBO (next row group) = 128 B.  An N = 32 MMA reads all of it, an N = 16 MMA on the same descriptor the Bhi rows.
// Followed by the exponent table [nks][4 bin pairs]{-mu1 s (2 bins), -mu2 s (2 bins)}: one LDS.128 feeds an FMUL2 and an FFMA2.
__global__ void gn_tc_table_kernel(const float *__restrict__ i0, const float *__restrict__ mus, int nE, int nks, float *__restrict__ tab)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nks * GT_KS) return;
    const int ks = e / GT_KS, kc = (e % GT_KS) / 4, kk = e % 4;
    float row[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) row[q] = 0.f;
    float t0 = 0.f, t1 = 0.f;
    if (e < nE) {
        const float m1 = mus[e], m2 = mus[nE + e], s1 = i0[e], s2 = i0[nE + e];
        t0 = -m1 * GnMath<float>::scale; t1 = -m2 * GnMath<float>::scale;
        const float f[6] = {1.f, m1, m2, m1 * m1, m1 * m2, m2 * m2};
#pragma unroll
        for (int q = 0; q < 6; ++q) { row[q] = s1 * f[q]; row[6 + q] = s2 * f[q]; }
    }
    float *b = tab + (size_t)ks * 256;
#pragma unroll
    for (int n = 0; n < 16; ++n) {
        const float hi = __uint_as_float((__float_as_uint(row[n]) + 0x1000u) & 0xffffe000u);      // nearest TF32 (11-bit significand)
        b[((kc * 4 + n / 8) * 8 + n % 8) * 4 + kk] = hi;
        b[((kc * 4 + 2 + n / 8) * 8 + n % 8) * 4 + kk] = row[n] - hi;                               // exact
    }
    float *x = tab + (size_t)nks * 256 + ks * 16 + ((e % GT_KS) / 2) * 4 + (e & 1);
    x[0] = t0;
    x[2] = t1;
}

// Air pixels (first measurement >= mask_thresh x its maximum) are zeroed by get_basismat_sinos whatever the iterations made of
// them (matdecomp.py:195-196,204-205): they are zeroed HERE, and only the indices of the other pixels are handed to the
// decomposition kernel (a quarter of a CT sinogram is air, in runs at both ends of every detector line).  The list keeps the
// pixels of a CTA's 1024 in order (coalescing survives); the order of the CTAs' pieces is whatever the atomics make it.
// scratch: [0] max key, [2] flagged count, [3] tile counter, [4 .. 4 + P) flagged list, [4 + P] number of listed pixels,
// [8 + P .. 8 + 2 P) the list.
__global__ void __launch_bounds__(256) gn_compact_kernel(const float *__restrict__ g1, long long P, float mask_thresh, void *scratch,
                                                         float *__restrict__ a_out, const GnMirrors mir)
{
    __shared__ unsigned warp_tot[8], base_sm;
    const float thr = mask_thresh * decode_max<float>(scratch);
    unsigned *count = (unsigned *)scratch + 4 + P, *list = (unsigned *)scratch + 8 + P;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long long blk0 = (long long)blockIdx.x * 1024; blk0 < P; blk0 += (long long)gridDim.x * 1024) {
        unsigned keep_bits[4], n_mine = 0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const long long p = blk0 + u * 256 + threadIdx.x;
            const bool keep = p < P && !(g1[p] >= thr);
            if (p < P && !keep) {
                const float2 z = make_float2(0.f, 0.f);
                ((float2 *)a_out)[p] = z;
                for (int d = 0; d < mir.n; ++d) ((float2 *)mir.p[d])[p] = z;
            }
            keep_bits[u] = __ballot_sync(0xffffffffu, keep);
            n_mine += __popc(keep_bits[u]);                  // (same in all lanes: the warp's count for this quarter)
        }
        if (lane == 0) warp_tot[warp] = n_mine;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned t = 0;
            for (int w = 0; w < 8; ++w) t += warp_tot[w];
            base_sm = atomicAdd(count, t);
        }
        __syncthreads();
        // the CTA's piece is laid out [warp][quarter][lane]: runs of up to 32 consecutive pixels
        unsigned ofs = base_sm;
        for (int w = 0; w < warp; ++w) ofs += warp_tot[w];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const long long p = blk0 + u * 256 + threadIdx.x;
            if ((keep_bits[u] >> lane) & 1u) list[ofs + __popc(keep_bits[u] & ((1u << lane) - 1u))] = (unsigned)p;
            ofs += __popc(keep_bits[u]);
        }
        __syncthreads();
    }
}

// Every pixel starts from a = (1e-6, 1e-6) (matdecomp.py:98), so the twelve energy sums of the FIRST Newton iteration are
// the same for all pixels: they are computed once here (float64, one CTA) and the kernels take their first step from them
// without evaluating a single exponential -- one iteration in n_iters of the work gone, same arithmetic otherwise.
// out[0..5] = measurement 1 {nu, sum i0 mu1 att, sum i0 mu2 att, sum i0 mu1^2 att, sum i0 mu1 mu2 att, sum i0 mu2^2 att},
// out[6..11] = measurement 2.
__global__ void __launch_bounds__(256) gn_first_sums_kernel(const float *__restrict__ i0, const float *__restrict__ mus, int nE,
                                                             float *__restrict__ out)
{
    __shared__ double part[12][256];
    double acc[12];
#pragma unroll
    for (int q = 0; q < 12; ++q) acc[q] = 0.0;
    for (int e = threadIdx.x; e < nE; e += blockDim.x) {
        const double m1 = mus[e], m2 = mus[nE + e], s1 = i0[e], s2 = i0[nE + e];
        const double att = exp(-1e-6 * (m1 + m2));
        const double f[6] = {1.0, m1, m2, m1 * m1, m1 * m2, m2 * m2};
#pragma unroll
        for (int q = 0; q < 6; ++q) { acc[q] += s1 * f[q] * att; acc[6 + q] += s2 * f[q] * att; }
    }
#pragma unroll
    for (int q = 0; q < 12; ++q) part[q][threadIdx.x] = acc[q];
    __syncthreads();
    if (threadIdx.x < 12) {
        double t = 0.0;
        for (int k = 0; k < 256; ++k) t += part[threadIdx.x][k];
        out[threadIdx.x] = (float)t;
    }
}

__device__ __forceinline__ f32x2 fmul2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 fsub2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

#ifdef GT_PROF
// Phase clocks of CTA 0 (debug build only): [0..7] compute warp 0: free-wait, k-steps, st-wait+arrive, done-wait, ld+newton,
// chunks, iterations, -; [8..15] MMA warp: full-wait, issue, commit, chunks, latency commit(done)->seen, -
__device__ unsigned long long gt_prof[16];
__device__ long long gt_prof_done_stamp;
__device__ unsigned long long gt_cta[4096 * 3];               // per CTA: globaltimer at birth, at death, SM id
#define GT_T(var) const long long var = clock64()
#define GT_ADD(slot, a, b) (gt_loc[(slot) & 7] += (unsigned long long)((b) - (a)))       // registers; flushed once at the end
#else
#define GT_T(var)
#define GT_ADD(slot, a, b)
#endif
__global__ void __launch_bounds__(GT_THREADS, 4) gn_tc_kernel(const float *__restrict__ g, const float *__restrict__ btab, int nks, long long P,
                                                           int n_iters, float mask_thresh, void *scratch, float *__restrict__ a_out,
                                                           const GnMirrors mir, const GnRider rider, int lbo_sbo_swapped, int compacted)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float *bsm = (float *)smem_raw;                           // [nks][hi 128 floats, lo 128 floats], then [nks][4] exponent float4s
    __shared__ __align__(8) unsigned long long bars[2 * GT_NBUF + 1];
    __shared__ uint32_t tmem_base_sm;
    __shared__ unsigned next_tile_sm[2];
    __shared__ int tile_air_sm[4];                            // per compute warp: all its pixels of the current tile are air
    const int tid = threadIdx.x, lane = tid & 31;
    // the warp index as a value the compiler knows to be the same in all lanes: the role branches below are then uniform
    // branches and the code inside them may use the uniform datapath
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const uint32_t bar_full = smem_u32(&bars[0]), bar_free = smem_u32(&bars[GT_NBUF]), bar_done = smem_u32(&bars[2 * GT_NBUF]);
#ifdef GT_PROF
    unsigned long long gt_loc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long long gt_birth = clock64();
    unsigned long long gt_birth_ns;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_birth_ns));
#endif

    for (int i = tid; i < nks * 68 + 3; i += GT_THREADS) ((float4 *)bsm)[i] = ((const float4 *)btab)[i];     // + the first-iteration sums
    if (tid == 0) {
        for (int b = 0; b < GT_NBUF; ++b) { mbar_init(bar_full + 8 * b, 4); mbar_init(bar_free + 8 * b, 1); }
        mbar_init(bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
#ifdef GT_PROF
        const long long ta0 = clock64();
#endif
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_sm)), "n"(GT_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
#ifdef GT_PROF
        if (lane == 0) {                                      // all CTAs: time spent waiting for tensor-memory columns
            const unsigned long long dt = (unsigned long long)(clock64() - ta0);
            atomicAdd(&gt_prof[13], dt);
            atomicMax(&gt_prof[14], dt);
            if (dt > 20000) atomicAdd(&gt_prof[15], 1ull);
        }
#endif
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // the B operand written above is read by the tensor core (async proxy)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_sm;

    // compacted: the pixels to decompose are listed in the scratch block (gn_compact_kernel); the others are zeroed already
    const unsigned *px_list = (const unsigned *)scratch + 8 + P;
    const long long n_px = compacted ? (long long)((const unsigned *)scratch)[4 + P] : P;
    const long long n_tiles = (n_px + 127) / 128;
    const int n_chunks = (nks + GT_CHUNK_KS - 1) / GT_CHUNK_KS;

    if (warp == 4) {
        // ---- MMA issuer warp: mirrors the loop structure of the compute warps, converged; one elected lane issues
        const uint64_t lbo = lbo_sbo_swapped ? 128 : 512, sbo = lbo_sbo_swapped ? 512 : 128;
        const uint64_t desc0 = ((lbo >> 4) << 16) | ((sbo >> 4) << 32) | (1ull << 46);
        const uint32_t bsm_addr = smem_u32(bsm);
        uint32_t cnt = 0, tpar = 0;
        for (long long tile = blockIdx.x; tile < n_tiles; tile = next_tile_sm[tpar], tpar ^= 1) {
            asm volatile("bar.sync 2, %0;" ::"n"(GT_THREADS) : "memory");      // the compute warps have voted on the tile
            const bool air_tile = (tile_air_sm[0] & tile_air_sm[1] & tile_air_sm[2] & tile_air_sm[3]) != 0;
            for (int it = 1; it < n_iters && !air_tile; ++it)  // (the first iteration needs no sums of its own, see below)
#pragma unroll 1
                for (int c = 0; c < n_chunks; ++c, ++cnt) {
                    const uint32_t b = cnt % GT_NBUF;
                    GT_T(m0);
                    mbar_wait(bar_full + 8 * b, (cnt / GT_NBUF) & 1);
                    tc_fence_after();
                    GT_T(m1);
                    const uint32_t a_buf = tmem + b * GT_CH;
#pragma unroll
                    for (int s = 0; s < GT_CHUNK_KS; ++s) {
                        const int ks = c * GT_CHUNK_KS + s;
                        if (ks < nks) {
                            const uint64_t db = desc0 | (uint64_t)(((bsm_addr + ks * 1024) & 0x3ffff) >> 4);
                            // columns 0..15 += hi x Bhi + lo x Bhi, columns 16..31 += hi x Blo
                            tc_mma_tf32(tmem + GT_DCOL, a_buf + s * GT_KCOLS, db, GT_IDESC32, ks > 0);
                            tc_mma_tf32(tmem + GT_DCOL, a_buf + s * GT_KCOLS + GT_KS, db, GT_IDESC16, 1);
                        }
                    }
                    GT_T(m2);
                    tc_commit(bar_free + 8 * b);                  // the chunk's A operand may be overwritten once these MMAs are done
                    if (c == n_chunks - 1) tc_commit(bar_done);   // ... and the sums of this iteration are complete
                    GT_T(m3);
                    GT_ADD(8, m0, m1); GT_ADD(9, m1, m2); GT_ADD(10, m2, m3); GT_ADD(11, 0, 1);
#ifdef GT_PROF
                    if (c == n_chunks - 1 && blockIdx.x == 0 && lane == 0) gt_prof_done_stamp = m3;
#endif
                }
            asm volatile("bar.sync 1, %0;" ::"n"(GT_THREADS) : "memory");      // the next tile of this CTA is known (see below)
        }
        __syncwarp();
    } else {
        // ---- compute warps: thread t owns pixel tile * 128 + t and lane t of tensor memory
        const float thr = mask_thresh > 0.f ? mask_thresh * decode_max<float>(scratch) : INFINITY;
        const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
        const uint32_t xsm_addr = smem_u32(bsm + (size_t)nks * 256);         // exponent pairs, four float4 (64 B) per k-step
        const float *first_sums = bsm + (size_t)nks * 272;                   // twelve sums of the first iteration
        const int rider_per_tile = n_iters > 1 ? n_iters - 1 : 1;
        const int rider_nk = rider_per_tile * (int)((n_tiles + gridDim.x - 1) / gridDim.x);
        int rider_k = 0;
        const bool ride = rider.n != 0;
        auto rider_slice = [&](int k) {
            const long long nthreads = (long long)gridDim.x * 128;
            for (long long i = (long long)blockIdx.x * 128 + tid + nthreads * k; i < rider.n16; i += nthreads * rider_nk) {
                const uint4 v = rider.src[i];
                for (int d = 0; d < rider.n; ++d) ((uint4 *)rider.dst[d])[i] = v;
            }
        };
        // Tiles are handed out dynamically (a counter in the scratch block): the CTAs of an SM progress at different rates,
        // and with a fixed share each the last of them ran alone for a quarter of the kernel (profiles/r02k: CTA deaths
        // 335 .. 466 us).  Thread 0 draws the CTA's next tile while the current one is being worked on.
        unsigned *tile_counter = (unsigned *)scratch + 3;
        uint32_t cnt = 0, done_cnt = 0, tpar = 0;
        for (long long tile = blockIdx.x; tile < n_tiles; tile = next_tile_sm[tpar], tpar ^= 1) {
            if (tid == 0) next_tile_sm[tpar] = gridDim.x + atomicAdd(tile_counter, 1u);
            const long long q = tile * 128 + tid;
            const long long p = q < n_px ? (compacted ? (long long)px_list[q] : q) : P;      // P: no pixel
            const float g1 = p < P ? g[p] : 1.f, g2 = p < P ? g[P + p] : 1.f;
            float a1 = 1e-6f, a2 = 1e-6f;                                     // matdecomp.py:98
            bool sane = true;
            // A tile whose 128 pixels are all air is not decomposed at all: get_basismat_sinos zeroes those pixels whatever the
            // iterations made of them (matdecomp.py:195-196,204-205), and in a CT sinogram they come in long runs at both ends
            // of every detector line.  (Pixels past the end of the sinogram count as air.)
            const bool air_px = p < P ? g1 >= thr : true;
            const bool air_warp = __all_sync(0xffffffffu, air_px);
            if (lane == 0) tile_air_sm[warp] = air_warp;
            asm volatile("bar.sync 2, %0;" ::"n"(GT_THREADS) : "memory");
            const bool air_tile = (tile_air_sm[0] & tile_air_sm[1] & tile_air_sm[2] & tile_air_sm[3]) != 0;
            if (n_iters > 0 && !air_tile) {
                // first iteration: a is the same for every pixel, and so are its energy sums (gn_first_sums_kernel)
                float A0[12];
#pragma unroll
                for (int q = 0; q < 12; ++q) A0[q] = first_sums[q];
                newton_update_fast(A0, g1, g2, a1, a2, sane);
            }
            for (int it = 1; it < n_iters && !air_tile; ++it) {
                if (ride && rider_k < rider_nk) rider_slice(rider_k++);       // (a CTA may draw more tiles than its even share)
                const f32x2 a1p = pack2(a1, a1), a2p = pack2(a2, a2);
#pragma unroll 1
                for (int c = 0; c < n_chunks; ++c, ++cnt) {
                    const uint32_t b = cnt % GT_NBUF;
                    GT_T(cb);
                    // A k-step: eight exponentials, split into TF32 head and tail, one 16-column store.  A full chunk is
                    // straight-line code (no branch between its k-steps), so that the table loads and exponent FMAs of the
                    // next k-step are scheduled under the MUFU latency of the current one; only the last, partly filled
                    // chunk of an energy grid takes the branchy path.
                    // A k-step: eight exponentials, each split into its TF32 head and tail -- 16 words for one 16-column store.
                    // A full chunk is straight-line code, so that the table loads and exponent FMAs of the next k-step are
                    // scheduled under the MUFU latency of the current one; only the last, partly filled chunk of an energy grid
                    // takes the branchy path.  The values are computed BEFORE waiting for the buffer: the wait (a ~100-cycle
                    // round trip to the barrier even when the buffer is free) then runs under the exponentials in flight.
                    auto kvals = [&](uint32_t xaddr, uint32_t (&v)[16]) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            float4 t;                                          // {-mu1 s (bins 2j, 2j+1), -mu2 s (bins 2j, 2j+1)}
                            asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w) : "r"(xaddr + 16 * j));
                            const f32x2 xx = ffma2(a2p, pack2(t.z, t.w), fmul2(a1p, pack2(t.x, t.y)));
                            float x0, x1;
                            unpack2(xx, x0, x1);
                            const float e0 = GnMath<float>::att(x0), e1 = GnMath<float>::att(x1);
                            v[2 * j] = __float_as_uint(e0) & 0xffffe000u;
                            v[2 * j + 1] = __float_as_uint(e1) & 0xffffe000u;
                            const f32x2 l2 = fsub2(pack2(e0, e1), pack2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1])));
                            float l0, l1;
                            unpack2(l2, l0, l1);
                            v[8 + 2 * j] = __float_as_uint(l0);
                            v[8 + 2 * j + 1] = __float_as_uint(l1);
                        }
                    };
                    const uint32_t xa = xsm_addr + (uint32_t)c * (GT_CHUNK_KS * 64);
                    const uint32_t col = lane_base + b * GT_CH;
                    if ((c + 1) * GT_CHUNK_KS <= nks) {
                        uint32_t v[GT_CHUNK_KS][16];
#pragma unroll
                        for (int s = 0; s < GT_CHUNK_KS; ++s) kvals(xa + 64 * s, v[s]);
                        GT_T(c0);
                        mbar_wait(bar_free + 8 * b, ((cnt / GT_NBUF) & 1) ^ 1);
                        tc_fence_after();
                        GT_T(c1);
                        GT_ADD(0, c0, c1);
#pragma unroll
                        for (int s = 0; s < GT_CHUNK_KS; ++s) tmem_st16(col + s * GT_KCOLS, v[s]);
                    } else {
                        mbar_wait(bar_free + 8 * b, ((cnt / GT_NBUF) & 1) ^ 1);
                        tc_fence_after();
                        for (int s = 0; c * GT_CHUNK_KS + s < nks; ++s) {
                            uint32_t v[16];
                            kvals(xa + 64 * s, v);
                            tmem_st16(col + s * GT_KCOLS, v);
                        }
                    }
                    GT_T(c2);
                    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_full + 8 * b);
                    GT_T(c3);
                    GT_ADD(1, cb, c2); GT_ADD(2, c2, c3); GT_ADD(5, 0, 1);
                }
                GT_T(d0);
                mbar_wait(bar_done, done_cnt++ & 1);
                tc_fence_after();
                GT_T(d1);
                float D[32], A[12];
                tmem_ld32(lane_base + GT_DCOL, D);
                tc_fence_before();
#pragma unroll
                for (int q = 0; q < 12; ++q) A[q] = D[q] + D[16 + q];
                newton_update_fast(A, g1, g2, a1, a2, sane);
#ifdef GT_PROF
                const long long d2 = clock64() + (long long)(__float_as_int(a1) & 0);      // after the update
                GT_ADD(3, d0, d1); GT_ADD(4, d1, d2); GT_ADD(6, 0, 1); GT_ADD(7, gt_prof_done_stamp, d1);
#endif
            }
            if (p < P) {
                const bool masked = g1 >= thr;                                    // matdecomp.py:196,204-205
                const float2 o = make_float2(masked ? 0.f : a1, masked ? 0.f : a2);
                ((float2 *)a_out)[p] = o;
                for (int d = 0; d < mir.n; ++d) ((float2 *)mir.p[d])[p] = o;
                const bool finite = fabsf(a1) < INFINITY && fabsf(a2) < INFINITY;
                if (!masked && !(sane && finite)) {                               // rare: hand the pixel to gn_fixup_kernel
                    unsigned *fl = (unsigned *)scratch;
                    fl[4 + atomicAdd(fl + 2, 1u)] = (unsigned)p;
                }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(GT_THREADS) : "memory");
        }
        if (ride)
            for (; rider_k < rider_nk; ++rider_k) rider_slice(rider_k);
    }
#ifdef GT_PROF
    if (warp == 4) gt_loc[4] = (unsigned long long)(clock64() - gt_birth);          // life of the CTA
    if (tid == 0 && blockIdx.x < 4096) {
        unsigned long long t, sm;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        asm volatile("{ .reg .u32 r; mov.u32 r, %%smid; cvt.u64.u32 %0, r; }" : "=l"(sm));
        gt_cta[3 * blockIdx.x] = gt_birth_ns; gt_cta[3 * blockIdx.x + 1] = t; gt_cta[3 * blockIdx.x + 2] = sm;
    }
    if (blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == 4))
        for (int q = 0; q < (warp == 4 ? 5 : 8); ++q) gt_prof[(warp == 4 ? 8 : 0) + q] += gt_loc[q];
#endif
    tc_fence_before();
    __syncthreads();
    if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(GT_COLS) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------------
// General form of optimize_sino_cpu (matdecomp.py:87-127 as written: any number of measurements and basis materials, and
// a spectrum PER DETECTOR BIN, i0 [nMeas][nBins][nE] -- :101-105,118 broadcast a per-bin i0, :95,125 invert an
// [nBins, nMats, nMats] stack).  get_basismat_sinos never needs it (two measurements, two materials, one spectrum tiled
// over the bins, :151) and takes the specialised kernels above; this one keeps the drop-in's signature honest.  One thread
// per pixel, arithmetic in float64 whatever the I/O type, the reference's exponent clip, Gaussian elimination with
// partial pivoting for the nMats x nMats Newton system.
constexpr int GG_MAX = 4;
template <typename R>
__global__ void __launch_bounds__(128) gn_general_kernel(const R *__restrict__ g, const R *__restrict__ i0, const R *__restrict__ mus,
                                                         int nMeas, int nMats, int nE, long long n_bins, long long i0_bin_stride,
                                                         long long P, int n_iters, R *__restrict__ a_out)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *mu_s = (double *)smem_raw;                        // [nMats][nE]
    for (int t = threadIdx.x; t < nMats * nE; t += blockDim.x) mu_s[t] = (double)mus[t];
    __syncthreads();
    const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const long long bin = p % n_bins;                          // sinogram pixels are [views][bins]
    double a[GG_MAX], gm[GG_MAX];
    for (int q = 0; q < GG_MAX; ++q) a[q] = 1e-6;              // matdecomp.py:98
    for (int m = 0; m < nMeas; ++m) gm[m] = (double)g[(long long)m * P + p];
    for (int it = 0; it < n_iters; ++it) {
        double nu[GG_MAX], gr[GG_MAX][GG_MAX], hs[GG_MAX][GG_MAX][GG_MAX];
        for (int m = 0; m < GG_MAX; ++m) {
            nu[m] = 0.0;
            for (int u = 0; u < GG_MAX; ++u) {
                gr[m][u] = 0.0;
                for (int w = 0; w < GG_MAX; ++w) hs[m][u][w] = 0.0;
            }
        }
        for (int e = 0; e < nE; ++e) {
            double x = 0.0;
            for (int u = 0; u < nMats; ++u) x -= a[u] * mu_s[u * nE + e];
            const double att = exp(fmin(fmax(x, -700.0), 700.0));                                   // :116
            for (int m = 0; m < nMeas; ++m) {
                const double w = (double)i0[((long long)m * n_bins + bin) * i0_bin_stride + e] * att;
                nu[m] += w;                                                                         // :118
                for (int u = 0; u < nMats; ++u) {
                    const double wu = w * mu_s[u * nE + e];
                    gr[m][u] += wu;                                                                 // -nu_grad, :119
                    for (int v = u; v < nMats; ++v) hs[m][u][v] += wu * mu_s[v * nE + e];           // :120
                }
            }
        }
        // dF_u = -sum_m r_m nu_grad_mu = sum_m r_m gr_mu;   H_uv = -sum_m (r_m hs_muv - g_m/nu_m^2 gr_mu gr_mv)    (:122-123)
        double H[GG_MAX][GG_MAX + 1];
        for (int u = 0; u < nMats; ++u)
            for (int v = 0; v <= nMats; ++v) H[u][v] = 0.0;
        for (int m = 0; m < nMeas; ++m) {
            const double r = gm[m] / nu[m] - 1.0, s2 = gm[m] / (nu[m] * nu[m]);
            for (int u = 0; u < nMats; ++u) {
                H[u][nMats] += r * gr[m][u];
                for (int v = 0; v < nMats; ++v) {
                    const double h = u <= v ? hs[m][u][v] : hs[m][v][u];
                    H[u][v] += -(r * h) + s2 * gr[m][u] * gr[m][v];
                }
            }
        }
        // solve H d = dF (augmented column), partial pivoting; a -= d   (:125)
        for (int c = 0; c < nMats; ++c) {
            int piv = c;
            for (int u = c + 1; u < nMats; ++u)
                if (fabs(H[u][c]) > fabs(H[piv][c])) piv = u;
            if (piv != c)
                for (int v = 0; v <= nMats; ++v) { const double t = H[c][v]; H[c][v] = H[piv][v]; H[piv][v] = t; }
            const double inv = 1.0 / H[c][c];
            for (int u = c + 1; u < nMats; ++u) {
                const double f = H[u][c] * inv;
                for (int v = c; v <= nMats; ++v) H[u][v] -= f * H[c][v];
            }
        }
        for (int c = nMats - 1; c >= 0; --c) {
            double d = H[c][nMats];
            for (int v = c + 1; v < nMats; ++v) d -= H[c][v] * H[v][nMats];
            H[c][nMats] = d / H[c][c];
        }
        for (int u = 0; u < nMats; ++u) a[u] -= H[u][nMats];
    }
    for (int u = 0; u < nMats; ++u) a_out[p * nMats + u] = (R)a[u];
}

// Float64 pass over the pixels the single-precision kernel flagged (scratch: [2] = count, [4 ...] = pixel ids): the same
// Newton iteration in double precision with the reference's exponent clip, on the same (float) inputs.  These are
// photon-starved or metal-crossing pixels whose iterates leave the float range; the reference's float64 result for them is
// finite, and a NaN left in a sinogram would spoil every pixel of the image reconstructed from it.
__global__ void __launch_bounds__(128) gn_fixup_kernel(const float *__restrict__ g, const float *__restrict__ i0,
                                                       const float *__restrict__ mus, int nE, long long P, int n_iters,
                                                       const unsigned *__restrict__ flagged, float *__restrict__ a_out,
                                                       const GnMirrors mir)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *tab = (double *)smem_raw;                                    // [nE][14]
    const unsigned count = flagged[2];
    if (count == 0) return;
    for (int e = threadIdx.x; e < nE; e += blockDim.x) {
        const double m1 = mus[e], m2 = mus[nE + e], s1 = i0[e], s2 = i0[nE + e];
        double *row = tab + e * 14;
        row[0] = -m1; row[1] = -m2;
        const double f[6] = {1.0, m1, m2, m1 * m1, m1 * m2, m2 * m2};
#pragma unroll
        for (int q = 0; q < 6; ++q) { row[2 + q] = s1 * f[q]; row[8 + q] = s2 * f[q]; }
    }
    __syncthreads();
    // one WARP per flagged pixel, the energy bins dealt round-robin to its lanes: a single thread would need ~0.2 ms for
    // one pixel of a C2 scan, and one flagged pixel is enough to put that latency on the critical path
    const int lane = threadIdx.x & 31;
    const unsigned warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    for (unsigned t = warp; t < count; t += n_warps) {
        const long long p = flagged[4 + t];
        const double g1 = g[p], g2 = g[P + p];
        double a1 = 1e-6, a2 = 1e-6;
        for (int it = 0; it < n_iters; ++it) {
            double A[12];
#pragma unroll
            for (int q = 0; q < 12; ++q) A[q] = 0.0;
            for (int e = lane; e < nE; e += 32) {
                const double *row = tab + e * 14;
                const double att = GnMath<double>::att(fma(a2, row[1], a1 * row[0]));
#pragma unroll
                for (int q = 0; q < 12; ++q) A[q] = fma(row[2 + q], att, A[q]);
            }
            // xor butterfly: x + y is commutative, so every lane ends with bitwise the same twelve sums
#pragma unroll
            for (int q = 0; q < 12; ++q)
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) A[q] += __shfl_xor_sync(0xffffffffu, A[q], o);
            const double nu1 = A[0], nu2 = A[6];
            const double r1 = g1 / nu1 - 1.0, r2 = g2 / nu2 - 1.0;
            const double s1 = g1 / (nu1 * nu1), s2 = g2 / (nu2 * nu2);
            const double d1 = r1 * A[1] + r2 * A[7];
            const double d2 = r1 * A[2] + r2 * A[8];
            const double h11 = -(r1 * A[3] + r2 * A[9]) + (s1 * A[1] * A[1] + s2 * A[7] * A[7]);
            const double h12 = -(r1 * A[4] + r2 * A[10]) + (s1 * A[1] * A[2] + s2 * A[7] * A[8]);
            const double h22 = -(r1 * A[5] + r2 * A[11]) + (s1 * A[2] * A[2] + s2 * A[8] * A[8]);
            const double w = h12 * h12;
            const double err = fma(-h12, h12, w);
            const double det = fma(h11, h22, -w) + err;
            const double inv = 1.0 / det;
            a1 -= (h22 * d1 - h12 * d2) * inv;
            a2 -= (h11 * d2 - h12 * d1) * inv;
        }
        if (lane == 0) {
            const float2 o = make_float2((float)a1, (float)a2);
            ((float2 *)a_out)[p] = o;
            for (int d = 0; d < mir.n; ++d) ((float2 *)mir.p[d])[p] = o;
        }
    }
}

template <typename R>
static int launch_max(const void *g, long long P, void *scratch, cudaStream_t st)
{
    DEXCT_CUDA_OK(cudaMemsetAsync(scratch, 0x80, 8, st));      // key of a very negative value for both key widths
    if (P > 0) {
        gn_max_kernel<R><<<(unsigned)min((long long)sm_count() * 4, (P + 255) / 256), 256, 0, st>>>((const R *)g, P, (R *)scratch);
        DEXCT_CUDA_OK(cudaGetLastError());
    }
    return DEXCT_OK;
}

template <typename R, int PX, int VAR = 0>
static int launch_gn(const void *g, const void *i0, const void *mus, int nE, long long P, int n_iters, double mask_thresh,
                     void *a_out, void *scratch, bool have_max, cudaStream_t st, const GnMirrors &mir, const GnRider &rider)
{
    const char *bs = getenv("DEXCT_GN_BLOCK");         // tuning knob (threads per CTA: 128 or 256)
    const int block = (bs && atoi(bs) == 128) ? 128 : 256;
    const bool ctab = sizeof(R) == 4 && nE + 2 <= GN_CTAB_ROWS;
    constexpr bool UNI = VAR >= 1;
    const size_t smem = ctab ? 0 : sizeof(R) * (size_t)(nE + 2) * GN_ROW;
    if (smem > 200 * 1024) {
        set_error("GN table needs %zu bytes of shared memory (limit 200 KiB): nE=%d", smem, nE);
        return DEXCT_ECAPACITY;
    }
    void (*kern)(const R *, const R *, const R *, int, long long, int, R, void *, R *, const GnMirrors, const GnRider) = gn_kernel<R, PX, false, 0>;
    if (ctab) kern = gn_kernel<R, PX, sizeof(R) == 4, VAR>;
    if (smem > 48 * 1024) DEXCT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    DEXCT_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, block, smem));
    if (per_sm < 1) per_sm = 1;
    const long long want = (P + (long long)block * PX - 1) / ((long long)block * PX);
    long long grid = (long long)sm_count() * per_sm;          // one resident wave; the grid-stride loop does the rest
    if (grid > want) grid = want;
    if (mask_thresh > 0 && !have_max) {
        if (int e = launch_max<R>(g, P, scratch, st)) return e;
    }
    if (sizeof(R) == 4) DEXCT_CUDA_OK(cudaMemsetAsync((char *)scratch + 8, 0, 8, st));     // count of flagged pixels
    // float64 pass over whatever the single-precision kernel flags (usually nothing: the kernel then returns at once)
    auto fixup = [&]() -> int {
        const size_t fsm = sizeof(double) * 14 * (size_t)nE;
        if (fsm > 200 * 1024) {
            set_error("GN table needs %zu bytes of shared memory (limit 200 KiB): nE=%d", fsm, nE);
            return DEXCT_ECAPACITY;
        }
        if (fsm > 48 * 1024)
            DEXCT_CUDA_OK(cudaFuncSetAttribute(gn_fixup_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsm));
        gn_fixup_kernel<<<sm_count() * 4, 128, fsm, st>>>((const float *)g, (const float *)i0, (const float *)mus, nE, P,
                                                          n_iters, (const unsigned *)scratch, (float *)a_out, mir);
        DEXCT_CUDA_OK(cudaGetLastError());
        return DEXCT_OK;
    };
    if (ctab) {
        ConstSlot slot(SLOT_GN, st);                          // exclusive use of gn_ctab until the kernel below has run
        if (!slot.ok()) return DEXCT_ECUDA;
        if (UNI) {
            gn_table_unified_kernel<<<(nE + 4 + 127) / 128, 128, 0, st>>>((const float *)i0, (const float *)mus, nE, (float *)slot.staging());
            DEXCT_CUDA_OK(cudaGetLastError());
            DEXCT_CUDA_OK(cudaMemcpyToSymbolAsync(gn_crow, slot.staging(), sizeof(float) * (size_t)(nE + 4) * GN_UROW, 0,
                                                  cudaMemcpyDeviceToDevice, st));
        } else {
        gn_table_kernel<<<(nE + 2 + 127) / 128, 128, 0, st>>>((const float *)i0, (const float *)mus, nE, (float *)slot.staging());
        DEXCT_CUDA_OK(cudaGetLastError());
        DEXCT_CUDA_OK(cudaMemcpyToSymbolAsync(gn_ctab, slot.staging(), sizeof(float) * (size_t)(nE + 2) * GN_CROW, 0,
                                              cudaMemcpyDeviceToDevice, st));
        DEXCT_CUDA_OK(cudaMemcpyToSymbolAsync(gn_cexp, (const float *)slot.staging() + (size_t)(nE + 2) * GN_CROW,
                                              sizeof(float) * (size_t)(nE + 2) * 2, 0, cudaMemcpyDeviceToDevice, st));
        }
        kern<<<(unsigned)grid, block, smem, st>>>((const R *)g, (const R *)i0, (const R *)mus, nE, P, n_iters, (R)mask_thresh,
                                                 scratch, (R *)a_out, mir, rider);
        DEXCT_CUDA_OK(cudaGetLastError());
        return fixup();
    }
    kern<<<(unsigned)grid, block, smem, st>>>((const R *)g, (const R *)i0, (const R *)mus, nE, P, n_iters, (R)mask_thresh,
                                             scratch, (R *)a_out, mir, rider);
    DEXCT_CUDA_OK(cudaGetLastError());
    return sizeof(R) == 4 ? fixup() : DEXCT_OK;
}

// Single precision through the tensor-core kernel (nE <= 8 GT_MAX_KS).  Same contract as launch_gn<float, .>.
static int launch_gn_tc(const void *g, const void *i0, const void *mus, int nE, long long P, int n_iters, double mask_thresh,
                        void *a_out, void *scratch, bool have_max, cudaStream_t st, const GnMirrors &mir, const GnRider &rider)
{
    const int nks = (nE + GT_KS - 1) / GT_KS;
    const size_t smem = (size_t)nks * (1024 + 64) + 48;     // B operand + exponent pairs + first-iteration sums
    DEXCT_CUDA_OK(cudaFuncSetAttribute(gn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // CTAs per SM: 512 columns of tensor memory / GT_COLS, and what fits the shared memory.  (The occupancy calculator
    // answers 1 for a kernel that allocates tensor memory -- profiles/r02c: a grid of 148 CTAs ran at 3x the time.)
    int per_sm = 512 / GT_COLS;
    while (per_sm > 1 && (size_t)per_sm * (smem + 2048) > 220 * 1024) --per_sm;
    const char *ps = getenv("DEXCT_GN_TC_CTAS");             // tuning knob
    if (ps && atoi(ps) >= 1 && atoi(ps) <= 512 / GT_COLS) per_sm = atoi(ps);
    const long long n_tiles = (P + 127) / 128;
    long long grid = (long long)sm_count() * per_sm;           // one resident wave; the tile-stride loop does the rest
    if (grid > n_tiles) grid = n_tiles;
    if (mask_thresh > 0 && !have_max) {
        if (int e = launch_max<float>(g, P, scratch, st)) return e;
    }
    DEXCT_CUDA_OK(cudaMemsetAsync((char *)scratch + 8, 0, 8, st));                              // count of flagged pixels, tile counter
    const char *sw = getenv("DEXCT_GN_TC_SWAP");                // debugging knob: LBO/SBO of the B descriptor exchanged
    {
        ConstSlot slot(SLOT_GN, st);                          // exclusive use of the staging image until the kernel has run
        if (!slot.ok()) return DEXCT_ECUDA;
        gn_tc_table_kernel<<<(nks * GT_KS + 127) / 128, 128, 0, st>>>((const float *)i0, (const float *)mus, nE, nks, (float *)slot.staging());
        DEXCT_CUDA_OK(cudaGetLastError());
        gn_first_sums_kernel<<<1, 256, 0, st>>>((const float *)i0, (const float *)mus, nE, (float *)slot.staging() + (size_t)nks * 272);
        DEXCT_CUDA_OK(cudaGetLastError());
        // with an air mask: zero the air pixels now and decompose only the others (gn_compact_kernel)
        const char *cp = getenv("DEXCT_GN_COMPACT");             // 0: decompose every tile that holds a non-air pixel, as before
        const int compacted = mask_thresh > 0 && !(cp && atoi(cp) == 0);
        if (compacted) {
            DEXCT_CUDA_OK(cudaMemsetAsync((unsigned *)scratch + 4 + P, 0, sizeof(unsigned), st));
            const long long nb = (P + 1023) / 1024;
            gn_compact_kernel<<<(unsigned)std::min<long long>(nb, (long long)sm_count() * 8), 256, 0, st>>>((const float *)g, P, (float)mask_thresh,
                                                                                                         scratch, (float *)a_out, mir);
            DEXCT_CUDA_OK(cudaGetLastError());
        }
        gn_tc_kernel<<<(unsigned)grid, GT_THREADS, smem, st>>>((const float *)g, (const float *)slot.staging(), nks, P, n_iters,
                                                              (float)mask_thresh, scratch, (float *)a_out, mir, rider, sw && atoi(sw), compacted);
        DEXCT_CUDA_OK(cudaGetLastError());
    }
    const size_t fsm = sizeof(double) * 14 * (size_t)nE;
    if (fsm > 48 * 1024) DEXCT_CUDA_OK(cudaFuncSetAttribute(gn_fixup_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsm));
    gn_fixup_kernel<<<sm_count() * 4, 128, fsm, st>>>((const float *)g, (const float *)i0, (const float *)mus, nE, P, n_iters,
                                                      (const unsigned *)scratch, (float *)a_out, mir);
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}

}  // namespace dexct

using namespace dexct;

#ifdef GT_PROF
extern "C" int dexct_debug_gn_tc_ctas(unsigned long long *out, int n)
{
    return cudaMemcpyFromSymbol(out, gt_cta, sizeof(unsigned long long) * 3 * (size_t)n) == cudaSuccess ? DEXCT_OK : DEXCT_ECUDA;
}
extern "C" int dexct_debug_gn_tc_prof(unsigned long long *out16, int reset)
{
    unsigned long long z[16] = {0};
    if (cudaMemcpyFromSymbol(out16, gt_prof, sizeof(z)) != cudaSuccess) return DEXCT_ECUDA;
    if (reset && cudaMemcpyToSymbol(gt_prof, z, sizeof(z)) != cudaSuccess) return DEXCT_ECUDA;
    return DEXCT_OK;
}
#endif

extern "C" size_t dexct_gn_scratch_bytes(int64_t P, int32_t dtype)
{
    if (P < 0) return 0;
    return dtype == DEXCT_F32 ? 32 + 2 * sizeof(unsigned) * (size_t)P : 16;    // keys + counters, flagged list, list of non-air pixels
}

extern "C" int dexct_gn_air_max(const void *g1, int64_t P, void *scratch, int32_t dtype, void *stream)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(scratch && (g1 || P == 0) && P >= 0, "g1 and scratch must be non-null, P >= 0");
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    cudaStream_t st = (cudaStream_t)stream;
    return dtype == DEXCT_F32 ? launch_max<float>(g1, P, scratch, st) : launch_max<double>(g1, P, scratch, st);
}

static int gn_decompose_impl(const void *g, const void *i0, const void *mus, int32_t nE, int64_t P, int32_t n_iters,
                             double mask_thresh, void *a_out, const GnMirrors &mir, const GnRider &rider, void *scratch,
                             int32_t scratch_has_max, int32_t dtype, void *stream)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(nE >= 1 && P >= 0 && n_iters >= 0, "need nE >= 1, P >= 0, n_iters >= 0");
    DEXCT_REQUIRE(P == 0 || (g && i0 && mus && a_out), "g, i0, mus and a_out must be non-null");
    DEXCT_REQUIRE(scratch || (mask_thresh <= 0 && dtype == DEXCT_F64),
                  "scratch (dexct_gn_scratch_bytes) must be non-null in single precision or when mask_thresh > 0");
    DEXCT_REQUIRE(P < (1ll << 32), "at most 2^32 - 1 pixels per call");
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    if (P == 0) return DEXCT_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const bool have = scratch_has_max != 0;
    if (dtype == DEXCT_F32) {
        // Default: the tensor-core kernel (tcgen05 + tensor memory), 1.44x the FP32-pipe kernel at C2 size; energy grids
        // beyond 448 bins (the B operand would not fit its staging slot) and DEXCT_GN_TC=0 take the FP32-pipe kernel.
        const char *tc = getenv("DEXCT_GN_TC");
        if ((!tc || atoi(tc)) && (nE + GT_KS - 1) / GT_KS <= GT_MAX_KS)
            return launch_gn_tc(g, i0, mus, nE, P, n_iters, mask_thresh, a_out, scratch, have, st, mir, rider);
        const char *px = getenv("DEXCT_GN_PX");          // tuning knob (pixels per thread)
        const int npx = px ? atoi(px) : 4;            // 4: the table row of a bin is fetched once for four pixels
        if (npx == 1) return launch_gn<float, 1>(g, i0, mus, nE, P, n_iters, mask_thresh, a_out, scratch, have, st, mir, rider);
        if (npx == 3) return launch_gn<float, 3>(g, i0, mus, nE, P, n_iters, mask_thresh, a_out, scratch, have, st, mir, rider);
        if (npx == 2) return launch_gn<float, 2>(g, i0, mus, nE, P, n_iters, mask_thresh, a_out, scratch, have, st, mir, rider);
        const char *vs = getenv("DEXCT_GN_VAR");         // tuning knob (loop variant)
        const int var = vs ? atoi(vs) : 0;
        if (var == 1) return launch_gn<float, 4, 1>(g, i0, mus, nE, P, n_iters, mask_thresh, a_out, scratch, have, st, mir, rider);
        if (var == 2) return launch_gn<float, 4, 2>(g, i0, mus, nE, P, n_iters, mask_thresh, a_out, scratch, have, st, mir, rider);
        return launch_gn<float, 4>(g, i0, mus, nE, P, n_iters, mask_thresh, a_out, scratch, have, st, mir, rider);
    }
    return launch_gn<double, 1>(g, i0, mus, nE, P, n_iters, mask_thresh, a_out, scratch, have, st, mir, rider);
}

extern "C" int dexct_gn_decompose(const void *g, const void *i0, const void *mus, int32_t nE, int64_t P, int32_t n_iters,
                                  double mask_thresh, void *a_out, void *scratch, int32_t scratch_has_max,
                                  int32_t dtype, void *stream)
{
    GnMirrors mir{};
    GnRider rider{};
    return gn_decompose_impl(g, i0, mus, nE, P, n_iters, mask_thresh, a_out, mir, rider, scratch, scratch_has_max, dtype, stream);
}

static int fill_mirrors(GnMirrors &mir, void *const *mirrors_host, int32_t n_mirrors)
{
    DEXCT_REQUIRE(n_mirrors >= 0 && n_mirrors <= DEXCT_MAX_PEERS && (n_mirrors == 0 || mirrors_host),
                  "0 <= n_mirrors <= DEXCT_MAX_PEERS and a pointer table required");
    mir.n = n_mirrors;
    for (int i = 0; i < n_mirrors; ++i) {
        DEXCT_REQUIRE(mirrors_host[i] && ((uintptr_t)mirrors_host[i] & 15) == 0, "mirror %d is null or not 16-byte aligned", i);
        mir.p[i] = mirrors_host[i];
    }
    return DEXCT_OK;
}

extern "C" int dexct_gn_decompose_mirrored(const void *g, const void *i0, const void *mus, int32_t nE, int64_t P,
                                           int32_t n_iters, double mask_thresh, void *a_out, void *const *mirrors_host,
                                           int32_t n_mirrors, void *scratch, int32_t scratch_has_max, int32_t dtype,
                                           void *stream)
{
    GnMirrors mir{};
    GnRider rider{};
    if (int e = fill_mirrors(mir, mirrors_host, n_mirrors)) return e;
    return gn_decompose_impl(g, i0, mus, nE, P, n_iters, mask_thresh, a_out, mir, rider, scratch, scratch_has_max, dtype, stream);
}

extern "C" int dexct_gn_decompose_fused_push(const void *g, const void *i0, const void *mus, int32_t nE, int64_t P,
                                             int32_t n_iters, double mask_thresh, void *a_out, void *const *mirrors_host,
                                             int32_t n_mirrors, const void *push_src, size_t push_bytes,
                                             void *const *push_dst_host, int32_t n_push_dst, void *scratch,
                                             int32_t scratch_has_max, int32_t dtype, void *stream)
{
    GnMirrors mir{};
    GnRider rider{};
    if (int e = fill_mirrors(mir, mirrors_host, n_mirrors)) return e;
    DEXCT_REQUIRE(n_push_dst >= 0 && n_push_dst <= DEXCT_MAX_PEERS, "0 <= n_push_dst <= DEXCT_MAX_PEERS");
    if (n_push_dst > 0 && push_bytes > 0) {
        DEXCT_REQUIRE(P > 0, "the push rides on the decomposition kernel: P must be positive");
        DEXCT_REQUIRE(push_src && push_dst_host && push_bytes % 16 == 0 && ((uintptr_t)push_src & 15) == 0,
                      "push_src must be non-null and 16-byte aligned, push_bytes a multiple of 16");
        rider.src = (const uint4 *)push_src;
        rider.n16 = (long long)(push_bytes / 16);
        rider.n = n_push_dst;
        for (int i = 0; i < n_push_dst; ++i) {
            DEXCT_REQUIRE(push_dst_host[i] && ((uintptr_t)push_dst_host[i] & 15) == 0, "push destination %d is null or not 16-byte aligned", i);
            rider.dst[i] = push_dst_host[i];
        }
    }
    return gn_decompose_impl(g, i0, mus, nE, P, n_iters, mask_thresh, a_out, mir, rider, scratch, scratch_has_max, dtype, stream);
}

extern "C" int dexct_gn_decompose_general(const void *g, const void *i0, const void *mus, int32_t nMeas, int32_t nMats, int32_t nE,
                                          int64_t n_bins, int32_t i0_per_bin, int64_t P, int32_t n_iters, void *a_out, int32_t dtype,
                                          void *stream)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(nMeas >= 1 && nMeas <= GG_MAX && nMats >= 1 && nMats <= GG_MAX,
                  "1 to %d measurements and basis materials are supported (nMeas=%d nMats=%d)", GG_MAX, nMeas, nMats);
    DEXCT_REQUIRE(nE >= 1 && P >= 0 && n_iters >= 0 && n_bins >= 1, "need nE >= 1, P >= 0, n_iters >= 0, n_bins >= 1");
    DEXCT_REQUIRE(P % n_bins == 0, "P = %lld pixels is not a whole number of views of %lld bins", (long long)P, (long long)n_bins);
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    if (P == 0) return DEXCT_OK;
    DEXCT_REQUIRE(g && i0 && mus && a_out, "g, i0, mus and a_out must be non-null");
    const size_t smem = sizeof(double) * (size_t)nMats * nE;
    if (smem > 200 * 1024) {
        set_error("basis table needs %zu bytes of shared memory (limit 200 KiB): nMats=%d nE=%d", smem, nMats, nE);
        return DEXCT_ECAPACITY;
    }
    cudaStream_t st = (cudaStream_t)stream;
    // one spectrum for all bins: the bin index is folded away by a zero "bins" extent
    const long long pix_bins = i0_per_bin ? n_bins : 1;
    const unsigned grid = (unsigned)((P + 127) / 128);
    if (dtype == DEXCT_F32) {
        if (smem > 48 * 1024) DEXCT_CUDA_OK(cudaFuncSetAttribute(gn_general_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        gn_general_kernel<float><<<grid, 128, smem, st>>>((const float *)g, (const float *)i0, (const float *)mus, nMeas, nMats, nE,
                                                        pix_bins, (long long)nE, P, n_iters, (float *)a_out);
    } else {
        if (smem > 48 * 1024) DEXCT_CUDA_OK(cudaFuncSetAttribute(gn_general_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        gn_general_kernel<double><<<grid, 128, smem, st>>>((const double *)g, (const double *)i0, (const double *)mus, nMeas, nMats, nE,
                                                         pix_bins, (long long)nE, P, n_iters, (double *)a_out);
    }
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// dexct_gn_decompose_static: the single-precision decomposition for callers that may not allocate (code under CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same kernels as the FP32-pipe path of dexct_gn_decompose with the table row staged in shared memory
// (gn_max_kernel, gn_kernel<float, 4, false, 0>, gn_fixup_kernel) -- that path needs no staging buffer -- and the scratch
// block (air-maximum key, flagged-pixel count and list: 16 + 4 P bytes) in a file-scope device array instead of caller
// memory.  Every GPU that loads the library holds the array (4 MiB + 256 B) until exit.  All calls share it and it keeps the
// last call's contents: the key and the counters are reset at the start of every call; calls that overlap on one GPU (two
// streams, two threads) race on it and must be serialised by the caller.
namespace dexct {
constexpr long long GN_STATIC_MAX_PIXELS = DEXCT_GN_STATIC_MAX_PIXELS;
static __device__ __align__(256) unsigned char gn_static_scratch[256 + 4 * GN_STATIC_MAX_PIXELS];
namespace {
std::mutex g_static_mtx;
void *g_static_ptr[64] = {};          // device address of gn_static_scratch, per device (the symbol itself is a host stand-in)
}  // namespace

// Device address of the array on the CURRENT device, looked up once per device with that device current.
static int gn_static_address(void **out)
{
    int dev = -1;
    DEXCT_CUDA_OK(cudaGetDevice(&dev));
    DEXCT_REQUIRE(dev >= 0 && dev < 64, "device index out of range");
    std::lock_guard<std::mutex> lock(g_static_mtx);
    if (!g_static_ptr[dev]) {
        void *ptr = nullptr;
        DEXCT_CUDA_OK(cudaGetSymbolAddress((void **)&ptr, gn_static_scratch));
        DEXCT_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned a null address for the static scratch array");
        g_static_ptr[dev] = ptr;
    }
    *out = g_static_ptr[dev];
    return DEXCT_OK;
}
}  // namespace dexct

extern "C" int64_t dexct_gn_static_max_pixels(void) { return dexct::GN_STATIC_MAX_PIXELS; }

extern "C" int dexct_gn_decompose_static(const void *g, const void *i0, const void *mus, int32_t nE, int64_t P, int32_t n_iters,
                                         double mask_thresh, void *a_out, void *stream)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(g && i0 && mus && a_out, "g, i0, mus and a_out must be non-null");
    DEXCT_REQUIRE(nE > 0 && P >= 0 && n_iters >= 0, "nE > 0, P >= 0, n_iters >= 0 required");
    if (P > GN_STATIC_MAX_PIXELS) {
        set_error("dexct_gn_decompose_static: %lld pixels exceed the static scratch array (DEXCT_GN_STATIC_MAX_PIXELS = %lld): "
                  "split the call or use dexct_gn_decompose with caller-provided scratch", (long long)P, GN_STATIC_MAX_PIXELS);
        return DEXCT_ECAPACITY;
    }
    if (P == 0) return DEXCT_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = sizeof(float) * (size_t)(nE + 2) * GN_ROW, fsm = sizeof(double) * 14 * (size_t)nE;
    if (smem > 200 * 1024 || fsm > 200 * 1024) {
        set_error("GN table needs %zu bytes of shared memory (limit 200 KiB): nE=%d", smem > fsm ? smem : fsm, nE);
        return DEXCT_ECAPACITY;
    }
    void *scratch = nullptr;
    if (int e = gn_static_address(&scratch)) return e;
    auto kern = gn_kernel<float, 4, false, 0>;
    if (smem > 48 * 1024) DEXCT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (fsm > 48 * 1024) DEXCT_CUDA_OK(cudaFuncSetAttribute(gn_fixup_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsm));
    int per_sm = 0;
    DEXCT_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, smem));
    if (per_sm < 1) per_sm = 1;
    const long long want = (P + 1023) / 1024;
    const long long grid = std::min<long long>((long long)sm_count() * per_sm, want);
    // what must start from a known state: the air-maximum key (launch_max resets it) and the counters
    if (mask_thresh > 0) {
        if (int e = launch_max<float>(g, P, scratch, st)) return e;
    } else {
        DEXCT_CUDA_OK(cudaMemsetAsync(scratch, 0x80, 8, st));
    }
    DEXCT_CUDA_OK(cudaMemsetAsync((char *)scratch + 8, 0, 8, st));
    const GnMirrors mir{};
    const GnRider rider{};
    kern<<<(unsigned)grid, 256, smem, st>>>((const float *)g, (const float *)i0, (const float *)mus, nE, P, n_iters, (float)mask_thresh,
                                            scratch, (float *)a_out, mir, rider);
    DEXCT_CUDA_OK(cudaGetLastError());
    gn_fixup_kernel<<<sm_count() * 4, 128, fsm, st>>>((const float *)g, (const float *)i0, (const float *)mus, nE, P, n_iters,
                                                      (const unsigned *)scratch, (float *)a_out, mir);
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}
