// Fan-beam filtered back-projection as a gather over views.  Replaces get_recon (/root/reference/main.py:134,168; the
// reference's implementation sits in the absent submodule, README.md:30-31 describes it as "fan-beam filtered back
// projection with a sinc window filter").  Row f1 of SURVEY.md section 8; conventions in DESIGN.md section 7.
//
// Equiangular fan-beam FBP (Kak & Slaney, ch. 3.4.1) for a full rotation:
//   R'(beta, gamma) = R(beta, gamma) SID cos(gamma)                                  pre-weight
//   Q(beta, .)      = dgamma  R'(beta, .) * g,   g(gamma) = 1/2 (gamma / sin gamma)^2 h(gamma)   filter
//   f(x, y)         = dbeta sum_beta Q(beta, gamma'(x, y, beta)) / L^2(x, y, beta)             back-projection
// h is the impulse response of the ramp filter |f| multiplied by the window sinc(f / (2 fc)) and cut at
// fc = ramp * Nyquist ("ramp_filter_percent_Nyquist", input/params.txt:35), in closed form:
//   h(t) = (2 fc^2 / pi) [ s(pi/4 + pi fc t) + s(pi/4 - pi fc t) ],   s(u) = sin^2(u) / u   (ramp = 1: Shepp-Logan)
// Multi-row sinograms are reconstructed row by row (one fan-beam slice per detector row).
#include "common.cuh"

namespace dexct {

static inline size_t fbp_align(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct FbpLayout {
    size_t off_cs, off_taps, off_q, total;
    int q_pitch;                 // elements per filtered line: n_ch + 2 (one zero on either side), rounded up to 4
};

static FbpLayout fbp_layout(const DexctScan &s, int view_count, int dtype)
{
    const size_t esz = dtype == DEXCT_F32 ? 4 : 8;
    FbpLayout L;
    size_t o = 0;
    L.off_cs = o;   o = fbp_align(o + 2 * esz * (size_t)view_count, 256);
    L.off_taps = o; o = fbp_align(o + esz * (size_t)(2 * s.n_ch - 1), 256);
    L.q_pitch = (int)fbp_align((size_t)s.n_ch + 2, 4);
    L.off_q = o;    o = fbp_align(o + esz * (size_t)view_count * s.n_rows * L.q_pitch, 256);
    L.total = o;
    return L;
}

__device__ __forceinline__ double fbp_s(double u) { return u == 0.0 ? 0.0 : sin(u) * sin(u) / u; }

// cos/sin of the view angles (double precision, rounded once) and the filter taps g[n], n = -(n_ch-1) .. n_ch-1.
template <typename R>
__global__ void fbp_setup_kernel(DexctScan s, int view_begin, int view_count, double ramp, R *__restrict__ cs,
                                 R *__restrict__ taps)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < view_count) {
        double sn, c;
        sincos(__dadd_rn(s.beta0, __dmul_rn((double)(view_begin + t), s.dbeta)), &sn, &c);
        cs[2 * t] = (R)c;
        cs[2 * t + 1] = (R)sn;
    }
    if (t < 2 * s.n_ch - 1) {
        const double PI = 3.14159265358979323846;
        const int n = t - (s.n_ch - 1);
        const double g = (double)n * s.dgamma;
        const double fc = ramp / (2.0 * s.dgamma);
        const double h = (2.0 * fc * fc / PI) * (fbp_s(0.25 * PI + PI * fc * g) + fbp_s(0.25 * PI - PI * fc * g));
        const double w = n == 0 ? 1.0 : (g / sin(g)) * (g / sin(g));
        taps[t] = (R)(0.5 * w * h);
    }
}

// One detector line (view, row) per CTA: pre-weight into shared memory, then the convolution with the taps.
// Thread t produces channels t, t + blockDim, ...: the tap reads of a warp are consecutive words (conflict-free) and
// the line element is a broadcast.
template <typename R>
__global__ void __launch_bounds__(256) fbp_filter_kernel(const R *__restrict__ sino, int n_ch, double gamma0, double dgamma,
                                                         double sid, const R *__restrict__ taps, R *__restrict__ q, int q_pitch)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    R *line = (R *)smem_raw;                 // [n_ch]
    R *g = line + n_ch;                      // [2 n_ch - 1]
    const long long ln = blockIdx.x;
    const R *src = sino + ln * n_ch;
    for (int c = threadIdx.x; c < n_ch; c += blockDim.x)
        line[c] = src[c] * (R)(sid * cos(gamma0 + (double)c * dgamma));
    for (int t = threadIdx.x; t < 2 * n_ch - 1; t += blockDim.x) g[t] = taps[t];
    __syncthreads();
    R *dst = q + ln * q_pitch;
    constexpr int U = 4;
    for (int c0 = threadIdx.x; c0 < n_ch; c0 += U * blockDim.x) {
        R acc[U];
        const R *gp[U];
#pragma unroll
        for (int k = 0; k < U; ++k) {
            acc[k] = (R)0;
            // channels beyond the line read taps of a neighbouring (valid) position; their sums are discarded
            const int c = min(c0 + k * (int)blockDim.x, n_ch - 1);
            gp[k] = g + c + (n_ch - 1);      // g[n = c - c'] sits at gp[k][-c']
        }
#pragma unroll 4
        for (int cp = 0; cp < n_ch; ++cp) {
            const R v = line[cp];
#pragma unroll
            for (int k = 0; k < U; ++k) acc[k] = fma(v, gp[k][-cp], acc[k]);
        }
#pragma unroll
        for (int k = 0; k < U; ++k) {
            const int c = c0 + k * (int)blockDim.x;
            if (c < n_ch) dst[1 + c] = acc[k] * (R)dgamma;
        }
    }
    if (threadIdx.x == 0) dst[0] = (R)0;
    for (int c = n_ch + 1 + threadIdx.x; c < q_pitch; c += blockDim.x) dst[c] = (R)0;
}

// Channel coordinate and 1/L^2 of a pixel for one view.  FAST (single precision, fan half-angle below atan(0.85) = 40 deg):
// reciprocal by MUFU.RCP plus one Newton step, atan by an odd degree-17 minimax polynomial on |t| <= 0.85 (absolute error
// 6e-8 rad evaluated in float, against 4.5e-8 for a correctly rounded atanf); t is clamped to that range first -- beyond it
// the ray misses the detector anyway.  Otherwise: IEEE division and atan2.
constexpr float FBP_T0 = 0.85f;
template <typename R, bool FAST> struct FbpMath {
    static __device__ __forceinline__ R gamma(R negB, R A) { return atan2(negB, A); }
    static __device__ __forceinline__ R inv(R x) { return (R)1 / x; }
};
template <> struct FbpMath<float, true> {
    static __device__ __forceinline__ float rcp(float x)
    {
        float r;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
        return r;
    }
    static __device__ __forceinline__ float gamma(float negB, float A)
    {
        const float r0 = rcp(A);
        const float r = fmaf(r0, fmaf(-A, r0, 1.0f), r0);
        const float t = fminf(fmaxf(negB * r, -FBP_T0), FBP_T0);
        const float u = t * t;
        float q = 0.0057596251515161295f;
        q = fmaf(q, u, -0.02522639020233572f);
        q = fmaf(q, u, 0.05444744168992351f);
        q = fmaf(q, u, -0.08292680841837932f);
        q = fmaf(q, u, 0.1093448867911483f);
        q = fmaf(q, u, -0.142630518713619f);
        q = fmaf(q, u, 0.19998517545978775f);
        q = fmaf(q, u, -0.33333295659610157f);
        return fmaf(t * u, q, t);
    }
    static __device__ __forceinline__ float inv(float x) { return rcp(x); }
};

// One image pixel per thread, all views of the range in a loop.  The 32 pixels of a warp are neighbours along x: for
// one view they fall on a short run of channels, so the two gathers per pixel hit a couple of L1 lines of a filtered
// line that lives in L2 for the whole launch.
template <typename R, bool FAST>
__global__ void __launch_bounds__(256) fbp_backproject_kernel(const R *__restrict__ q, int q_pitch, const R *__restrict__ cs,
                                                              int view_count, int n_rows, int n_ch, R sid, R gamma0,
                                                              R inv_dgamma, int n, R px, R scale, int accumulate,
                                                              R *__restrict__ image)
{
    typedef FbpMath<R, FAST> M;
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;
    const int row = blockIdx.z;
    if (ix >= n || iy >= n) return;
    const R half = (R)0.5 * (R)(n - 1);
    const R x = ((R)ix - half) * px, y = ((R)iy - half) * px;
    const R umax = (R)n_ch, ubias = -gamma0 * inv_dgamma;
    const R *line = q + (long long)row * q_pitch + 1;
    const long long view_stride = (long long)n_rows * q_pitch;
    R acc0 = (R)0, acc1 = (R)0;
    auto one = [&](int v, R &acc) {
        const R cb = cs[2 * v], sb = cs[2 * v + 1];
        const R A = sid - fma(y, sb, x * cb);
        const R negB = fma(x, sb, -(y * cb));
        const R gam = M::gamma(negB, A);
        R u = fma(gam, inv_dgamma, ubias);
        u = fmin(fmax(u, (R)-1), umax);
        const R fl = fmin(floor(u), umax - (R)1);
        const R f = u - fl;
        const R *p = line + v * view_stride + (int)fl;
        const R q0 = p[0], q1 = p[1];
        acc = fma(fma(f, q1 - q0, q0), M::inv(fma(A, A, negB * negB)), acc);
    };
    int v = 0;
    for (; v + 1 < view_count; v += 2) {              // two independent accumulation chains
        one(v, acc0);
        one(v + 1, acc1);
    }
    if (v < view_count) one(v, acc0);
    R *out = image + ((long long)row * n + iy) * n + ix;
    const R r = (acc0 + acc1) * scale;
    *out = accumulate ? *out + r : r;
}

template <typename R>
__global__ void fbp_hu_kernel(const R *__restrict__ image, long long n, R mu_water, R *__restrict__ hu)
{
    const R k = (R)1000 / mu_water;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        hu[i] = (image[i] - mu_water) * k;
}

template <typename R>
static int fbp_run(const DexctScan &s, const void *sino, int view_begin, int view_count, int n, double fov, double ramp,
                   void *image, int accumulate, void *scratch, int dtype, cudaStream_t st)
{
    const FbpLayout L = fbp_layout(s, view_count, dtype);
    char *w = (char *)scratch;
    R *cs = (R *)(w + L.off_cs), *taps = (R *)(w + L.off_taps), *q = (R *)(w + L.off_q);
    const int nset = max(view_count, 2 * s.n_ch - 1);
    fbp_setup_kernel<R><<<(nset + 127) / 128, 128, 0, st>>>(s, view_begin, view_count, ramp, cs, taps);
    DEXCT_CUDA_OK(cudaGetLastError());
    const size_t smem = sizeof(R) * (size_t)(3 * s.n_ch - 1);
    if (smem > 200 * 1024) {
        set_error("FBP filter needs %zu bytes of shared memory (limit 200 KiB): n_ch=%d", smem, s.n_ch);
        return DEXCT_ECAPACITY;
    }
    if (smem > 48 * 1024)
        DEXCT_CUDA_OK(cudaFuncSetAttribute(fbp_filter_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    fbp_filter_kernel<R><<<(unsigned)((long long)view_count * s.n_rows), 256, smem, st>>>(
        (const R *)sino, s.n_ch, s.gamma0, s.dgamma, s.sid, taps, q, L.q_pitch);
    DEXCT_CUDA_OK(cudaGetLastError());
    const dim3 block(32, 8), grid((n + 31) / 32, (n + 7) / 8, s.n_rows);
    // fast path: the whole detector (plus the one-channel taper) inside the polynomial's range
    const double gmax = fmax(fabs(s.gamma0), fabs(s.gamma0 + (s.n_ch - 1) * s.dgamma)) + 2.0 * s.dgamma;
    const bool fast = sizeof(R) == 4 && gmax < atan((double)FBP_T0);
    auto kern = fbp_backproject_kernel<R, false>;
    if (fast) kern = fbp_backproject_kernel<R, sizeof(R) == 4>;
    kern<<<grid, block, 0, st>>>(q, L.q_pitch, cs, view_count, s.n_rows, s.n_ch, (R)s.sid, (R)s.gamma0, (R)(1.0 / s.dgamma), n,
                                 (R)(fov / n), (R)s.dbeta, accumulate, (R *)image);
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}

}  // namespace dexct

using namespace dexct;

extern "C" size_t dexct_fbp_workspace_bytes(const DexctScan *scan, int32_t view_count, int32_t dtype)
{
    if (!scan || scan->n_ch <= 0 || scan->n_rows <= 0 || view_count <= 0 || (dtype != DEXCT_F32 && dtype != DEXCT_F64)) return 0;
    return fbp_layout(*scan, view_count, dtype).total;
}

extern "C" int dexct_fbp_fan(const DexctScan *scan, const void *sino, int32_t view_begin, int32_t view_count,
                             int32_t n_matrix, double fov, double ramp, void *image, int32_t accumulate, void *scratch,
                             int32_t dtype, void *stream)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(scan, "scan must be non-null");
    DEXCT_REQUIRE(scan->n_views > 0 && scan->n_rows > 0 && scan->n_ch > 1, "scan sizes must be positive (n_ch >= 2)");
    DEXCT_REQUIRE(view_begin >= 0 && view_count >= 0 && view_begin + view_count <= scan->n_views,
                  "view range [%d, %d) outside the scan's %d views", view_begin, view_begin + view_count, scan->n_views);
    DEXCT_REQUIRE(n_matrix > 0 && fov > 0, "n_matrix and fov must be positive");
    DEXCT_REQUIRE(ramp > 0 && ramp <= 1, "ramp (cut-off as a fraction of Nyquist) must be in (0, 1], got %g", ramp);
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    DEXCT_REQUIRE(scan->sid > 0 && scan->dgamma > 0 && scan->dbeta != 0, "SID, dgamma must be positive and dbeta non-zero");
    // every pixel must lie in front of the source for every view: half the diagonal of the square inside the source circle
    DEXCT_REQUIRE(0.70710678118654752 * fov < scan->sid, "FOV %g cm does not fit inside the source circle (SID %g cm)", fov, scan->sid);
    DEXCT_REQUIRE(image, "image must be non-null");
    cudaStream_t st = (cudaStream_t)stream;
    if (view_count == 0) {
        if (!accumulate) {
            const size_t esz = dtype == DEXCT_F32 ? 4 : 8;
            DEXCT_CUDA_OK(cudaMemsetAsync(image, 0, esz * (size_t)scan->n_rows * n_matrix * n_matrix, st));
        }
        return DEXCT_OK;
    }
    DEXCT_REQUIRE(sino && scratch, "sino and scratch must be non-null");
    DexctScan s = *scan;
    s.dbeta = scan->dbeta;
    if (dtype == DEXCT_F32)
        return fbp_run<float>(s, sino, view_begin, view_count, n_matrix, fov, ramp, image, accumulate, scratch, dtype, st);
    return fbp_run<double>(s, sino, view_begin, view_count, n_matrix, fov, ramp, image, accumulate, scratch, dtype, st);
}

extern "C" int dexct_fbp_hu(const void *image, int64_t n, double mu_water, void *hu, int32_t dtype, void *stream)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(n >= 0 && mu_water > 0, "n >= 0 and mu_water > 0 required");
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    if (n == 0) return DEXCT_OK;
    DEXCT_REQUIRE(image && hu, "image and hu must be non-null");
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = (unsigned)min((long long)sm_count() * 8, (long long)((n + 255) / 256));
    if (dtype == DEXCT_F32) fbp_hu_kernel<float><<<grid, 256, 0, st>>>((const float *)image, n, (float)mu_water, (float *)hu);
    else fbp_hu_kernel<double><<<grid, 256, 0, st>>>((const double *)image, n, mu_water, (double *)hu);
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}
