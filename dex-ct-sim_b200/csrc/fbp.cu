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
#include <algorithm>
#include "common.cuh"
#include "tc05.cuh"

namespace dexct {

static inline size_t fbp_align(size_t x, size_t a) { return (x + a - 1) / a * a; }
constexpr int FBP_RB_MAX = 16;   // detector rows one thread of the multi-row back-projection carries
constexpr int FBP_WARPROWS_MIN = 24;                          // from here on lanes run along the rows (warprows kernel)
static inline int fbp_rows_per_lane(int n_rows) { return n_rows >= 96 ? 4 : (n_rows >= 48 ? 2 : 1); }

struct FbpLayout {
    size_t off_cs, off_taps, off_q, total;
    int q_pitch;                 // elements per filtered line: n_ch + 2 (one zero on either side), rounded up to 4
    int row_pitch;               // rows per channel plane of the row-fastest layout: n_rows rounded up to FBP_RB_MAX
};

static FbpLayout fbp_layout(const DexctScan &s, int view_count, int dtype)
{
    const size_t esz = dtype == DEXCT_F32 ? 4 : 8;
    FbpLayout L;
    size_t o = 0;
    L.off_cs = o;   o = fbp_align(o + 2 * esz * (size_t)view_count, 256);
    L.off_taps = o; o = fbp_align(o + esz * (size_t)(2 * s.n_ch - 1), 256);
    L.q_pitch = (int)fbp_align((size_t)s.n_ch + 2, 4);
    // multi-row scans keep the filtered data row-fastest: [view][channel plane 0 .. n_ch+1][row_pitch] (see
    // fbp_backproject_rows_kernel); the buffer is sized for whichever layout is larger
    L.row_pitch = (int)fbp_align((size_t)s.n_rows, s.n_rows >= FBP_WARPROWS_MIN ? 32 * fbp_rows_per_lane(s.n_rows) : FBP_RB_MAX);
    const size_t by_line = (size_t)view_count * s.n_rows * L.q_pitch;
    const size_t by_row = (size_t)view_count * ((size_t)s.n_ch + 2) * L.row_pitch;
    L.off_q = o;    o = fbp_align(o + esz * (by_line > by_row ? by_line : by_row), 256);
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

// Short scans (total rotation below a full turn): Parker's weights in the form that takes any range pi + 2 delta with
// delta >= half the fan angle; conjugate of ray (b, gamma) is (b + pi + 2 gamma, -gamma) for a source moving towards +beta,
// a scan towards -beta is the mirror image (gamma -> -gamma).  The full-turn formula carries 1/2 for the double coverage, so
// the weights are applied twice over.  delta < 0: full turn, weight 1.  Same expressions as oracle/fbp_oracle.py parker().
struct FbpParker { double delta, absdb; int view0, sign; };
__device__ __forceinline__ double fbp_parker2(const FbpParker &p, long long view_rel, double gamma)
{
    if (p.delta < 0.0) return 1.0;
    const double PI = 3.14159265358979323846;
    const double g = p.sign >= 0 ? gamma : -gamma;
    const double b = (double)(p.view0 + view_rel) * p.absdb;
    double w = 1.0;
    if (b < 2.0 * p.delta - 2.0 * g) { const double sn = sin(0.25 * PI * b / fmax(p.delta - g, 1e-300)); w = sn * sn; }
    if (b > PI - 2.0 * g) { const double sn = sin(0.25 * PI * (PI + 2.0 * p.delta - b) / fmax(p.delta + g, 1e-300)); w = sn * sn; }
    return 2.0 * w;
}

// One detector line (view, row) per CTA: pre-weight into shared memory, then the convolution with the taps.
// Thread t produces channels t, t + blockDim, ...: the tap reads of a warp are consecutive words (conflict-free) and
// the line element is a broadcast.
template <typename R>
__global__ void __launch_bounds__(256) fbp_filter_kernel(const R *__restrict__ sino, int n_ch, double gamma0, double dgamma,
                                                         double sid, const R *__restrict__ taps, R *__restrict__ q, int q_pitch,
                                                         int n_rows, int row_pitch, const FbpParker pk)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    R *line = (R *)smem_raw;                 // [n_ch]
    R *g = line + n_ch;                      // [2 n_ch - 1]
    const long long ln = blockIdx.x;
    const R *src = sino + ln * n_ch;
    for (int c = threadIdx.x; c < n_ch; c += blockDim.x)
        line[c] = src[c] * (R)(sid * cos(gamma0 + (double)c * dgamma) * fbp_parker2(pk, ln / n_rows, gamma0 + (double)c * dgamma));
    for (int t = threadIdx.x; t < 2 * n_ch - 1; t += blockDim.x) g[t] = taps[t];
    __syncthreads();
    // row_pitch == 0: line-major q[line][q_pitch]; else row-fastest q[view][plane][row_pitch] (plane = channel + 1)
    const long long cstride = row_pitch ? row_pitch : 1;
    R *dst = row_pitch ? q + (ln / n_rows) * (long long)(n_ch + 2) * row_pitch + (ln % n_rows) : q + ln * q_pitch;
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
            if (c < n_ch) dst[(1 + c) * cstride] = acc[k] * (R)dgamma;
        }
    }
    if (threadIdx.x == 0) dst[0] = (R)0;
    const int end = row_pitch ? n_ch + 2 : q_pitch;
    for (int c = n_ch + 1 + threadIdx.x; c < end; c += blockDim.x) dst[c * cstride] = (R)0;
}

// Filter, sliding-window form: a CTA convolves 8 consecutive detector lines; thread (r, cb) produces the 8 channels
// [8 cb, 8 cb + 8) of line r.  For 8 consecutive input channels c' it needs the 15 taps g[c - c'] of its window; moving on
// by 8 shifts the window by 8, so each step loads 8 new taps and 8 line values (four 16-byte shared loads, all multicast
// and conflict-free: the 8 lines sit 4 banks apart, the 4 channel blocks of a warp 8 banks apart) for 64 FMAs -- the
// one-line kernel above spends 5 shared loads on 4 FMAs and is bound by shared-memory bandwidth.
// Taps in shared memory: G[i] = g[n = i - off], off = n_chp + 7 (n_chp = n_ch rounded up to 8), zero outside |n| < n_ch.
constexpr int FBP_FL = 8;                                     // lines per CTA
template <typename R>
__global__ void __launch_bounds__(256) fbp_filter8_kernel(const R *__restrict__ sino, long long n_lines, int n_ch, double gamma0,
                                                          double dgamma, double sid, const R *__restrict__ taps,
                                                          R *__restrict__ q, int q_pitch, int n_rows, int row_pitch, const FbpParker pk)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int n_chp = (n_ch + 7) & ~7;
    const int lstride = n_chp + 4 + ((32 - (n_chp & 31)) & 31);         // = 4 (mod 32): the 8 lines 4 banks apart
    const int off = n_chp + 7;
    R *line = (R *)smem_raw;                                             // [8][lstride]
    R *G = line + FBP_FL * lstride;                                      // [2 n_chp + 8]
    const long long ln0 = (long long)blockIdx.x * FBP_FL;
    for (int i = threadIdx.x; i < FBP_FL * n_chp; i += blockDim.x) {
        const int r = i / n_chp, c = i - r * n_chp;
        R v = (R)0;
        if (c < n_ch && ln0 + r < n_lines)
            v = sino[(ln0 + r) * n_ch + c] *
                (R)(sid * cos(gamma0 + (double)c * dgamma) * fbp_parker2(pk, (ln0 + r) / n_rows, gamma0 + (double)c * dgamma));
        line[r * lstride + c] = v;
    }
    for (int i = threadIdx.x; i < 2 * n_chp + 8; i += blockDim.x) {
        const int n = i - off;
        G[i] = (n > -n_ch && n < n_ch) ? taps[n + n_ch - 1] : (R)0;
    }
    __syncthreads();
    const int r = threadIdx.x & 7;
    const long long ln = ln0 + r;
    // row_pitch == 0: line-major q[line][q_pitch]; else row-fastest q[view][plane][row_pitch] (plane = channel + 1)
    const long long cstride = row_pitch ? row_pitch : 1;
    R *dst = nullptr;
    if (ln < n_lines) dst = row_pitch ? q + (ln / n_rows) * (long long)(n_ch + 2) * row_pitch + (ln % n_rows) : q + ln * q_pitch;
    const R *lr = line + r * lstride;
    for (int cb = threadIdx.x >> 3; cb < (n_chp >> 3); cb += blockDim.x >> 3) {
        const int c0 = cb << 3;
        R acc[8], win[16];
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = (R)0;
        // window for c' in [cp, cp + 8): win[m] = g[c0 - cp - 7 + m], m = 0 .. 14; it enters the loop holding its upper half
        const R *gp = G + c0 + off - 7;                                  // &G[index of win[0] at cp = 0], a multiple of 8
#pragma unroll
        for (int m = 0; m < 8; ++m) win[8 + m] = gp[8 + m];
        for (int cp = 0; cp < n_chp; cp += 8) {
            R lv[8];
#pragma unroll
            for (int m = 0; m < 8; ++m) { win[m] = gp[m - cp]; lv[m] = lr[cp + m]; }
#pragma unroll
            for (int j = 0; j < 8; ++j)
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] = fma(lv[j], win[k - j + 7], acc[k]);
#pragma unroll
            for (int m = 0; m < 8; ++m) win[8 + m] = win[m];
        }
        if (dst) {
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (c0 + k < n_ch) dst[(1 + c0 + k) * cstride] = acc[k] * (R)dgamma;
        }
    }
    if (dst && (threadIdx.x >> 3) == 0) {
        dst[0] = (R)0;
        const int end = row_pitch ? n_ch + 2 : q_pitch;
        for (int c = n_ch + 1; c < end; ++c) dst[c * cstride] = (R)0;
    }
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

// Full rotation, view count a multiple of 4, even image size: the pixel turned by a quarter turn sees at view v + N/4 exactly
// what the pixel itself sees at view v (same A, B, hence the same channel coordinate and 1/L^2).  A thread therefore owns the
// four quarter-turn images of one pixel of the first quadrant and computes each (pixel, view) geometry once for all four:
// per step j it works out the geometry at the views j, j + N/4, j + N/2, j + 3N/4 and feeds the sixteen (pixel k, view)
// combinations from them -- 12.5 instead of 22 instructions per pixel-view.  `turn` is +1 when N/4 views advance the source
// by +90 degrees (dbeta > 0), -1 otherwise.
template <typename R, bool FAST>
__global__ void __launch_bounds__(256) fbp_backproject_sym4_kernel(const R *__restrict__ q, int q_pitch, const R *__restrict__ cs,
                                                                   int n_views, int n_rows, int n_ch, R sid, R gamma0,
                                                                   R inv_dgamma, int n, R px, R scale, int accumulate, int turn,
                                                                   R *__restrict__ image)
{
    typedef FbpMath<R, FAST> M;
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;
    const int row = blockIdx.z;
    const int nh = n >> 1;
    if (ix >= nh || iy >= nh) return;
    const R half = (R)0.5 * (R)(n - 1);
    const R x = ((R)ix - half) * px, y = ((R)iy - half) * px;
    const R umax = (R)n_ch, ubias = -gamma0 * inv_dgamma;
    const int nq = n_views >> 2;
    const long long view_stride = (long long)n_rows * q_pitch;
    const R *line = q + (long long)row * q_pitch + 1;
    R acc[4] = {(R)0, (R)0, (R)0, (R)0};
    for (int j = 0; j < nq; ++j) {
        int off[4];
        R f[4], w[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            const int v = j + m * nq;
            const R cb = cs[2 * v], sb = cs[2 * v + 1];
            const R A = sid - fma(y, sb, x * cb);
            const R negB = fma(x, sb, -(y * cb));
            const R gam = M::gamma(negB, A);
            R u = fma(gam, inv_dgamma, ubias);
            u = fmin(fmax(u, (R)-1), umax);
            const R fl = fmin(floor(u), umax - (R)1);
            f[m] = u - fl;
            off[m] = (int)fl;
            w[m] = M::inv(fma(A, A, negB * negB));
        }
        const R *pj = line + j * view_stride;
#pragma unroll
        for (int m = 0; m < 4; ++m)
#pragma unroll
            for (int k = 0; k < 4; ++k) {                    // pixel k (k quarter turns) at view j + ((m + k) & 3) N/4
                const R *p = pj + (long long)((m + k) & 3) * nq * view_stride + off[m];
                const R q0 = p[0], q1 = p[1];
                acc[k] = fma(fma(f[m], q1 - q0, q0), w[m], acc[k]);
            }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        // k quarter turns of (ix, iy): +90 degrees maps (ix, iy) -> (n-1-iy, ix), -90 degrees -> (iy, n-1-ix)
        const int kk = turn > 0 ? k : (4 - k) & 3;
        const int jx = kk == 0 ? ix : kk == 1 ? n - 1 - iy : kk == 2 ? n - 1 - ix : iy;
        const int jy = kk == 0 ? iy : kk == 1 ? ix : kk == 2 ? n - 1 - iy : n - 1 - ix;
        R *out = image + ((long long)row * n + jy) * n + jx;
        const R r = acc[k] * scale;
        *out = accumulate ? *out + r : r;
    }
}

// Multi-row scans: every detector row is its own fan-beam slice with the SAME in-plane geometry, so the channel coordinate,
// the interpolation weight and 1/L^2 of a (pixel, view) pair are computed once and applied to RB rows.  The filtered data
// are row-fastest (q[view][plane][row_pitch], plane = channel + 1 with a zero plane on either side), so the RB values of
// one channel are RB consecutive words: two 16-byte loads per four rows.  Per (pixel, view, row) this costs ~5 instructions
// instead of ~22.
template <typename R, int RB> struct FbpVec;
template <int RB> struct FbpVec<float, RB> {
    static __device__ __forceinline__ void load(const float *p, float (&v)[RB])
    {
#pragma unroll
        for (int k = 0; k < RB; k += 4) {
            const float4 t = __ldg((const float4 *)(p + k));
            v[k] = t.x; v[k + 1] = t.y; v[k + 2] = t.z; v[k + 3] = t.w;
        }
    }
};
template <int RB> struct FbpVec<double, RB> {
    static __device__ __forceinline__ void load(const double *p, double (&v)[RB])
    {
#pragma unroll
        for (int k = 0; k < RB; k += 2) {
            const double2 t = __ldg((const double2 *)(p + k));
            v[k] = t.x; v[k + 1] = t.y;
        }
    }
};

template <typename R, bool FAST, int RB>
__global__ void __launch_bounds__(256) fbp_backproject_rows_kernel(const R *__restrict__ q, int row_pitch, const R *__restrict__ cs,
                                                                   int view_count, int n_rows, int n_ch, R sid, R gamma0,
                                                                   R inv_dgamma, int n, R px, R scale, int accumulate,
                                                                   R *__restrict__ image)
{
    typedef FbpMath<R, FAST> M;
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;
    const int r0 = blockIdx.z * RB;
    if (ix >= n || iy >= n) return;
    const R half = (R)0.5 * (R)(n - 1);
    const R x = ((R)ix - half) * px, y = ((R)iy - half) * px;
    const R umax = (R)n_ch, ubias = -gamma0 * inv_dgamma;
    const long long view_stride = (long long)(n_ch + 2) * row_pitch;
    const R *base = q + row_pitch + r0;                       // plane 1 = channel 0
    R acc[RB];
#pragma unroll
    for (int k = 0; k < RB; ++k) acc[k] = (R)0;
    for (int v = 0; v < view_count; ++v) {
        const R cb = cs[2 * v], sb = cs[2 * v + 1];
        const R A = sid - fma(y, sb, x * cb);
        const R negB = fma(x, sb, -(y * cb));
        const R gam = M::gamma(negB, A);
        R u = fma(gam, inv_dgamma, ubias);
        u = fmin(fmax(u, (R)-1), umax);
        const R fl = fmin(floor(u), umax - (R)1);
        const R f = u - fl;
        const R w = M::inv(fma(A, A, negB * negB));
        const R *p = base + v * view_stride + (long long)(int)fl * row_pitch;
        R q0[RB], q1[RB];
        FbpVec<R, RB>::load(p, q0);
        FbpVec<R, RB>::load(p + row_pitch, q1);
#pragma unroll
        for (int k = 0; k < RB; ++k) acc[k] = fma(fma(f, q1[k] - q0[k], q0[k]), w, acc[k]);
    }
#pragma unroll
    for (int k = 0; k < RB; ++k) {
        if (r0 + k < n_rows) {
            R *out = image + ((long long)(r0 + k) * n + iy) * n + ix;
            const R r = acc[k] * scale;
            *out = accumulate ? *out + r : r;
        }
    }
}

// Many rows (>= 24): lanes run along the detector ROWS.  A warp owns a strip of 16 pixels along x and 32 RPL rows; for
// each view lanes 0..15 work out channel index, interpolation weight and 1/L^2 of "their" pixel and park them in shared
// memory; then, pixel by pixel, every lane fetches the two filtered samples of its RPL rows -- one 128 RPL-byte run per
// channel plane, i.e. fully coalesced -- and accumulates.  The thread-per-pixel kernels above touch 32 different cache
// lines per load instruction (neighbouring pixels fall on different channels) and are bound by L1 tag throughput.
template <typename R, int RPL> struct FbpRowVec;
template <typename R> struct FbpRowVec<R, 1> { static __device__ __forceinline__ void load(const R *p, R (&v)[1]) { v[0] = __ldg(p); } };
template <> struct FbpRowVec<float, 2> {
    static __device__ __forceinline__ void load(const float *p, float (&v)[2]) { const float2 t = __ldg((const float2 *)p); v[0] = t.x; v[1] = t.y; }
};
template <> struct FbpRowVec<float, 4> {
    static __device__ __forceinline__ void load(const float *p, float (&v)[4])
    {
        const float4 t = __ldg((const float4 *)p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
};
template <> struct FbpRowVec<double, 2> {
    static __device__ __forceinline__ void load(const double *p, double (&v)[2]) { const double2 t = __ldg((const double2 *)p); v[0] = t.x; v[1] = t.y; }
};
template <> struct FbpRowVec<double, 4> {
    static __device__ __forceinline__ void load(const double *p, double (&v)[4])
    {
        const double2 a = __ldg((const double2 *)p), b = __ldg((const double2 *)p + 1);
        v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
    }
};

constexpr int FBP_WPX = 8;                                    // pixels of a warp's strip
constexpr int FBP_VPI = 32 / FBP_WPX;                         // views per iteration: every lane computes one (pixel, view) pair
template <typename R, bool FAST, int RPL>
__global__ void __launch_bounds__(256) fbp_backproject_warprows_kernel(const R *__restrict__ q, int row_pitch, const R *__restrict__ cs,
                                                                       int view_count, int n_rows, int n_ch, R sid, R gamma0,
                                                                       R inv_dgamma, int n, R px, R scale, int accumulate,
                                                                       R *__restrict__ image)
{
    typedef FbpMath<R, FAST> M;
    struct Geo { long long off; R f, w; };
    __shared__ Geo geo[8][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ix0 = blockIdx.x * FBP_WPX, iy = blockIdx.y * 8 + warp;
    const int r0 = blockIdx.z * 32 * RPL + lane * RPL;        // first row of this lane (row_pitch covers it)
    if (iy >= n) return;                                      // warp-uniform
    const R half = (R)0.5 * (R)(n - 1);
    const R x = ((R)min(ix0 + (lane & (FBP_WPX - 1)), n - 1) - half) * px, y = ((R)iy - half) * px;
    const int vsub = lane / FBP_WPX;
    const R umax = (R)n_ch, ubias = -gamma0 * inv_dgamma;
    const long long view_stride = (long long)(n_ch + 2) * row_pitch;
    const R *base = q + row_pitch + r0;                       // plane 1 = channel 0
    R acc[FBP_WPX][RPL];
#pragma unroll
    for (int j = 0; j < FBP_WPX; ++j)
#pragma unroll
        for (int k = 0; k < RPL; ++k) acc[j][k] = (R)0;
    for (int v0 = 0; v0 < view_count; v0 += FBP_VPI) {
        {
            const int v = min(v0 + vsub, view_count - 1);
            const R cb = cs[2 * v], sb = cs[2 * v + 1];
            const R A = sid - fma(y, sb, x * cb);
            const R negB = fma(x, sb, -(y * cb));
            const R gam = M::gamma(negB, A);
            R u = fma(gam, inv_dgamma, ubias);
            u = fmin(fmax(u, (R)-1), umax);
            const R fl = fmin(floor(u), umax - (R)1);
            Geo g;
            g.off = v * view_stride + (long long)((int)fl * row_pitch);
            g.f = u - fl;
            g.w = v0 + vsub < view_count ? M::inv(fma(A, A, negB * negB)) : (R)0;     // views past the end weigh nothing
            geo[warp][lane] = g;
        }
        __syncwarp();
#pragma unroll
        for (int t = 0; t < 32; ++t) {
            const Geo g = geo[warp][t];
            const R *p = base + g.off;
            R q0[RPL], q1[RPL];
            FbpRowVec<R, RPL>::load(p, q0);
            FbpRowVec<R, RPL>::load(p + row_pitch, q1);
#pragma unroll
            for (int k = 0; k < RPL; ++k)
                acc[t % FBP_WPX][k] = fma(fma(g.f, q1[k] - q0[k], q0[k]), g.w, acc[t % FBP_WPX][k]);
        }
        __syncwarp();
    }
#pragma unroll
    for (int j = 0; j < FBP_WPX; ++j) {
        if (ix0 + j >= n) break;
#pragma unroll
        for (int k = 0; k < RPL; ++k) {
            if (r0 + k < n_rows) {
                R *out = image + ((long long)(r0 + k) * n + iy) * n + ix0 + j;
                const R r = acc[j][k] * scale;
                *out = accumulate ? *out + r : r;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Filter on the tensor cores (single precision, row-fastest output, full-turn scans).  The convolution of every detector
// line with the 2 n_ch - 1 taps is a product with a Toeplitz matrix: Q[line][c] = dgamma sum_c' X'[line][c'] g[c - c'] --
// the one genuine contraction of the reconstruction (n_ch^2 FMAs per line, 72 % of the FMA pipe in fbp_filter8_kernel).
// Here it runs as tcgen05.mma kind::tf32 with the "3xTF32" split of gn_tc_kernel (gn.cu):
//   * a CTA (one per SM, 512 columns of tensor memory) owns a slab of 256 output channels [n0, n0 + 256) and walks over
//     tiles of 128 detector lines (thread t of the four compute warps <-> line t <-> lane t of tensor memory);
//   * B needs no work per step: for input channels [k0, k0 + 8) it is the 256 x 8 window B[n][k] = g[n0 + n - k0 - k] of the
//     taps, and the windows of successive k0 are the SAME strip of 8-column rows shifted by 8 rows.  The strip
//     S[r][k] = g[n0 + r - (Kp - 8) - k], r = 0 .. 255 + Kp - 8, is built once per CTA in shared memory in the K-major
//     no-swizzle core-matrix layout (8 rows x 16 bytes; split into TF32 heads and tails: 2 x 64 KiB for 1800 channels), and
//     the matrix descriptor of a step simply starts (Kp - 8 - k0) / 8 row groups into it;
//   * A: every compute warp stages 32 lines x 32 input channels through shared memory (coalesced loads, one line per thread
//     out), applies the pre-weight SID cos(gamma), splits heads and tails and stores them into tensor memory
//     (tcgen05.st.x16 per 8 channels), double buffered in chunks of four k-steps;
//   * the MMA warp issues per k-step  D += hi x Bhi,  D += lo x Bhi,  D += hi x Blo  (M128 N256 K8, 128 cycles each: the
//     tensor pipe is the bound here -- 3 x 128 cycles per 8 channels, 128 lines and 256 outputs);
//   * after the last k-step the threads read their line of D (128 lanes x 256 columns) back and write it row-fastest.
constexpr int FT_N = 256;                    // output channels per CTA slab = accumulator columns
constexpr int FT_CKS = 4;                    // k-steps (of 8 input channels) per A chunk
constexpr int FT_CH = FT_CKS * 8;            // input channels per chunk
constexpr int FT_ACOL = FT_N;                // first A column in tensor memory
constexpr int FT_THREADS = 160;
constexpr int FT_LSTRIDE = 36;               // floats per staged line (32 + 4: conflict-free LDS.128 by line)
constexpr uint32_t FT_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(FT_N >> 3) << 17) | ((128u >> 4) << 24);

__global__ void __launch_bounds__(FT_THREADS, 1) fbp_filter_tc_kernel(const float *__restrict__ sino, long long n_lines, int n_ch,
                                                                       double gamma0, double dgamma, double sid,
                                                                       const float *__restrict__ taps, float *__restrict__ q,
                                                                       int n_rows, int row_pitch, int n_slabs)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int Kp = (n_ch + 7) & ~7;
    const int R = FT_N + Kp - 8;                              // rows of the tap strip
    float *s_hi = (float *)smem_raw;                          // [R / 8][2 K halves][8 rows][4]
    float *s_lo = s_hi + (size_t)R * 8;
    float *wgt = s_lo + (size_t)R * 8;                        // [Kp] pre-weights SID cos(gamma), zero past n_ch
    float *stage = wgt + Kp;                                  // [4 warps][32 lines][FT_LSTRIDE]
    __shared__ __align__(8) unsigned long long bars[5];
    __shared__ uint32_t tmem_base_sm;
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const uint32_t bar_full = smem_u32(&bars[0]), bar_free = smem_u32(&bars[2]), bar_done = smem_u32(&bars[4]);

    const int slab = blockIdx.x % n_slabs, group = blockIdx.x / n_slabs, n_groups = gridDim.x / n_slabs;
    const int n0 = slab * FT_N;
    // the tap strip of this slab, heads and tails
    for (int e = tid; e < R * 8; e += FT_THREADS) {
        const int r = e >> 3, k = e & 7;
        const int n = n0 + r - (Kp - 8) - k;                  // tap index c - c'
        const float g = (n > -n_ch && n < n_ch) ? taps[n + (n_ch - 1)] : 0.f;
        const float hi = __uint_as_float((__float_as_uint(g) + 0x1000u) & 0xffffe000u);
        const int at = ((r >> 3) * 2 + (k >> 2)) * 32 + (r & 7) * 4 + (k & 3);
        s_hi[at] = hi;
        s_lo[at] = g - hi;
    }
    for (int c = tid; c < Kp; c += FT_THREADS) wgt[c] = c < n_ch ? (float)(sid * cos(gamma0 + (double)c * dgamma)) : 0.f;
    if (tid == 0) {
        for (int b = 0; b < 2; ++b) { mbar_init(bar_full + 8 * b, 4); mbar_init(bar_free + 8 * b, 1); }
        mbar_init(bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base_sm)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_sm;
    const long long n_tiles = (n_lines + 127) / 128;
    const int n_ks = Kp / 8, n_chunks = (n_ks + FT_CKS - 1) / FT_CKS;

    if (warp == 4) {
        // descriptors: K-major, no swizzle, core matrices of 8 rows x 16 B; next K half 128 B on, next row group 256 B on
        const uint64_t desc0 = ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46);
        const uint32_t hi_addr = smem_u32(s_hi), lo_addr = smem_u32(s_lo);
        uint32_t cnt = 0;
        for (long long tile = group; tile < n_tiles; tile += n_groups)
#pragma unroll 1
            for (int c = 0; c < n_chunks; ++c, ++cnt) {
                const uint32_t b = cnt & 1;
                mbar_wait(bar_full + 8 * b, (cnt >> 1) & 1);
                tc_fence_after();
                const uint32_t a_buf = tmem + FT_ACOL + b * (FT_CKS * 16);
#pragma unroll
                for (int s = 0; s < FT_CKS; ++s) {
                    const int ks = c * FT_CKS + s;
                    if (ks < n_ks) {
                        const uint32_t row_ofs = (uint32_t)((Kp - 8 - 8 * ks) >> 3) * 256;        // bytes into the strip
                        const uint64_t dhi = desc0 | (uint64_t)(((hi_addr + row_ofs) & 0x3ffff) >> 4);
                        const uint64_t dlo = desc0 | (uint64_t)(((lo_addr + row_ofs) & 0x3ffff) >> 4);
                        tc_mma_tf32(tmem, a_buf + s * 16, dhi, FT_IDESC, ks > 0);
                        tc_mma_tf32(tmem, a_buf + s * 16 + 8, dhi, FT_IDESC, 1);
                        tc_mma_tf32(tmem, a_buf + s * 16, dlo, FT_IDESC, 1);
                    }
                }
                tc_commit(bar_free + 8 * b);
                if (c == n_chunks - 1) tc_commit(bar_done);
            }
        __syncwarp();
    } else {
        float *my_stage = stage + warp * 32 * FT_LSTRIDE;
        const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
        const long long plane = row_pitch;                   // elements between channel planes of the row-fastest layout
        uint32_t cnt = 0, done_cnt = 0;
        for (long long tile = group; tile < n_tiles; tile += n_groups) {
            const long long line0 = tile * 128 + warp * 32;   // first line of this warp
            const long long my_line = line0 + lane;
            // 32 lines x 32 channels per chunk: lane = channel on the way in (coalesced), lane = line on the way out; the
            // loads of the NEXT chunk are in flight while this one is split and stored
            float in[32];
            auto fetch = [&](int k0) {
#pragma unroll
                for (int l = 0; l < 32; ++l) {
                    const long long ln = line0 + l;
                    in[l] = (ln < n_lines && k0 + lane < n_ch) ? __ldg(sino + ln * n_ch + k0 + lane) : 0.f;
                }
            };
            fetch(0);
#pragma unroll 1
            for (int c = 0; c < n_chunks; ++c, ++cnt) {
                const uint32_t b = cnt & 1;
                const int k0 = c * FT_CH;
                __syncwarp();
#pragma unroll
                for (int l = 0; l < 32; ++l) my_stage[l * FT_LSTRIDE + lane] = in[l];
                __syncwarp();
                if (c + 1 < n_chunks) fetch(k0 + FT_CH);
                mbar_wait(bar_free + 8 * b, ((cnt >> 1) & 1) ^ 1);
                tc_fence_after();
#pragma unroll
                for (int s = 0; s < FT_CKS; ++s) {
                    if (c * FT_CKS + s < n_ks) {
                        const float4 x0 = *(const float4 *)(my_stage + lane * FT_LSTRIDE + 8 * s);
                        const float4 x1 = *(const float4 *)(my_stage + lane * FT_LSTRIDE + 8 * s + 4);
                        const float4 w0 = *(const float4 *)(wgt + k0 + 8 * s), w1 = *(const float4 *)(wgt + k0 + 8 * s + 4);
                        const float v[8] = {x0.x * w0.x, x0.y * w0.y, x0.z * w0.z, x0.w * w0.w, x1.x * w1.x, x1.y * w1.y, x1.z * w1.z, x1.w * w1.w};
                        uint32_t a[16];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            a[j] = __float_as_uint(v[j]) & 0xffffe000u;
                            a[8 + j] = __float_as_uint(v[j] - __uint_as_float(a[j]));
                        }
                        tmem_st16(lane_base + FT_ACOL + b * (FT_CKS * 16) + s * 16, a);
                    }
                }
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_full + 8 * b);
            }
            // the filtered line of this thread: 256 columns of its lane, written row-fastest
            mbar_wait(bar_done, done_cnt++ & 1);
            tc_fence_after();
            const bool live = my_line < n_lines;
            const long long view = live ? my_line / n_rows : 0;
            const int row = live ? (int)(my_line - view * n_rows) : 0;
            float *dst = q + view * (long long)(n_ch + 2) * plane + row;
#pragma unroll 1
            for (int cb = 0; cb < FT_N; cb += 32) {
                float d[32];
                tmem_ld32(lane_base + cb, d);
                if (live) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int ch = n0 + cb + j;
                        if (ch < n_ch) dst[(long long)(1 + ch) * plane] = d[j] * (float)dgamma;
                    }
                }
            }
            tc_fence_before();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
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
    // total rotation of the scan: one full turn (no weighting), or a short scan of at least pi + fan (Parker weights)
    const double PI = 3.14159265358979323846;
    const double total = fabs(s.dbeta) * (double)s.n_views;
    const double half_fan = fmax(fabs(s.gamma0), fabs(s.gamma0 + (s.n_ch - 1) * s.dgamma));
    FbpParker pk{-1.0, fabs(s.dbeta), view_begin, s.dbeta >= 0 ? 1 : -1};
    if (fabs(total - 2.0 * PI) > 1e-6) {
        if (!(total < 2.0 * PI && 0.5 * (total - PI) >= half_fan - 1e-12)) {
            set_error("total rotation %.6f rad: get_recon takes one full turn (2 pi) or a short scan of at least pi + fan angle = "
                      "%.6f rad and less than 2 pi", total, PI + 2.0 * half_fan);
            return DEXCT_EINVAL;
        }
        pk.delta = 0.5 * (total - PI);
    }
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
    // rows >= 8: row-fastest filtered data and the kernel that shares a (pixel, view)'s geometry between RB rows
    const char *rb_env = getenv("DEXCT_FBP_RB");              // tuning knob: 0 forces the one-row kernel
    int rb = s.n_rows >= 16 ? 16 : (s.n_rows >= 8 ? 8 : 0);
    if (rb_env && atoi(rb_env) != 32) rb = atoi(rb_env) == 16 ? 16 : (atoi(rb_env) == 8 ? 8 : 0);   // 32: automatic choice
    const int row_pitch = rb ? L.row_pitch : 0;
    const bool warprows = rb && s.n_rows >= FBP_WARPROWS_MIN && !(rb_env && atoi(rb_env) != 32);
    if (rb && s.n_rows % L.row_pitch)                         // rows of the last block that no line fills
        DEXCT_CUDA_OK(cudaMemsetAsync(q, 0, sizeof(R) * (size_t)view_count * (s.n_ch + 2) * L.row_pitch, st));
    {
        const int n_chp = (s.n_ch + 7) & ~7;
        const int lstride = n_chp + 4 + ((32 - (n_chp & 31)) & 31);
        const size_t smem8 = sizeof(R) * ((size_t)FBP_FL * lstride + 2 * n_chp + 8);
        const char *f_env = getenv("DEXCT_FBP_FILTER");       // tuning knob: 1 forces the one-line kernel
        const long long n_lines = (long long)view_count * s.n_rows;
        // single precision, row-fastest output, full turn, at least one slab of channels: the tensor-core filter
        const char *tc_env = getenv("DEXCT_FBP_TC");          // 0: the FP32-pipe filter kernels
        const int Kp = (s.n_ch + 7) & ~7;
        const size_t smem_tc = sizeof(float) * ((size_t)(FT_N + Kp - 8) * 16 + Kp + 4 * 32 * FT_LSTRIDE);
        if (sizeof(R) == 4 && row_pitch && pk.delta < 0 && s.n_ch >= 64 && smem_tc <= 220 * 1024 && !(tc_env && atoi(tc_env) == 0) &&
            !(f_env && atoi(f_env) == 1)) {
            const int n_slabs = (s.n_ch + FT_N - 1) / FT_N;
            const long long n_tiles = (n_lines + 127) / 128;
            int groups = (int)std::min<long long>(n_tiles, std::max(1, sm_count() / n_slabs));
            // planes 0 and n_ch + 1 (the zero taper) and the rows no line fills
            DEXCT_CUDA_OK(cudaMemsetAsync(q, 0, sizeof(R) * (size_t)view_count * (s.n_ch + 2) * L.row_pitch, st));
            DEXCT_CUDA_OK(cudaFuncSetAttribute(fbp_filter_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tc));
            fbp_filter_tc_kernel<<<(unsigned)(groups * n_slabs), FT_THREADS, smem_tc, st>>>(
                (const float *)sino, n_lines, s.n_ch, s.gamma0, s.dgamma, s.sid, (const float *)taps, (float *)q, s.n_rows, row_pitch, n_slabs);
        } else if (smem8 <= 200 * 1024 && !(f_env && atoi(f_env) == 1)) {
            if (smem8 > 48 * 1024)
                DEXCT_CUDA_OK(cudaFuncSetAttribute(fbp_filter8_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem8));
            fbp_filter8_kernel<R><<<(unsigned)((n_lines + FBP_FL - 1) / FBP_FL), 256, smem8, st>>>(
                (const R *)sino, n_lines, s.n_ch, s.gamma0, s.dgamma, s.sid, taps, q, L.q_pitch, s.n_rows, row_pitch, pk);
        } else {
            fbp_filter_kernel<R><<<(unsigned)n_lines, 256, smem, st>>>(
                (const R *)sino, s.n_ch, s.gamma0, s.dgamma, s.sid, taps, q, L.q_pitch, s.n_rows, row_pitch, pk);
        }
    }
    DEXCT_CUDA_OK(cudaGetLastError());
    // fast path: the whole detector (plus the one-channel taper) inside the polynomial's range
    const double gmax = fmax(fabs(s.gamma0), fabs(s.gamma0 + (s.n_ch - 1) * s.dgamma)) + 2.0 * s.dgamma;
    const bool fast = sizeof(R) == 4 && gmax < atan((double)FBP_T0);
    if (warprows) {
        const int rpl = fbp_rows_per_lane(s.n_rows);
        const dim3 grid((n + FBP_WPX - 1) / FBP_WPX, (n + 7) / 8, (s.n_rows + 32 * rpl - 1) / (32 * rpl));
        auto kern = fbp_backproject_warprows_kernel<R, false, 1>;
        if (rpl == 4) kern = fast ? fbp_backproject_warprows_kernel<R, sizeof(R) == 4, 4> : fbp_backproject_warprows_kernel<R, false, 4>;
        else if (rpl == 2) kern = fast ? fbp_backproject_warprows_kernel<R, sizeof(R) == 4, 2> : fbp_backproject_warprows_kernel<R, false, 2>;
        else if (fast) kern = fbp_backproject_warprows_kernel<R, sizeof(R) == 4, 1>;
        kern<<<grid, 256, 0, st>>>(q, row_pitch, cs, view_count, s.n_rows, s.n_ch, (R)s.sid, (R)s.gamma0, (R)(1.0 / s.dgamma),
                                   n, (R)(fov / n), (R)fabs(s.dbeta), accumulate, (R *)image);
        DEXCT_CUDA_OK(cudaGetLastError());
        return DEXCT_OK;
    }
    if (rb) {
        const dim3 block(32, 8), grid((n + 31) / 32, (n + 7) / 8, (s.n_rows + rb - 1) / rb);
        auto kern = fbp_backproject_rows_kernel<R, false, 8>;
        if (rb == 16) kern = fast ? fbp_backproject_rows_kernel<R, sizeof(R) == 4, 16> : fbp_backproject_rows_kernel<R, false, 16>;
        else if (fast) kern = fbp_backproject_rows_kernel<R, sizeof(R) == 4, 8>;
        kern<<<grid, block, 0, st>>>(q, row_pitch, cs, view_count, s.n_rows, s.n_ch, (R)s.sid, (R)s.gamma0, (R)(1.0 / s.dgamma),
                                     n, (R)(fov / n), (R)fabs(s.dbeta), accumulate, (R *)image);
        DEXCT_CUDA_OK(cudaGetLastError());
        return DEXCT_OK;
    }
    // whole full-rotation scans with 4 | n_views and an even image: quarter-turn symmetry (fbp_backproject_sym4_kernel)
    const double turn_angle = s.dbeta * (double)(s.n_views / 4);
    const char *sym_env = getenv("DEXCT_FBP_SYM");            // tuning knob: 0 forces the general kernel
    if (view_begin == 0 && view_count == s.n_views && s.n_views % 4 == 0 && n % 2 == 0 &&
        fabs(fabs(turn_angle) - 1.5707963267948966) < 1e-9 && !(sym_env && atoi(sym_env) == 0)) {
        const dim3 block(32, 8), grid((n / 2 + 31) / 32, (n / 2 + 7) / 8, s.n_rows);
        auto kern = fbp_backproject_sym4_kernel<R, false>;
        if (fast) kern = fbp_backproject_sym4_kernel<R, sizeof(R) == 4>;
        kern<<<grid, block, 0, st>>>(q, L.q_pitch, cs, s.n_views, s.n_rows, s.n_ch, (R)s.sid, (R)s.gamma0, (R)(1.0 / s.dgamma),
                                     n, (R)(fov / n), (R)fabs(s.dbeta), accumulate, turn_angle > 0 ? 1 : -1, (R *)image);
        DEXCT_CUDA_OK(cudaGetLastError());
        return DEXCT_OK;
    }
    const dim3 block(32, 8), grid((n + 31) / 32, (n + 7) / 8, s.n_rows);
    auto kern = fbp_backproject_kernel<R, false>;
    if (fast) kern = fbp_backproject_kernel<R, sizeof(R) == 4>;
    kern<<<grid, block, 0, st>>>(q, L.q_pitch, cs, view_count, s.n_rows, s.n_ch, (R)s.sid, (R)s.gamma0, (R)(1.0 / s.dgamma), n,
                                 (R)(fov / n), (R)fabs(s.dbeta), accumulate, (R *)image);
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
