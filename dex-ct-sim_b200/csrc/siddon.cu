// Forward projection kernels: workspace preparation, Siddon traversal with fused polyenergetic detection,
// detection from stored lengths, and the debug trace.  Replaces get_sino (/root/reference/main.py:120).
#include "siddon_core.cuh"
#include <algorithm>

namespace dexct {

// ---------------------------------------------------------------------------------------------------------
// Workspace preparation

__global__ void tables_kernel(DexctScan s, double *cosb, double *sinb, double *cosg, double *sing, double *zdet)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < s.n_views) {
        const double b = __dadd_rn(s.beta0, __dmul_rn((double)t, s.dbeta));
        double sn, cs;
        sincos(b, &sn, &cs);
        cosb[t] = cs; sinb[t] = sn;
    }
    if (t < s.n_ch) {
        const double g = __dadd_rn(s.gamma0, __dmul_rn((double)t, s.dgamma));
        double sn, cs;
        sincos(g, &sn, &cs);
        cosg[t] = cs; sing[t] = sn;
    }
    if (t < s.n_rows) zdet[t] = __dadd_rn(s.z0, __dmul_rn((double)t, s.dzdet));
}

// y-major layout: columns are rows of the image, minor axis x is already contiguous in the input.
// ym[(k+1)][j][ii], ii = i + 1 in [0, nx+1]: low byte = label(i, j, k), high byte = label(i+1, j, k); 0 outside.
__global__ void repack_ym_kernel(const uint8_t *__restrict__ labels, int nx, int ny, int nz,
                                 uint16_t *__restrict__ ym, int pitch, long long plane)
{
    const long long row = blockIdx.x;                      // k*ny + j
    const int k = (int)(row / ny), j = (int)(row % ny);
    const uint8_t *src = labels + row * nx;
    uint16_t *dst = ym + (long long)(k + 1) * plane + (long long)j * pitch;
    for (int ii = threadIdx.x; ii < nx + 2; ii += blockDim.x) {
        const int i = ii - 1;
        const unsigned lo = (i >= 0 && i < nx) ? src[i] : 0u;
        const unsigned hi = (i + 1 < nx) ? src[i + 1] : 0u;
        dst[ii] = (uint16_t)(lo | (hi << 8));
    }
}

// x-major layout: columns are image columns, minor axis y must become contiguous -> tiled transpose.
// xm[(k+1)][i][jj], jj = j + 1: low byte = label(i, j, k), high byte = label(i, j+1, k).
__global__ void repack_xm_kernel(const uint8_t *__restrict__ labels, int nx, int ny, int nz,
                                 uint16_t *__restrict__ xm, int pitch, long long plane)
{
    __shared__ uint8_t tile[33][36];                       // [j_local (0..32)][i_local], one extra j row
    const int k = blockIdx.z;
    const int i0 = blockIdx.x * 32, j0 = blockIdx.y * 32 - 1;      // tile covers j in [j0, j0+33)
    const uint8_t *src = labels + (long long)k * ny * nx;
    for (int t = threadIdx.y; t < 33; t += blockDim.y) {
        const int j = j0 + t, i = i0 + threadIdx.x;
        tile[t][threadIdx.x] = (j >= 0 && j < ny && i < nx) ? src[(long long)j * nx + i] : (uint8_t)0;
    }
    __syncthreads();
    uint16_t *dst = xm + (long long)(k + 1) * plane;
    for (int t = threadIdx.y; t < 32; t += blockDim.y) {
        const int i = i0 + t;
        const int jj = blockIdx.y * 32 + threadIdx.x;      // = j + 1 with j = j0 + threadIdx.x
        if (i < nx && jj < ny + 2) {
            const unsigned lo = tile[threadIdx.x][t], hi = tile[threadIdx.x + 1][t];
            dst[(long long)i * pitch + jj] = (uint16_t)(lo | (hi << 8));
        }
    }
}

// Skip layouts (2-D scans).  One per (major axis, sign s of the minor slope), indexed like one plane of the pair layouts.
// The word at (column i, jj = j + 1) tells a ray with 0 <= s*m <= 1 whose lower minor index inside column i is j what it
// can meet from there on:
//   bit 15 set  : voxels (i, j) and (i, j+1) carry the same label l = bits 0..6, and so does every voxel the ray can touch
//                 in columns i .. i+S-1, S = bits 8..14 (1..127): the whole stretch is one Siddon run of label l and the
//                 kernel accounts it in a single step;
//   bit 15 clear: bits 0..6 = label(i, j), bits 8..14 = label(i, j+1) -- an ordinary column, split at the crossing.
// "Can touch" is the cone of slopes [0, 1] (times s): at column i+1 the lower index is j or j+s, so
//   S(i, j) = 1 + min(S'(i+1, j), S'(i+1, j+s)),  S' = S where the label is the same and 0 elsewhere,
// a sweep from the last column to the first.  Labels must be < 128 for these layouts to be used (siddon_poly checks K).
constexpr int SKIP_EPT = 4;          // entries per thread of the sweep
constexpr int SKIP_THREADS = 1024;
constexpr int SKIP_MAX_ENTRIES = SKIP_EPT * SKIP_THREADS;

struct SkipBuildArgs {
    const uint16_t *pairs[2];        // plane k = 0 of the x-major / y-major pair layout
    uint16_t *out[4];
    int np[2], nq[2], pitch[2];
};

// The layouts for NEGATIVE slopes are built on the image mirrored along the minor axis (row U = nq-1-j), where the
// slope is positive again: the walk then needs no min/max of the two boundary positions.  Mirrored pair (U, U+1) is the
// byte-swapped original pair (j-1, j) with j = nq-1-U.
// The cone recurrence S(i, j) = 1 + min(S'(i+1, j), S'(i+1, j+1)) equals  S(i, j) = min(R(i, j), 1 + S'(i+1, j+1))  with
// R(i, j) the run of homogeneous words of one label along the columns from (i, j) (proof: unroll the minimum along the
// diagonal; checked against the recurrence on random images).  Both are 1-D scans with no coupling between rows resp.
// diagonals, so the layouts are built by two passes of independent threads -- no barrier per column (the barrier-per-column
// sweep took 0.29 ms for the C2 slice, all of it synchronisation and load latency).
// Pass 1: thread = (layout, entry j), columns swept from the last to the first; R goes into the layout itself.
__global__ void __launch_bounds__(128) skip_runs_kernel(SkipBuildArgs A)
{
    const int lay = blockIdx.y, ax = lay >> 1;
    const bool mirror = (lay & 1) != 0;
    const uint16_t *__restrict__ pairs = A.pairs[ax];
    uint16_t *__restrict__ out = A.out[lay];
    const int np = A.np[ax], nq = A.nq[ax], n_ent = nq + 1, pitch = A.pitch[ax];
    const int jj = blockIdx.x * blockDim.x + threadIdx.x;
    if (jj >= n_ent) return;
    const uint16_t *src = pairs + (mirror ? nq - jj : jj);
    unsigned run = 0, lab = 0xffu;
    constexpr int B = 16;                                  // columns fetched per batch: B independent loads in flight
    for (int i0 = np - 1; i0 >= 0; i0 -= B) {
        unsigned w[B];
#pragma unroll
        for (int b = 0; b < B; ++b) w[b] = i0 - b >= 0 ? src[(long long)(i0 - b) * pitch] : 0u;
#pragma unroll
        for (int b = 0; b < B; ++b) {
            if (i0 - b < 0) break;
            const unsigned b0 = w[b] & 0x7f, b1 = (w[b] >> 8) & 0x7f;
            const bool hom = b0 == b1;
            run = hom ? min(1u + (b0 == lab ? run : 0u), 127u) : 0u;
            lab = hom ? b0 : 0xffu;
            out[(long long)(i0 - b) * pitch + jj] = (uint16_t)run;
        }
    }
}

// Pass 2: thread = (layout, diagonal d = j - i); all threads step through the columns together so that a warp touches
// adjacent entries of one column.  Reads R from the layout and replaces it by the final word.
__global__ void __launch_bounds__(128) skip_cone_kernel(SkipBuildArgs A)
{
    const int lay = blockIdx.y, ax = lay >> 1;
    const bool mirror = (lay & 1) != 0;
    const uint16_t *__restrict__ pairs = A.pairs[ax];
    uint16_t *__restrict__ out = A.out[lay];
    const int np = A.np[ax], nq = A.nq[ax], n_ent = nq + 1, pitch = A.pitch[ax];
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= np + n_ent - 1) return;
    const int d = t - (np - 1);
    unsigned S = 0, lab = 0xffu;                           // state of (i + 1, j + 1)
    constexpr int B = 16;
    for (int i0 = np - 1; i0 >= 0; i0 -= B) {
        unsigned w[B], R[B];
#pragma unroll
        for (int b = 0; b < B; ++b) {
            const int i = i0 - b, jj = i + d;
            const bool in = i >= 0 && jj >= 0 && jj < n_ent;  // (the diagonal has not entered / has left the layout)
            w[b] = in ? pairs[(long long)i * pitch + (mirror ? nq - jj : jj)] : 0u;
            R[b] = in ? out[(long long)i * pitch + jj] : 0u;
        }
#pragma unroll
        for (int b = 0; b < B; ++b) {
            const int i = i0 - b, jj = i + d;
            if (i < 0 || jj < 0 || jj >= n_ent) continue;
            const unsigned b0 = w[b] & 0x7f, b1 = (w[b] >> 8) & 0x7f;
            const unsigned lo = mirror ? b1 : b0, hi = mirror ? b0 : b1;
            const bool hom = lo == hi;
            S = hom ? min(R[b], 1u + (lo == lab ? S : 0u)) : 0u;
            lab = hom ? lo : 0xffu;
            out[(long long)i * pitch + jj] = (uint16_t)(hom ? (0x8000u | (S << 8) | lo) : (lo | (hi << 8)));
        }
    }
}

// 3-D skip layouts.  A cone-beam ray climbs at most mu_z = max |dz/dp| planes per column (about 0.07 for CT cone angles),
// so over a run of S columns it stays within 1 + S mu_z planes of the plane it starts in.  The 2-D sweep above is therefore
// run once per z plane on the labels ANDed over a slab of planes: at level l the word is homogeneous only if the 2 l + 3
// planes k-l-1 .. k+l+1 carry the same label at both voxels of the pair, and runs of that level are capped at
// cap_l = floor((l + 1) / mu_z) columns.  The word stores the longest run over the levels; its meaning for the walk is the
// same as in 2-D (every voxel the ray can touch in the next S columns has label l), whatever its z slope.
#ifndef SKIP_LEVELS_N
#define SKIP_LEVELS_N 4
#endif
constexpr int SKIP_LEVELS = SKIP_LEVELS_N;       // z radii 1..SKIP_LEVELS planes (at most 8: SiddonParams::skip_cap)

// rz [k][i][j] = number of planes on BOTH sides of k (capped at SKIP_LEVELS) that carry the label of voxel (i, j, k);
// outside the volume counts as label 0.  Reads the low bytes of a pair layout, so i is the layout's column (major) index.
__global__ void rz_kernel(const uint16_t *__restrict__ pairs, int np, int nq, int nz, int pitch, long long plane,
                          uint8_t *__restrict__ rz)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= nq) return;
    const uint16_t *src = pairs + plane + (long long)i * pitch + (j + 1);
    uint8_t *dst = rz + (long long)i * nq + j;
    const long long vplane = (long long)np * nq;
    unsigned prev = 0, run = SKIP_LEVELS;                  // below the volume: air, homogeneous
    for (int k = 0; k < nz; ++k) {
        const unsigned l = src[k * plane] & 0xffu;
        run = (l == prev) ? min(run + 1u, (unsigned)SKIP_LEVELS) : 0u;
        dst[k * vplane] = (uint8_t)run;
        prev = l;
    }
    prev = 0; run = SKIP_LEVELS;
    for (int k = nz - 1; k >= 0; --k) {
        const unsigned l = src[k * plane] & 0xffu;
        run = (l == prev) ? min(run + 1u, (unsigned)SKIP_LEVELS) : 0u;
        dst[k * vplane] = (uint8_t)min(run, (unsigned)dst[k * vplane]);
        prev = l;
    }
}

struct SkipBuild3Args {
    const uint16_t *pairs[2];        // whole pair layouts (z aprons included)
    const uint8_t *rz[2];            // [nz][np][nq] per axis
    uint16_t *out[4];
    int np[2], nq[2], pitch[2];
    long long plane[2];
    int cap[SKIP_LEVELS];
};

__global__ void __launch_bounds__(SKIP_THREADS) skip_build3_kernel(SkipBuild3Args A)
{
    extern __shared__ uint16_t sk_buf[];                   // [SKIP_LEVELS][2][n_ent + 2]
    const int lay = blockIdx.x, ax = lay >> 1, k = blockIdx.y;
    const bool mirror = (lay & 1) != 0;
    const int np = A.np[ax], nq = A.nq[ax], n_ent = nq + 1, pitch = A.pitch[ax];
    const uint16_t *__restrict__ pairs = A.pairs[ax] + (long long)(k + 1) * A.plane[ax];
    const uint8_t *__restrict__ rz = A.rz[ax] + (long long)k * np * nq;
    uint16_t *__restrict__ out = A.out[lay] + (long long)(k + 1) * A.plane[ax];
    const int ld = n_ent + 2;
    for (int t = threadIdx.x; t < 2 * SKIP_LEVELS * ld; t += blockDim.x) sk_buf[t] = 0x00ff;
    __syncthreads();
    int cur = 0;
    for (int i = np - 1; i >= 0; --i) {
        unsigned word[SKIP_EPT], rzp[SKIP_EPT];
#pragma unroll
        for (int u = 0; u < SKIP_EPT; ++u) {
            const int jj = threadIdx.x + u * SKIP_THREADS;
            word[u] = 0; rzp[u] = 0;
            if (jj < n_ent) {
                const int oj = mirror ? nq - jj : jj;      // entry of the un-mirrored layout: pair (oj - 1, oj)
                word[u] = pairs[(long long)i * pitch + oj];
                const unsigned ra = oj - 1 >= 0 ? rz[(long long)i * nq + oj - 1] : SKIP_LEVELS;
                const unsigned rb = oj < nq ? rz[(long long)i * nq + oj] : SKIP_LEVELS;
                rzp[u] = min(ra, rb);
            }
        }
#pragma unroll
        for (int u = 0; u < SKIP_EPT; ++u) {
            const int jj = threadIdx.x + u * SKIP_THREADS;
            if (jj < n_ent) {
                const unsigned b0 = word[u] & 0x7f, b1 = (word[u] >> 8) & 0x7f;
                const unsigned lo = mirror ? b1 : b0, hi = mirror ? b0 : b1;
                unsigned best = 0;
#pragma unroll
                for (int l = 0; l < SKIP_LEVELS; ++l) {
                    const uint16_t *prev = sk_buf + (2 * l + cur) * ld + 1;
                    uint16_t *next = sk_buf + (2 * l + (cur ^ 1)) * ld + 1;
                    const unsigned pa = prev[jj], pb = prev[jj + 1];
                    const unsigned sa = (pa & 0xff) == lo ? pa >> 8 : 0u, sb = (pb & 0xff) == lo ? pb >> 8 : 0u;
                    const bool hom = lo == hi && rzp[u] >= (unsigned)(l + 1);
                    const unsigned S = hom ? min(1u + min(sa, sb), (unsigned)A.cap[l]) : 0u;
                    next[jj] = (uint16_t)(hom ? ((S << 8) | lo) : 0x00ffu);
                    best = max(best, S);
                }
                out[(long long)i * pitch + jj] = (uint16_t)(best > 0 ? (0x8000u | (best << 8) | lo) : (lo | (hi << 8)));
            }
        }
        __syncthreads();
        cur ^= 1;
    }
}

// ---------------------------------------------------------------------------------------------------------
// Polyenergetic detection epilogue (forward model of matdecomp.py:116-118 with K materials):
// counts_s = sum_e W[s][e] exp(-sum_m L_m mu[m][e]).

// Tables are staged as mu * EXP_SCALE so that the float path needs no multiply ahead of MUFU.EX2:
// att = 2^(-t) with t = sum_m L_m (mu_m log2 e).  The double path keeps natural units and calls exp().
template <typename R> struct DetMath;
template <> struct DetMath<float> {
    static constexpr float scale = 1.4426950408889634f;
    static __device__ __forceinline__ float att(float t)
    {
        float y;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(-t));
        return y;
    }
    static __device__ __forceinline__ float fma(float a, float b, float c) { return fmaf(a, b, c); }
};
template <> struct DetMath<double> {
    static constexpr double scale = 1.0;
    static __device__ __forceinline__ double att(double t) { return exp(-t); }
    static __device__ __forceinline__ double fma(double a, double b, double c) { return ::fma(a, b, c); }
};

constexpr int MAX_SPEC = 4;

// Lm: per-thread lengths, element m at Lm[m * stride].  mu/W: tables (shared memory).
template <typename R>
__device__ __forceinline__ void detect(const R *Lm, int stride, int K, const R *mu, const R *W, int nE, int nSpec,
                                       R *out /*[nSpec]*/)
{
    R acc[MAX_SPEC];
#pragma unroll
    for (int s = 0; s < MAX_SPEC; ++s) acc[s] = (R)0;
    for (int e = 0; e < nE; ++e) {
        R t = (R)0;
        for (int m = 0; m < K; ++m) t = DetMath<R>::fma(Lm[m * stride], mu[m * nE + e], t);
        const R att = DetMath<R>::att(t);
#pragma unroll
        for (int s = 0; s < MAX_SPEC; ++s)
            if (s < nSpec) acc[s] = DetMath<R>::fma(W[s * nE + e], att, acc[s]);
    }
#pragma unroll
    for (int s = 0; s < MAX_SPEC; ++s) out[s] = acc[s];
}

// Same with the lengths in registers (K <= KT), the common case: one broadcast LDS + one FMA per (m, e).
template <typename R, int KT>
__device__ __forceinline__ void detect_reg(const R (&Lr)[KT], const R *mu, const R *W, int nE, int nSpec, R *out)
{
    R acc[MAX_SPEC];
#pragma unroll
    for (int s = 0; s < MAX_SPEC; ++s) acc[s] = (R)0;
    for (int e = 0; e < nE; ++e) {
        R t = (R)0;
#pragma unroll
        for (int m = 0; m < KT; ++m) t = DetMath<R>::fma(Lr[m], mu[m * nE + e], t);
        const R att = DetMath<R>::att(t);
#pragma unroll
        for (int s = 0; s < MAX_SPEC; ++s)
            if (s < nSpec) acc[s] = DetMath<R>::fma(W[s * nE + e], att, acc[s]);
    }
#pragma unroll
    for (int s = 0; s < MAX_SPEC; ++s) out[s] = acc[s];
}

struct DetectArgs {
    const void *mu, *W;       // [K][nE], [nSpec][nE]
    int K, nE, nSpec;
    void *counts, *logs, *L_out;
};

// Stage mu (zero-padded to KP rows), W and the unattenuated ("air") signal of each spectrum in shared memory.
template <typename R>
__device__ __forceinline__ void stage_tables(const DetectArgs &d, int KP, R *mu_s, R *W_s, R *air_s)
{
    const R *mu = (const R *)d.mu, *W = (const R *)d.W;
    for (int t = threadIdx.x; t < KP * d.nE; t += blockDim.x) mu_s[t] = t < d.K * d.nE ? mu[t] * DetMath<R>::scale : (R)0;
    for (int t = threadIdx.x; t < d.nSpec * d.nE; t += blockDim.x) W_s[t] = W[t];
    __syncthreads();
    if (threadIdx.x < d.nSpec) {
        R a = (R)0;
        for (int e = 0; e < d.nE; ++e) a += W_s[threadIdx.x * d.nE + e];
        air_s[threadIdx.x] = a;
    }
    __syncthreads();
}

template <typename R> __device__ __forceinline__ R neg_log_ratio(R c, R air);
template <> __device__ __forceinline__ float neg_log_ratio<float>(float c, float air) { return -logf(c / air); }
template <> __device__ __forceinline__ double neg_log_ratio<double>(double c, double air) { return -log(c / air); }

// ---------------------------------------------------------------------------------------------------------
// Production kernel: one ray per thread, adjacent lanes = adjacent channels (so a warp reads adjacent minor-axis
// voxels of the same column: one or two 128-byte lines per request in the pair layout).
//
// Per-material column weights accumulate in shared memory, acc[label][thread].  Whole columns whose two (2-D) or four
// (3-D) candidate voxels all carry the label of the current RUN never touch it: they bump a register counter that is
// flushed when the run ends.  Inside homogeneous regions -- most of any anatomy -- a column then costs one FMA, one
// floor, one 2-byte load and a compare; columns are processed four at a time so that the four loads are in flight
// together.

template <typename R> struct AccSink {
    R *acc;            // acc[label * BLOCK + tid] lives in shared memory; this points at [tid]
    int stride;
    __device__ __forceinline__ void operator()(int, int, int, int label, R w) const
    {
        R *p = acc + label * stride;
        *p = Rn<R>::add(*p, w);
    }
};

template <typename R> struct RunState {
    unsigned curp;     // label of the current run replicated in both bytes of a pair; 0xffffffff = no run
    R run;             // whole columns in the run (an exact integer)
};

template <typename R>
__device__ __forceinline__ void run_flush(RunState<R> &st, const AccSink<R> &sink)
{
    if (st.run > (R)0) sink(0, 0, 0, (int)(st.curp & 0xff), st.run);
    st.run = (R)0;
}

// A column that is not (known to be) inside the current run: close the run, account the column voxel by voxel, and
// start a new run if the column itself is homogeneous.
template <typename R, bool IS3D>
__device__ __forceinline__ void slow_column(const RayLoop<R> &L, int i, unsigned pr, unsigned pr1, RunState<R> &st,
                                            AccSink<R> &sink)
{
    run_flush(st, sink);
    slab_column<R, IS3D>(L, i, sink);
    const unsigned lo = pr & 0xff, homog = lo | (lo << 8);
    st.curp = (pr == homog && (!IS3D || pr1 == homog)) ? homog : 0xffffffffu;
}

// Interior columns [ia, ib] (each fully inside the ray's clipped range, so every column weighs exactly 1).
// q at a column boundary is the same expression fma(m, boundary - i_c, cq) that slab_column evaluates, so the
// voxel choice is bit-identical to the general routine (and to the CPU twin).
template <typename R, bool IS3D, bool MPOS, int CH>
__device__ __forceinline__ void walk_interior(const RayLoop<R> &L, int ia, int ib, RunState<R> &st, AccSink<R> &sink)
{
    typedef Rn<R> A;
    R p = (R)(ia - L.i_c);
    int jprev = FloorInt<R>::of(A::fma(L.m, p, L.cq));
    int kprev = IS3D ? FloorInt<R>::of(A::fma(L.mz, p, L.cw)) : 0;
    const uint16_t *col = L.pairs + (long long)ia * L.pitch;
    int i = ia;
    for (; i + CH - 1 <= ib; i += CH) {
        unsigned pr[CH], pr1[CH];
#pragma unroll
        for (int u = 0; u < CH; ++u) {
            p = A::add(p, (R)1);
            const int jb = FloorInt<R>::of(A::fma(L.m, p, L.cq));
            int j = MPOS ? jprev : jb;                                   // = floor(min(q_a, q_b))
            jprev = jb;
            j = min(max(j, L.jmin), L.jmax);
            const uint16_t *e = col + (long long)u * L.pitch + j;
            if (IS3D) {
                const int kb = FloorInt<R>::of(A::fma(L.mz, p, L.cw));
                int k = L.mzpos ? kprev : kb;
                kprev = kb;
                k = min(max(k, L.kmin), L.kmax);
                e += (long long)k * L.plane;
                pr1[u] = e[L.plane];
            } else {
                pr1[u] = 0;
            }
            pr[u] = *e;
        }
#pragma unroll
        for (int u = 0; u < CH; ++u) {
            if (pr[u] == st.curp && (!IS3D || pr1[u] == st.curp)) st.run = A::add(st.run, (R)1);
            else slow_column<R, IS3D>(L, i + u, pr[u], pr1[u], st, sink);
        }
        col += (long long)CH * L.pitch;
    }
    if (CH > 1 && i <= ib) walk_interior<R, IS3D, MPOS, 1>(L, i, ib, st, sink);
}

template <typename R, bool IS3D>
__device__ __forceinline__ void walk_ray(const RayLoop<R> &L, int i0, int i1, AccSink<R> &sink)
{
    RunState<R> st{0xffffffffu, (R)0};
    slab_column<R, IS3D>(L, i0, sink);                                  // first and last column may be partial
    if (i1 > i0) {
        if (i1 - 1 >= i0 + 1) {
            if (L.mpos) walk_interior<R, IS3D, true, 4>(L, i0 + 1, i1 - 1, st, sink);
            else        walk_interior<R, IS3D, false, 4>(L, i0 + 1, i1 - 1, st, sink);
            run_flush(st, sink);
        }
        slab_column<R, IS3D>(L, i1, sink);
    }
}

template <typename R, bool IS3D, int KT>
__global__ void __launch_bounds__(128) siddon_poly_kernel(SiddonParams P, int view_begin, long long n_rays, DetectArgs d)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int KP = KT > 0 ? KT : d.K;
    R *acc_s = (R *)smem_raw;                              // [KP][blockDim]
    R *mu_s = acc_s + (size_t)KP * blockDim.x;             // [KP][nE]
    R *W_s = mu_s + (size_t)KP * d.nE;                     // [nSpec][nE]
    R *air_s = W_s + (size_t)d.nSpec * d.nE;               // [nSpec]
    stage_tables<R>(d, KP, mu_s, W_s, air_s);

    const int tid = threadIdx.x;
    for (int m = 0; m < KP; ++m) acc_s[m * blockDim.x + tid] = (R)0;

    const long long ray = (long long)blockIdx.x * blockDim.x + tid;
    if (ray >= n_rays) return;
    const int c = (int)(ray % P.scan.n_ch);
    const long long line = ray / P.scan.n_ch;
    const int r = (int)(line % P.scan.n_rows);
    const int v = view_begin + (int)(line / P.scan.n_rows);

    RaySetup s;
    ray_setup<IS3D>(P, v, r, c, s);
    R seglen = (R)0;
    if (s.hit) {
        RayLoop<R> L;
        L.init(P, s);
        AccSink<R> sink{acc_s + tid, (int)blockDim.x};
        walk_ray<R, IS3D>(L, s.i0, s.i1, sink);
        seglen = (R)s.seglen;
    }

    R out[MAX_SPEC];
    if (KT > 0) {
        R Lr[KT > 0 ? KT : 1];
#pragma unroll
        for (int m = 0; m < (KT > 0 ? KT : 1); ++m) Lr[m] = acc_s[m * blockDim.x + tid] * seglen;
        if (d.L_out) {
            R *Lo = (R *)d.L_out + ray * d.K;
#pragma unroll
            for (int m = 0; m < (KT > 0 ? KT : 1); ++m)
                if (m < d.K) Lo[m] = Lr[m];
        }
        detect_reg<R, (KT > 0 ? KT : 1)>(Lr, mu_s, W_s, d.nE, d.nSpec, out);
    } else {
        for (int m = 0; m < d.K; ++m) acc_s[m * blockDim.x + tid] *= seglen;
        if (d.L_out) {
            R *Lo = (R *)d.L_out + ray * d.K;
            for (int m = 0; m < d.K; ++m) Lo[m] = acc_s[m * blockDim.x + tid];
        }
        detect<R>(acc_s + tid, blockDim.x, d.K, mu_s, W_s, d.nE, d.nSpec, out);
    }
#pragma unroll
    for (int sp = 0; sp < MAX_SPEC; ++sp) {
        if (sp < d.nSpec) {
            ((R *)d.counts)[sp * n_rays + ray] = out[sp];
            if (d.logs) ((R *)d.logs)[sp * n_rays + ray] = neg_log_ratio<R>(out[sp], air_s[sp]);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// 2-D production kernel over the skip layouts: one ray per thread, each lane walks its own ray and takes whole
// homogeneous stretches in one step (about 28 steps per ray instead of 450 columns on the C2 phantom).

// floor(x) as an integer and as a real, and small non-negative integers as reals, without the conversion unit in
// single precision (the quarter-rate F2I/I2F would otherwise bound the loop): x + 1.5*2^23 rounded towards -inf has
// floor(x) in its low mantissa bits, exactly, for |x| < 2^22.
template <typename R> struct FastInt;
template <> struct FastInt<float> {
    static constexpr float MAGIC = 12582912.0f;
    static __device__ __forceinline__ float bias(float x) { return __fadd_rd(x, MAGIC); }          // floor(x) + MAGIC
    static __device__ __forceinline__ float unbias(float t) { return __fsub_rn(t, MAGIC); }
    static __device__ __forceinline__ int to_int(float t) { return __float_as_int(t) - 0x4B400000; }
    static __device__ __forceinline__ float biased(int j) { return __int_as_float(j + 0x4B400000); }
    static __device__ __forceinline__ float real(int n) { return __fsub_rn(__int_as_float(n + 0x4B400000), MAGIC); }
};
template <> struct FastInt<double> {
    static __device__ __forceinline__ double bias(double x) { return floor(x); }
    static __device__ __forceinline__ double unbias(double t) { return t; }
    static __device__ __forceinline__ int to_int(double t) { return __double2int_rd(t); }
    static __device__ __forceinline__ double biased(int j) { return (double)j; }
    static __device__ __forceinline__ double real(int n) { return (double)n; }
};

// Packed single-precision pairs (FFMA2 on sm_100a: two FMAs per issue slot).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float x, float y)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(x), "f"(y));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float &x, float &y) { asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(v)); }
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

// Single-precision detection with the tables packed by ENERGY PAIRS in shared memory:
//   mu2 [nE2][KT]  float2 = mu_m * log2(e) at energies 2 e2, 2 e2 + 1  (rows of materials -> LDS.128 of two materials)
//   W2  [nE2][NSP] float2 = W_s at energies 2 e2, 2 e2 + 1
// An odd nE is padded with a zero-weight bin.  Per energy pair: KT FFMA2 + 2 MUFU.EX2 + NSP FFMA2.
// PM: bit h set = material pair (2h, 2h+1) takes part; a cleared bit promises that both lengths are zero.
template <int KT, int NSP, unsigned PM>
__device__ __forceinline__ void detect_packed_pm(const float (&Lr)[KT], const f32x2 *mu2, const f32x2 *W2, int nE2, float *out)
{
    f32x2 L2[KT], acc[NSP];
#pragma unroll
    for (int m = 0; m < KT; ++m) L2[m] = pack2(Lr[m], Lr[m]);
#pragma unroll
    for (int sp = 0; sp < NSP; ++sp) acc[sp] = 0ull;
#pragma unroll 2
    for (int e2 = 0; e2 < nE2; ++e2) {
        const ulonglong2 *row = (const ulonglong2 *)(mu2 + (size_t)e2 * KT);
        f32x2 t0 = 0ull, t1 = 0ull;
#pragma unroll
        for (int h = 0; h < KT / 2; ++h) {
            if (!((PM >> h) & 1u)) continue;                 // compile time: adding 0 * mu would not change t0, t1
            const ulonglong2 q = row[h];
            t0 = ffma2(L2[2 * h], q.x, t0);
            t1 = ffma2(L2[2 * h + 1], q.y, t1);
        }
        float xa, ya, xb, yb;
        unpack2(t0, xa, ya);
        unpack2(t1, xb, yb);
        const f32x2 att = pack2(DetMath<float>::att(xa + xb), DetMath<float>::att(ya + yb));
        const ulonglong2 *wrow = (const ulonglong2 *)(W2 + (size_t)e2 * NSP);
#pragma unroll
        for (int h = 0; h < NSP / 2; ++h) {
            const ulonglong2 q = wrow[h];
            acc[2 * h] = ffma2(q.x, att, acc[2 * h]);
            acc[2 * h + 1] = ffma2(q.y, att, acc[2 * h + 1]);
        }
    }
#pragma unroll
    for (int sp = 0; sp < MAX_SPEC; ++sp) {
        float x = 0.f, y = 0.f;
        if (sp < NSP) unpack2(acc[sp], x, y);
        out[sp] = x + y;
    }
}

// Material sparsity: the 32 rays of a warp are neighbouring channels of one view and cross the same few materials (C2
// phantom: 3.8 of 8 labels, 2.45 of the 4 label pairs per warp on average; one warp in five sees air only).  A pair whose
// two lengths are zero in EVERY lane adds exact zeros to the exponent, so the warp votes once per ray batch and runs the
// instance of the loop that leaves those pairs out -- bit-identical results, ~40 % fewer table loads and FFMA2.
template <int KT, int NSP>
__device__ __forceinline__ void detect_packed(const float (&Lr)[KT], const f32x2 *mu2, const f32x2 *W2, int nE2, float *out)
{
    if constexpr (KT == 8) {
        const unsigned lanes = __activemask();
        unsigned pm = 0;
#pragma unroll
        for (int h = 1; h < 4; ++h)
            if (__any_sync(lanes, Lr[2 * h] != 0.f || Lr[2 * h + 1] != 0.f)) pm |= 1u << (h - 1);
        switch (pm) {
        case 0: detect_packed_pm<KT, NSP, 0x1u>(Lr, mu2, W2, nE2, out); break;
        case 1: detect_packed_pm<KT, NSP, 0x3u>(Lr, mu2, W2, nE2, out); break;
        case 2: detect_packed_pm<KT, NSP, 0x5u>(Lr, mu2, W2, nE2, out); break;
        case 3: detect_packed_pm<KT, NSP, 0x7u>(Lr, mu2, W2, nE2, out); break;
        case 4: detect_packed_pm<KT, NSP, 0x9u>(Lr, mu2, W2, nE2, out); break;
        case 5: detect_packed_pm<KT, NSP, 0xBu>(Lr, mu2, W2, nE2, out); break;
        case 6: detect_packed_pm<KT, NSP, 0xDu>(Lr, mu2, W2, nE2, out); break;
        default: detect_packed_pm<KT, NSP, 0xFu>(Lr, mu2, W2, nE2, out); break;
        }
    } else {
        detect_packed_pm<KT, NSP, 0xFFFFFFFFu>(Lr, mu2, W2, nE2, out);
    }
}

template <int KT, int NSP>
__device__ __forceinline__ void stage_packed(const DetectArgs &d, f32x2 *mu2, f32x2 *W2, float *air_s)
{
    const float *mu = (const float *)d.mu, *W = (const float *)d.W;
    const int nE2 = (d.nE + 1) / 2;
    for (int t = threadIdx.x; t < nE2 * KT; t += blockDim.x) {
        const int e2 = t / KT, m = t % KT, e = 2 * e2;
        const float a = (m < d.K) ? mu[m * d.nE + e] * DetMath<float>::scale : 0.f;
        const float b = (m < d.K && e + 1 < d.nE) ? mu[m * d.nE + e + 1] * DetMath<float>::scale : 0.f;
        mu2[t] = pack2(a, b);
    }
    for (int t = threadIdx.x; t < nE2 * NSP; t += blockDim.x) {
        const int e2 = t / NSP, sp = t % NSP, e = 2 * e2;
        const float a = (sp < d.nSpec) ? W[sp * d.nE + e] : 0.f;
        const float b = (sp < d.nSpec && e + 1 < d.nE) ? W[sp * d.nE + e + 1] : 0.f;
        W2[t] = pack2(a, b);
    }
    if (threadIdx.x < MAX_SPEC) {
        float a = 0.f;
        if (threadIdx.x < d.nSpec)
            for (int e = 0; e < d.nE; ++e) a += W[threadIdx.x * d.nE + e];
        air_s[threadIdx.x] = a;
    }
    __syncthreads();
}

// Walk of one ray over its skip layout.  Everything is expressed in the ray's own (possibly mirrored) minor coordinate,
// in which the slope am = |m| is non-negative: q(p) = am p + c0, lower voxel j = floor(q(a)), upper voxel j + 1 entered
// where q crosses j + 1.  W points at (column 0, mirrored row of j_c) of the layout for this ray's axis and slope sign.
// No clamp on j: the ray is clipped to the volume, so j stays inside [-1, nq] -- the zero apron and the row padding of
// the layout.
// SINK: what else happens at every step.  Production passes NoStepSink (nothing, compiled out); the walk-trace kernel passes
// a recorder, so that the sequence checked against the oracle is the sequence of THIS code, not of a look-alike.
struct NoStepSink {
    static constexpr bool active = false;
    template <typename R>
    __device__ __forceinline__ void operator()(int, int, int, const int *, const int *, const int *, const R *) const {}
};

// One 16-bit skip word, sign-extended, through the read-only path; the address is one IMAD.WIDE from a 32-bit index.
__device__ __forceinline__ int skip_word(const uint16_t *__restrict__ W, int o)
{
    int w;
    asm("{ .reg .u64 p; .reg .s16 h; mad.wide.s32 p, %2, 2, %1; ld.global.nc.s16 h, [p]; cvt.s32.s16 %0, h; }"
        : "=r"(w) : "l"(W), "r"(o));
    return w;
}

template <typename R, int STRIDE, typename SINK>
__device__ __forceinline__ void walk_skip(const uint16_t *__restrict__ W, int pitch, int i0, int i1, int i_c, R am, R c0,
                                          R ua, R ub, R inv_am, R *acc, int stride, SINK &sink)
{
    typedef Rn<R> A;
    typedef FastInt<R> F;
    const int st = STRIDE > 0 ? STRIDE : stride;                          // compile-time accumulator stride when known
    const R c0m1 = A::sub(c0, (R)1);
    R fi = A::sub(F::real(i0 - i_c + (1 << 20)), (R)(1 << 20));          // column coordinate of column i, as a real
    const R fend = A::sub(F::real(i1 - i_c + (1 << 20)), (R)(1 << 20));
    int ofs = i0 * pitch;
    while (fi <= fend) {
        const R a = A::max(fi, ua);
        const R b1 = A::min(A::add(fi, (R)1), ub);
        const R qa = A::fma(am, a, c0);
        const R qbm1 = A::fma(am, b1, c0m1);                              // q(b1) - 1
        const R t = F::bias(qa);
        const int w = skip_word(W, ofs + F::to_int(t));
        const bool hom = w < 0;
        const int f0 = w & 0x7f, f1 = (w >> 8) & 0x7f;
        const int step = hom ? f1 : 1;
        const R fnext = A::add(fi, F::real(step));
        const R wtot = A::sub(A::min(fnext, ub), a);
        // share of the upper voxel: the column length beyond the crossing of minor plane j + 1
        R whi = A::max(A::mul(A::sub(qbm1, F::unbias(t)), inv_am), (R)0);
        whi = hom ? (R)0 : whi;
        const R wlo = A::sub(wtot, whi);
        R *p0 = acc + f0 * st, *p1 = acc + (hom ? f0 : f1) * st;
        const R v0 = *p0, v1 = *p1;
        *p1 = A::add(v1, whi);                                            // p1 == p0 only when whi == 0: order matters
        *p0 = A::add(v0, wlo);
        if constexpr (SINK::active) {
            const int jr = F::to_int(t);
            const int js[3] = {jr, jr + 1, 0}, ks[3] = {0, 0, 0}, ls[3] = {f0, hom ? f0 : f1, 0};
            const R ws[3] = {wlo, whi, (R)0};
            sink(ofs / pitch, hom ? step : 0, 2, js, ks, ls, ws);
        }
        ofs += step * pitch;
        fi = fnext;
    }
}

template <typename R, int KT, int NSP>
__global__ void __launch_bounds__(128) siddon_skip_kernel(SiddonParams P, int view_begin, long long n_rays, DetectArgs d)
{
    constexpr bool PACKED = (sizeof(R) == 4) && KT > 0;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int KP = KT > 0 ? KT : d.K;
    const int tid = threadIdx.x;
    R *acc_s = (R *)smem_raw;                              // [KP][blockDim]
    unsigned char *tab = smem_raw + sizeof(R) * (size_t)KP * blockDim.x;
    // generic tables
    R *mu_s = (R *)tab, *W_s = mu_s + (size_t)KP * d.nE, *air_s = W_s + (size_t)d.nSpec * d.nE;
    // packed tables (float, KT > 0)
    const int nE2 = (d.nE + 1) / 2;
    f32x2 *mu2 = (f32x2 *)tab, *W2 = mu2 + (size_t)nE2 * (KT > 0 ? KT : 1);
    if (PACKED) {
        air_s = (R *)(W2 + (size_t)nE2 * NSP);
        stage_packed<(KT > 0 ? KT : 2), NSP>(d, mu2, W2, (float *)air_s);
    } else {
        stage_tables<R>(d, KP, mu_s, W_s, air_s);
    }

    const long long n_blk = (n_rays + blockDim.x - 1) / blockDim.x;
    for (long long blk = blockIdx.x; blk < n_blk; blk += gridDim.x) {
        const long long ray = blk * blockDim.x + tid;
        if (ray >= n_rays) continue;
        const int c = (int)(ray % P.scan.n_ch);
        const int v = view_begin + (int)(ray / P.scan.n_ch);
        for (int m = 0; m < KP; ++m) acc_s[m * blockDim.x + tid] = (R)0;

        RaySetup s;
        ray_setup<false>(P, v, 0, c, s);
        R seglen = (R)0;
        if (s.hit) {
            const bool mpos = s.m >= 0.0;
            const int pitch = s.xmajor ? P.pitch_xm : P.pitch_ym;
            const uint16_t *W = P.skip[(s.xmajor ? 0 : 2) + (mpos ? 0 : 1)] + (mpos ? s.j_c : s.nq - 1 - s.j_c) + 1;
            const double c0 = mpos ? s.cq : 1.0 - s.cq;
            NoStepSink none;
            if (blockDim.x == 128)
                walk_skip<R, 128>(W, pitch, s.i0, s.i1, s.i_c, (R)fabs(s.m), (R)c0, (R)s.ua, (R)s.ub, (R)s.inv_am, acc_s + tid, 128, none);
            else
                walk_skip<R, 0>(W, pitch, s.i0, s.i1, s.i_c, (R)fabs(s.m), (R)c0, (R)s.ua, (R)s.ub, (R)s.inv_am, acc_s + tid,
                                (int)blockDim.x, none);
            seglen = (R)s.seglen;
        }

        R out[MAX_SPEC];
        if (KT > 0) {
            R Lr[KT > 0 ? KT : 1];
#pragma unroll
            for (int m = 0; m < (KT > 0 ? KT : 1); ++m) Lr[m] = acc_s[m * blockDim.x + tid] * seglen;
            if (d.L_out) {
                R *Lo = (R *)d.L_out + ray * d.K;
#pragma unroll
                for (int m = 0; m < (KT > 0 ? KT : 1); ++m)
                    if (m < d.K) Lo[m] = Lr[m];
            }
            if constexpr (PACKED) detect_packed<KT, NSP>(Lr, mu2, W2, nE2, out);
            else detect_reg<R, (KT > 0 ? KT : 1)>(Lr, mu_s, W_s, d.nE, d.nSpec, out);
        } else {
            for (int m = 0; m < d.K; ++m) acc_s[m * blockDim.x + tid] *= seglen;
            if (d.L_out) {
                R *Lo = (R *)d.L_out + ray * d.K;
                for (int m = 0; m < d.K; ++m) Lo[m] = acc_s[m * blockDim.x + tid];
            }
            detect<R>(acc_s + tid, blockDim.x, d.K, mu_s, W_s, d.nE, d.nSpec, out);
        }
#pragma unroll
        for (int sp = 0; sp < MAX_SPEC; ++sp) {
            if (sp < d.nSpec) {
                ((R *)d.counts)[sp * n_rays + ray] = out[sp];
                if (d.logs) ((R *)d.logs)[sp * n_rays + ray] = neg_log_ratio<R>(out[sp], air_s[sp]);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// 3-D production kernel over the 3-D skip layouts (multi-row / cone-beam scans).  Same walk as walk_skip with the z
// coordinate carried along: w(p) = mz p + cw, plane k = floor(min(w(a), w(b))).  A homogeneous word covers every plane the
// ray can reach during its run (see skip_build3_kernel), so it is accounted in one step exactly as in 2-D; an ordinary
// column is split at its minor-plane crossing and its z-plane crossing into at most three pieces, whose labels come from the
// words of planes k and k + 1 (a homogeneous word of plane k + 1 simply names the label of both voxels of its pair).

__device__ __forceinline__ void skip_labels(int w, int &lo, int &hi)
{
    lo = w & 0x7f;
    hi = w < 0 ? lo : (w >> 8) & 0x7f;
}

template <typename R> __device__ __forceinline__ void acc_add(R *acc, int stride, int label, R w)
{
    R *p = acc + label * stride;
    *p = Rn<R>::add(*p, w);
}

// Every step serves all lanes of the warp branch-free, but how much of the general step is needed is decided per step by
// a warp vote (measured on C4, tools/stat_sid3_steps.py: 83 % of the steps are homogeneous stretches; an ordinary column
// changes z plane in 7 % of the cases):
//   tier 1 -- no active lane has an ordinary column in which its ray changes z plane: the ray of every lane stays in plane k,
//             the step is the 2-D step on the word of plane k (two pieces: before / after the minor-plane crossing);
//   tier 2 -- otherwise: the word of plane k + 1 is fetched as well and the column is split into three pieces.
// Lanes on a homogeneous stretch pass through either tier with zero weights for the pieces they do not have.
// STRIDE: distance of the per-material accumulators in shared memory (elements); 0 = `stride` at run time.
template <typename R, int STRIDE, typename SINK>
__device__ __forceinline__ void walk_skip3(const uint16_t *__restrict__ W, int pitch, int plane, int i0, int i1, int i_c, R am,
                                           R c0, R mz, R cw, R ua, R ub, R inv_am, R inv_amz, R *acc, int stride, SINK &sink)
{
    typedef Rn<R> A;
    typedef FastInt<R> F;
    const int st = STRIDE > 0 ? STRIDE : stride;
    const R c0m1 = A::sub(c0, (R)1);
    const bool mzpos = mz >= (R)0;
    R fi = A::sub(F::real(i0 - i_c + (1 << 20)), (R)(1 << 20));
    const R fend = A::sub(F::real(i1 - i_c + (1 << 20)), (R)(1 << 20));
    int ofs = i0 * pitch;
    while (fi <= fend) {
        const R a = A::max(fi, ua);
        const R b1 = A::min(A::add(fi, (R)1), ub);
        const R qa = A::fma(am, a, c0);
        const R qbm1 = A::fma(am, b1, c0m1);
        const R wa = A::fma(mz, a, cw), wb = A::fma(mz, b1, cw);
        const R wmin = A::min(wa, wb), wmax = A::max(wa, wb);
        const R tj = F::bias(qa), tk = F::bias(wmin), tkx = F::bias(wmax);
        const int o = ofs + F::to_int(tj) + F::to_int(tk) * plane;
        const int w0 = skip_word(W, o);
        const bool hom = w0 < 0;
        const int l00 = w0 & 0x7f, f1 = (w0 >> 8) & 0x7f;
        const int step = hom ? f1 : 1;
        const int l01 = hom ? l00 : f1;
        const R fnext = A::add(fi, F::real(step));
        const R wtot = A::sub(A::min(fnext, ub), a);
        // column length beyond the crossing of minor plane j + 1
        R qhi = A::min(A::max(A::mul(A::sub(qbm1, F::unbias(tj)), inv_am), (R)0), wtot);
        qhi = hom ? (R)0 : qhi;
        const bool zsplit = !hom && (tkx != tk);                           // an ordinary column in which the ray changes plane
        if (!__any_sync(__activemask(), zsplit)) {
            // Bit for bit what tier 2 books for a lane whose ray stays in its plane (so that a ray's result does not depend on
            // which rays share its warp -- a view range on its own equals the same views inside the whole scan): first the
            // lower voxel gets tq = wtot - qhi, then the upper one wtot - tq.  Both sums are read before either is written
            // (one round trip to shared memory instead of two); in 3-D an ordinary word may carry the same label twice (its
            // pair is uniform but the planes around it are not), so the two addresses can coincide with a non-zero second
            // piece: the second sum then starts from the first.
            const R tq = A::sub(wtot, qhi), whi = A::sub(wtot, tq);
            R *p0 = acc + l00 * st, *p1 = acc + l01 * st;
            const R v0 = *p0, v1 = *p1;
            const R s0 = A::add(v0, tq);
            *p0 = s0;
            *p1 = A::add(l01 == l00 ? s0 : v1, whi);
            if constexpr (SINK::active) {
                const int jr = F::to_int(tj), kr = F::to_int(tk);
                const int js[3] = {jr, jr + 1, jr + 1}, ks[3] = {kr, kr, kr}, ls[3] = {l00, l01, l01};
                const R ws[3] = {tq, whi, (R)0};
                sink(ofs / pitch, hom ? step : 0, 3, js, ks, ls, ws);
            }
        } else {
            const int w1 = skip_word(W, o + plane);
            int l10, l11;                                                  // label(j, k+1), (j+1, k+1)
            skip_labels(w1, l10, l11);
            // column length spent in plane k + 1
            R zhi = A::min(A::max(A::mul(A::sub(A::sub(wmax, (R)1), F::unbias(tk)), inv_amz), (R)0), wtot);
            zhi = hom ? (R)0 : zhi;
            const R tq = A::sub(wtot, qhi);                                // p-length before the minor crossing (am >= 0)
            const R tz = mzpos ? A::sub(wtot, zhi) : zhi;                  // p-length before the z crossing
            const R t1 = A::min(tq, tz), t2 = A::max(tq, tz);
            const bool qfirst = tq < tz;
            const int lA0 = mzpos ? l00 : l10, lA1 = mzpos ? l01 : l11;    // plane the ray is in first / second along +p
            const int lB0 = mzpos ? l10 : l00, lB1 = mzpos ? l11 : l01;
            acc_add<R>(acc, st, hom ? l00 : lA0, t1);
            acc_add<R>(acc, st, hom ? l00 : (qfirst ? lA1 : lB0), A::sub(t2, t1));
            acc_add<R>(acc, st, hom ? l00 : lB1, A::sub(wtot, t2));
            if constexpr (SINK::active) {
                const int jr = F::to_int(tj), kr = F::to_int(tk);
                const int kA = mzpos ? kr : kr + 1, kB = mzpos ? kr + 1 : kr;
                const int js[3] = {jr, qfirst ? jr + 1 : jr, jr + 1}, ks[3] = {kA, qfirst ? kA : kB, kB};
                const int ls[3] = {hom ? l00 : lA0, hom ? l00 : (qfirst ? lA1 : lB0), hom ? l00 : lB1};
                const R ws[3] = {t1, A::sub(t2, t1), A::sub(wtot, t2)};
                sink(ofs / pitch, hom ? step : 0, 3, js, ks, ls, ws);
            }
        }
        ofs += step * pitch;
        fi = fnext;
    }
}

// MINB: resident CTAs per SM the register allocation is held to (9: the 56 registers the kernel takes unconstrained).  Volumes whose
// layouts are far larger than L2 walk on DRAM latency and gain from more warps (1024^3: 127.6 -> 121.7 ms at 10 CTAs);
// L2-resident ones lose (512^3: 11.7 -> 12.4 ms) -- launch_skip chooses.
template <typename R, int KT, int NSP, int MINB>
__global__ void __launch_bounds__(128, MINB) siddon_skip3_kernel(SiddonParams P, int view_begin, long long n_rays, DetectArgs d)
{
    constexpr bool PACKED = (sizeof(R) == 4) && KT > 0;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int KP = KT > 0 ? KT : d.K;
    const int tid = threadIdx.x;
    R *acc_s = (R *)smem_raw;                              // [KP][blockDim]
    unsigned char *tab = smem_raw + sizeof(R) * (size_t)KP * blockDim.x;
    R *mu_s = (R *)tab, *W_s = mu_s + (size_t)KP * d.nE, *air_s = W_s + (size_t)d.nSpec * d.nE;
    const int nE2 = (d.nE + 1) / 2;
    f32x2 *mu2 = (f32x2 *)tab, *W2 = mu2 + (size_t)nE2 * (KT > 0 ? KT : 1);
    if (PACKED) {
        air_s = (R *)(W2 + (size_t)nE2 * NSP);
        stage_packed<(KT > 0 ? KT : 2), NSP>(d, mu2, W2, (float *)air_s);
    } else {
        stage_tables<R>(d, KP, mu_s, W_s, air_s);
    }

    const long long n_blk = (n_rays + blockDim.x - 1) / blockDim.x;
    for (long long blk = blockIdx.x; blk < n_blk; blk += gridDim.x) {
        // Launch order: bands of `row_band` detector rows outermost, then views, then the rows of the band, then channels.
        // The CTAs resident at any one time then work on a few rows of a dozen consecutive views -- a slab of a few dozen
        // z planes of ONE layout, which stays in L2 from view to view -- instead of on all rows of one view, i.e. the whole
        // volume (1024^3: 272 GB of DRAM reads per scan and an L2 hit rate of 39 % in view-major order, profiles/r02x).
        const long long q_ray = blk * blockDim.x + tid;
        if (q_ray >= n_rays) continue;
        const int c = (int)(q_ray % P.scan.n_ch);
        const long long q = q_ray / P.scan.n_ch;                           // position of the detector line in launch order
        const long long per_band = (n_rays / P.scan.n_ch / P.scan.n_rows) * P.row_band;   // lines of a band: views x rows
        const int band = (int)(q / per_band);
        const long long rem = q - band * per_band;
        const int v_rel = (int)(rem / P.row_band);
        const int r = band * P.row_band + (int)(rem - (long long)v_rel * P.row_band);
        const int v = view_begin + v_rel;
        const long long ray = ((long long)v_rel * P.scan.n_rows + r) * P.scan.n_ch + c;     // where its results go
        for (int m = 0; m < KP; ++m) acc_s[m * blockDim.x + tid] = (R)0;

        RaySetup s;
        ray_setup<true>(P, v, r, c, s);
        R seglen = (R)0;
        if (s.hit) {
            const bool mpos = s.m >= 0.0;
            const int pitch = s.xmajor ? P.pitch_xm : P.pitch_ym;
            const int plane = (int)(s.xmajor ? P.plane_xm : P.plane_ym);
            const uint16_t *W = P.skip[(s.xmajor ? 0 : 2) + (mpos ? 0 : 1)] + (long long)(s.k_c + 1) * plane +
                                (mpos ? s.j_c : s.nq - 1 - s.j_c) + 1;
            const double c0 = mpos ? s.cq : 1.0 - s.cq;
            NoStepSink none;
            if (blockDim.x == 128)
                walk_skip3<R, 128>(W, pitch, plane, s.i0, s.i1, s.i_c, (R)fabs(s.m), (R)c0, (R)s.mz, (R)s.cw, (R)s.ua, (R)s.ub,
                                   (R)s.inv_am, (R)s.inv_amz, acc_s + tid, 128, none);
            else
                walk_skip3<R, 0>(W, pitch, plane, s.i0, s.i1, s.i_c, (R)fabs(s.m), (R)c0, (R)s.mz, (R)s.cw, (R)s.ua, (R)s.ub,
                                 (R)s.inv_am, (R)s.inv_amz, acc_s + tid, (int)blockDim.x, none);
            seglen = (R)s.seglen;
        }

        R out[MAX_SPEC];
        if (KT > 0) {
            R Lr[KT > 0 ? KT : 1];
#pragma unroll
            for (int m = 0; m < (KT > 0 ? KT : 1); ++m) Lr[m] = acc_s[m * blockDim.x + tid] * seglen;
            if (d.L_out) {
                R *Lo = (R *)d.L_out + ray * d.K;
#pragma unroll
                for (int m = 0; m < (KT > 0 ? KT : 1); ++m)
                    if (m < d.K) Lo[m] = Lr[m];
            }
            if constexpr (PACKED) detect_packed<KT, NSP>(Lr, mu2, W2, nE2, out);
            else detect_reg<R, (KT > 0 ? KT : 1)>(Lr, mu_s, W_s, d.nE, d.nSpec, out);
        } else {
            for (int m = 0; m < d.K; ++m) acc_s[m * blockDim.x + tid] *= seglen;
            if (d.L_out) {
                R *Lo = (R *)d.L_out + ray * d.K;
                for (int m = 0; m < d.K; ++m) Lo[m] = acc_s[m * blockDim.x + tid];
            }
            detect<R>(acc_s + tid, blockDim.x, d.K, mu_s, W_s, d.nE, d.nSpec, out);
        }
#pragma unroll
        for (int sp = 0; sp < MAX_SPEC; ++sp) {
            if (sp < d.nSpec) {
                ((R *)d.counts)[sp * n_rays + ray] = out[sp];
                if (d.logs) ((R *)d.logs)[sp * n_rays + ray] = neg_log_ratio<R>(out[sp], air_s[sp]);
            }
        }
    }
}

// Detection from stored lengths L [n_rays][K].
template <typename R>
__global__ void __launch_bounds__(128) poly_detect_kernel(const R *__restrict__ Lin, long long n_rays, DetectArgs d)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    R *acc_s = (R *)smem_raw;                              // [K][blockDim]
    R *mu_s = acc_s + (size_t)d.K * blockDim.x;
    R *W_s = mu_s + (size_t)d.K * d.nE;
    R *air_s = W_s + (size_t)d.nSpec * d.nE;
    stage_tables<R>(d, d.K, mu_s, W_s, air_s);
    const int tid = threadIdx.x;
    const long long ray0 = (long long)blockIdx.x * blockDim.x;
    // coalesced load of this block's [blockDim][K] slab, transposed into acc_s[m][tid]
    const long long base = ray0 * d.K;
    const long long lim = n_rays * d.K;
    for (int t = tid; t < (int)blockDim.x * d.K; t += blockDim.x) {
        const long long gidx = base + t;
        const int rr = t / d.K, m = t % d.K;
        acc_s[m * blockDim.x + rr] = gidx < lim ? Lin[gidx] : (R)0;
    }
    __syncthreads();
    const long long ray = ray0 + tid;
    if (ray >= n_rays) return;
    R out[MAX_SPEC];
    detect<R>(acc_s + tid, blockDim.x, d.K, mu_s, W_s, d.nE, d.nSpec, out);
#pragma unroll
    for (int sp = 0; sp < MAX_SPEC; ++sp) {
        if (sp < d.nSpec) {
            ((R *)d.counts)[sp * n_rays + ray] = out[sp];
            if (d.logs) ((R *)d.logs)[sp * n_rays + ray] = neg_log_ratio<R>(out[sp], air_s[sp]);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Debug trace: same set-up, same column routine, pieces written out in ray order.

template <typename R> struct TraceSink {
    int64_t vox[3];
    int lab[3];
    R w[3];
    int n;
    bool xmajor;
    int i, nx, ny, nq, nz, j_c, k_c;
    __device__ __forceinline__ void operator()(int slot, int jrel, int krel, int label, R wt)
    {
        const int j = jrel + j_c, k = krel + k_c;
        const bool inside = j >= 0 && j < nq && k >= 0 && k < nz;
        vox[slot] = inside ? (xmajor ? ((int64_t)k * ny + j) * nx + i : ((int64_t)k * ny + i) * nx + j) : -1;
        lab[slot] = label;
        w[slot] = wt;
        n = slot + 1;
    }
};

template <typename R, bool IS3D>
__global__ void siddon_trace_kernel(SiddonParams P, const int64_t *__restrict__ ray_ids, int n_rays, int max_seg,
                                    int64_t *__restrict__ vox, uint8_t *__restrict__ lab, R *__restrict__ len,
                                    int32_t *__restrict__ n_seg)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_rays) return;
    const int64_t ray = ray_ids[t];
    const int c = (int)(ray % P.scan.n_ch);
    const int64_t line = ray / P.scan.n_ch;
    const int r = (int)(line % P.scan.n_rows);
    const int v = (int)(line / P.scan.n_rows);
    RaySetup s;
    ray_setup<IS3D>(P, v, r, c, s);
    int n = 0;
    if (s.hit) {
        RayLoop<R> L;
        L.init(P, s);
        const R seglen = (R)s.seglen;
        TraceSink<R> sink;
        sink.xmajor = s.xmajor; sink.nx = P.vol.nx; sink.ny = P.vol.ny; sink.nq = s.nq; sink.nz = P.vol.nz;
        sink.j_c = s.j_c; sink.k_c = s.k_c;
        const int step = s.fwd ? 1 : -1;
        for (int i = s.fwd ? s.i0 : s.i1; s.fwd ? (i <= s.i1) : (i >= s.i0); i += step) {
            sink.n = 0; sink.i = i;
            slab_column<R, IS3D>(L, i, sink);
            for (int q = 0; q < sink.n; ++q) {
                const int slot = s.fwd ? q : sink.n - 1 - q;
                if (!(sink.w[slot] > (R)0) || sink.vox[slot] < 0) continue;
                if (vox && n < max_seg) {
                    const size_t o = (size_t)t * max_seg + n;
                    vox[o] = sink.vox[slot];
                    lab[o] = (uint8_t)sink.lab[slot];
                    len[o] = Rn<R>::mul(sink.w[slot], seglen);
                }
                ++n;
            }
        }
    }
    n_seg[t] = n;
}

// Debug: the steps of the PRODUCTION walks (walk_skip / walk_skip3 over the skip layouts) of selected rays, recorded by a
// sink inside the very same routine the production kernels inline.  Per step: first column, run length (0 = an ordinary
// column split into pieces, S > 0 = a homogeneous stretch of S columns taken in one go), and up to three pieces (voxel,
// label, length); a piece outside the volume (the zero apron) has voxel -1.
template <typename R> struct WalkTraceSink {
    static constexpr bool active = true;
    int32_t *col, *run;
    int64_t *vox;
    uint8_t *lab;
    R *len;
    int n, max_steps;
    bool xmajor, mpos;
    int nx, ny, nq, nz, j_c, k_c;
    R seglen;
    __device__ __forceinline__ void operator()(int i, int run_len, int np, const int *js, const int *ks, const int *ls, const R *ws)
    {
        if (col && n < max_steps) {
            col[n] = i;
            run[n] = run_len;
            for (int q = 0; q < 3; ++q) {
                const int j = mpos ? j_c + js[q] : j_c - js[q], k = k_c + ks[q];
                const bool inside = q < np && j >= 0 && j < nq && k >= 0 && k < nz;
                vox[3 * n + q] = inside ? (xmajor ? ((int64_t)k * ny + j) * nx + i : ((int64_t)k * ny + i) * nx + j) : -1;
                lab[3 * n + q] = q < np ? (uint8_t)ls[q] : 0;
                len[3 * n + q] = q < np ? Rn<R>::mul(ws[q], seglen) : (R)0;
            }
        }
        ++n;
    }
};

template <typename R, bool IS3D>
__global__ void siddon_walk_trace_kernel(SiddonParams P, const int64_t *__restrict__ ray_ids, int n_rays, int max_steps, int K,
                                         int32_t *__restrict__ col, int32_t *__restrict__ run, int64_t *__restrict__ vox,
                                         uint8_t *__restrict__ lab, R *__restrict__ len, int32_t *__restrict__ n_steps,
                                         int32_t *__restrict__ meta, R *__restrict__ acc_scratch)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_rays) return;
    const int64_t ray = ray_ids[t];
    const int c = (int)(ray % P.scan.n_ch);
    const int64_t line = ray / P.scan.n_ch;
    const int r = (int)(line % P.scan.n_rows);
    const int v = (int)(line / P.scan.n_rows);
    RaySetup s;
    ray_setup<IS3D>(P, v, r, c, s);
    R *acc = acc_scratch + (size_t)t * K;                      // per-material sums of the walk (stride 1), returned scaled
    for (int m = 0; m < K; ++m) acc[m] = (R)0;
    WalkTraceSink<R> sink;
    sink.n = 0;
    if (s.hit) {
        const bool mpos = s.m >= 0.0;
        const int pitch = s.xmajor ? P.pitch_xm : P.pitch_ym;
        const size_t o = (size_t)t * max_steps;
        sink.col = col ? col + o : nullptr; sink.run = run + o; sink.vox = vox + 3 * o; sink.lab = lab + 3 * o; sink.len = len + 3 * o;
        sink.max_steps = max_steps; sink.xmajor = s.xmajor; sink.mpos = mpos;
        sink.nx = P.vol.nx; sink.ny = P.vol.ny; sink.nq = s.nq; sink.nz = P.vol.nz; sink.j_c = s.j_c; sink.k_c = s.k_c;
        sink.seglen = (R)s.seglen;
        const double c0 = mpos ? s.cq : 1.0 - s.cq;
        if (IS3D) {
            const int plane = (int)(s.xmajor ? P.plane_xm : P.plane_ym);
            const uint16_t *W = P.skip[(s.xmajor ? 0 : 2) + (mpos ? 0 : 1)] + (long long)(s.k_c + 1) * plane +
                                (mpos ? s.j_c : s.nq - 1 - s.j_c) + 1;
            walk_skip3<R, 0>(W, pitch, plane, s.i0, s.i1, s.i_c, (R)fabs(s.m), (R)c0, (R)s.mz, (R)s.cw, (R)s.ua, (R)s.ub,
                          (R)s.inv_am, (R)s.inv_amz, acc, 1, sink);
        } else {
            const uint16_t *W = P.skip[(s.xmajor ? 0 : 2) + (mpos ? 0 : 1)] + (mpos ? s.j_c : s.nq - 1 - s.j_c) + 1;
            walk_skip<R, 0>(W, pitch, s.i0, s.i1, s.i_c, (R)fabs(s.m), (R)c0, (R)s.ua, (R)s.ub, (R)s.inv_am, acc, 1, sink);
        }
        for (int m = 0; m < K; ++m) acc[m] = Rn<R>::mul(acc[m], (R)s.seglen);
    }
    n_steps[t] = sink.n;
    meta[t] = (s.hit ? 8 : 0) | (s.hit && s.m >= 0.0 ? 4 : 0) | (s.fwd ? 2 : 0) | (s.xmajor ? 1 : 0);
}

// ---------------------------------------------------------------------------------------------------------
// Host side

static int make_params(const DexctScan *scan, const DexctVolume *vol, const void *workspace, SiddonParams &P)
{
    DEXCT_REQUIRE(scan && vol && workspace, "scan, vol and workspace must be non-null");
    DEXCT_REQUIRE(scan->n_views > 0 && scan->n_rows > 0 && scan->n_ch > 0, "scan sizes must be positive");
    DEXCT_REQUIRE(vol->nx > 0 && vol->ny > 0 && vol->nz > 0, "volume sizes must be positive");
    DEXCT_REQUIRE(vol->dx > 0 && vol->dy > 0 && vol->dz > 0, "voxel sizes must be positive");
    DEXCT_REQUIRE(scan->sid > 0 && scan->sdd > 0, "SID and SDD must be positive");
    const WsLayout L = ws_layout(*scan, *vol);
    const char *w = (const char *)workspace;
    P.scan = *scan; P.vol = *vol;
    P.cosb = (const double *)(w + L.off_cosb); P.sinb = (const double *)(w + L.off_sinb);
    P.cosg = (const double *)(w + L.off_cosg); P.sing = (const double *)(w + L.off_sing);
    P.zdet = (const double *)(w + L.off_zdet);
    P.xm = (const uint16_t *)(w + L.off_xm); P.ym = (const uint16_t *)(w + L.off_ym);
    P.pitch_xm = L.pitch_xm; P.pitch_ym = L.pitch_ym; P.plane_xm = L.plane_xm; P.plane_ym = L.plane_ym;
    P.inv_dx = 1.0 / vol->dx; P.inv_dy = 1.0 / vol->dy; P.inv_dz = 1.0 / vol->dz;
    const bool skip = L.has_skip && vol->nx + 1 <= SKIP_MAX_ENTRIES && vol->ny + 1 <= SKIP_MAX_ENTRIES;
    for (int a = 0; a < 4; ++a) P.skip[a] = skip ? (const uint16_t *)(w + L.off_skip[a]) : nullptr;
    // steepest z slope of any ray of the scan, in planes per column: |gz| / |gp| with |gp| >= SDD / sqrt(dx^2 + dy^2)
    const double zmax = fmax(fabs(scan->z0), fabs(scan->z0 + (scan->n_rows - 1) * scan->dzdet));
    const double muz = zmax / vol->dz * sqrt(vol->dx * vol->dx + vol->dy * vol->dy) / scan->sdd;
    // the slab traversal (siddon_core.cuh) crosses at most ONE z plane per column: steeper scans -- thin slices under a tall
    // detector or a short source-detector distance -- would silently drop voxels
    DEXCT_REQUIRE(scan->n_rows == 1 || muz <= 1.0,
                  "cone too steep for the slab traversal: a ray climbs up to %.3f slices per column (detector half-height %.4g cm "
                  "at SDD %.4g cm, dz %.4g cm, dx %.4g cm, dy %.4g cm); at most 1 is supported -- use thicker slices or fewer rows",
                  muz, zmax, scan->sdd, vol->dz, vol->dx, vol->dy);
    for (int l = 0; l < SKIP_LEVELS; ++l) {
        double cap = muz > 0 ? floor((l + 1) / muz) : 127.0;
        P.skip_cap[l] = (int)fmin(fmax(cap, 1.0), 127.0);
    }
    return DEXCT_OK;
}

// Threads per CTA: 128, halved (down to one warp) while the per-thread material accumulators (acc_bytes each) plus the tables
// do not fit into 200 KiB of shared memory -- float64 scans with more than ~180 materials.
static int cta_threads(size_t acc_bytes, size_t tab_bytes)
{
    int block = 128;
    while (block > 32 && acc_bytes * block + tab_bytes > 200 * 1024) block >>= 1;
    return block;
}

template <typename R, bool IS3D, int KT, int NSP>
static int launch_skip(const SiddonParams &P, int view_begin, long long n_rays, const DetectArgs &d, cudaStream_t st)
{
    const int KP = KT > 0 ? KT : d.K;
    const bool packed = sizeof(R) == 4 && KT > 0;
    const int nE2 = (d.nE + 1) / 2;
    const size_t tab = packed ? 8 * (size_t)nE2 * (KT + NSP) + sizeof(R) * MAX_SPEC
                              : sizeof(R) * ((size_t)KP * d.nE + (size_t)d.nSpec * d.nE + MAX_SPEC);
    const int block = cta_threads(sizeof(R) * (size_t)KP, tab);
    const size_t smem = sizeof(R) * (size_t)KP * block + tab;
    if (smem > 200 * 1024) {
        set_error("material/energy tables need %zu bytes of shared memory (limit 200 KiB): K=%d nE=%d", smem, d.K, d.nE);
        return DEXCT_ECAPACITY;
    }
    void (*kern)(SiddonParams, int, long long, DetectArgs) = siddon_skip_kernel<R, KT, NSP>;
    if (IS3D) {
        // skip layouts far beyond L2 (126 MB): the walk waits on DRAM, more resident warps help
        const double layout_bytes = 4.0 * 2.0 * (double)P.vol.nx * P.vol.ny * P.vol.nz;
        const char *mb = getenv("DEXCT_SID3_MINB");           // tuning knob: 9, 10 or 12
        const int minb = mb ? atoi(mb) : (layout_bytes > 4.0e9 ? 10 : 9);
        kern = siddon_skip3_kernel<R, KT, NSP, 9>;
        if (sizeof(R) == 4 && minb == 10) kern = siddon_skip3_kernel<R, KT, NSP, 10>;
        if (sizeof(R) == 4 && minb == 12) kern = siddon_skip3_kernel<R, KT, NSP, 12>;
    }
    if (smem > 48 * 1024) DEXCT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    DEXCT_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, block, smem));
    if (per_sm < 1) per_sm = 1;
    const long long n_blk = (n_rays + block - 1) / block;
    const long long grid = std::min<long long>(n_blk, (long long)sm_count() * per_sm);     // resident CTAs stride over ray blocks
    SiddonParams Pl = P;
    Pl.row_band = 1;
    if (IS3D) {                                               // the largest divisor of the row count up to 8 (siddon_skip3_kernel)
        const char *rb = getenv("DEXCT_ROW_BAND");            // tuning knob
        const int want = rb && atoi(rb) >= 1 ? atoi(rb) : 8;
        for (int b = 1; b <= want && b <= P.scan.n_rows; ++b)
            if (P.scan.n_rows % b == 0) Pl.row_band = b;
    }
    kern<<<(unsigned)grid, block, smem, st>>>(Pl, view_begin, n_rays, d);
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}

template <typename R, bool IS3D, int KT>
static int launch_poly(const SiddonParams &P, int view_begin, long long n_rays, const DetectArgs &d, cudaStream_t st)
{
    // scans with 7-bit labels walk the skip layouts
    if (P.skip[0] != nullptr && d.K <= 128)
        return d.nSpec <= 2 ? launch_skip<R, IS3D, KT, 2>(P, view_begin, n_rays, d, st)
                            : launch_skip<R, IS3D, KT, 4>(P, view_begin, n_rays, d, st);
    const int KP = KT > 0 ? KT : d.K;
    const size_t tab = sizeof(R) * ((size_t)KP * d.nE + (size_t)d.nSpec * d.nE + MAX_SPEC);
    const int block = cta_threads(sizeof(R) * (size_t)KP, tab);
    const size_t smem = sizeof(R) * (size_t)KP * block + tab;
    if (smem > 200 * 1024) {
        set_error("material/energy tables need %zu bytes of shared memory (limit 200 KiB): K=%d nE=%d", smem, d.K, d.nE);
        return DEXCT_ECAPACITY;
    }
    auto kern = siddon_poly_kernel<R, IS3D, KT>;
    if (smem > 48 * 1024) DEXCT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long grid = (n_rays + block - 1) / block;
    kern<<<(unsigned)grid, block, smem, st>>>(P, view_begin, n_rays, d);
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}

template <typename R, bool IS3D>
static int dispatch_k(const SiddonParams &P, int view_begin, long long n_rays, const DetectArgs &d, cudaStream_t st)
{
    if (d.K <= 4) return launch_poly<R, IS3D, 4>(P, view_begin, n_rays, d, st);
    if (d.K <= 8) return launch_poly<R, IS3D, 8>(P, view_begin, n_rays, d, st);
    if (d.K <= 16) return launch_poly<R, IS3D, 16>(P, view_begin, n_rays, d, st);
    return launch_poly<R, IS3D, 0>(P, view_begin, n_rays, d, st);
}

}  // namespace dexct

using namespace dexct;

extern "C" size_t dexct_siddon_workspace_bytes(const DexctScan *scan, const DexctVolume *vol)
{
    if (!scan || !vol || scan->n_views <= 0 || scan->n_ch <= 0 || scan->n_rows <= 0 || vol->nx <= 0 || vol->ny <= 0 ||
        vol->nz <= 0)
        return 0;
    return ws_layout(*scan, *vol).total;
}

extern "C" int dexct_siddon_prepare(const DexctScan *scan, const DexctVolume *vol, const uint8_t *labels,
                                    void *workspace, void *stream)
{
    if (int e = check_device()) return e;
    SiddonParams P;
    if (int e = make_params(scan, vol, workspace, P)) return e;
    DEXCT_REQUIRE(labels, "labels must be non-null");
    cudaStream_t st = (cudaStream_t)stream;
    const WsLayout L = ws_layout(*scan, *vol);
    char *w = (char *)workspace;
    const int nmax = max(scan->n_views, max(scan->n_ch, scan->n_rows));
    tables_kernel<<<(nmax + 127) / 128, 128, 0, st>>>(*scan, (double *)(w + L.off_cosb), (double *)(w + L.off_sinb),
                                                      (double *)(w + L.off_cosg), (double *)(w + L.off_sing),
                                                      (double *)(w + L.off_zdet));
    DEXCT_CUDA_OK(cudaGetLastError());
    // apron planes (k = -1, nz) and pitch padding must read as label 0
    DEXCT_CUDA_OK(cudaMemsetAsync(w + L.off_xm, 0, L.total - L.off_xm, st));
    repack_ym_kernel<<<(unsigned)((long long)vol->nz * vol->ny), 128, 0, st>>>(labels, vol->nx, vol->ny, vol->nz,
                                                                               (uint16_t *)(w + L.off_ym), L.pitch_ym, L.plane_ym);
    DEXCT_CUDA_OK(cudaGetLastError());
    dim3 grid((vol->nx + 31) / 32, (vol->ny + 2 + 31) / 32, vol->nz);
    repack_xm_kernel<<<grid, dim3(32, 8), 0, st>>>(labels, vol->nx, vol->ny, vol->nz, (uint16_t *)(w + L.off_xm),
                                                  L.pitch_xm, L.plane_xm);
    DEXCT_CUDA_OK(cudaGetLastError());
    if (P.skip[0] && !L.skip_3d) {
        SkipBuildArgs A;
        A.pairs[0] = (const uint16_t *)(w + L.off_xm) + L.plane_xm;
        A.pairs[1] = (const uint16_t *)(w + L.off_ym) + L.plane_ym;
        for (int a = 0; a < 4; ++a) A.out[a] = (uint16_t *)(w + L.off_skip[a]);
        A.np[0] = vol->nx; A.nq[0] = vol->ny; A.pitch[0] = L.pitch_xm;
        A.np[1] = vol->ny; A.nq[1] = vol->nx; A.pitch[1] = L.pitch_ym;
        const int n_max = max(vol->nx, vol->ny);
        skip_runs_kernel<<<dim3((n_max + 1 + 127) / 128, 4), 128, 0, st>>>(A);
        DEXCT_CUDA_OK(cudaGetLastError());
        skip_cone_kernel<<<dim3((2 * n_max + 127) / 128, 4), 128, 0, st>>>(A);
        DEXCT_CUDA_OK(cudaGetLastError());
    } else if (P.skip[0]) {
        SkipBuild3Args A;
        A.pairs[0] = (const uint16_t *)(w + L.off_xm);
        A.pairs[1] = (const uint16_t *)(w + L.off_ym);
        A.rz[0] = (const uint8_t *)(w + L.off_rz[0]);
        A.rz[1] = (const uint8_t *)(w + L.off_rz[1]);
        for (int a = 0; a < 4; ++a) A.out[a] = (uint16_t *)(w + L.off_skip[a]);
        A.np[0] = vol->nx; A.nq[0] = vol->ny; A.pitch[0] = L.pitch_xm; A.plane[0] = L.plane_xm;
        A.np[1] = vol->ny; A.nq[1] = vol->nx; A.pitch[1] = L.pitch_ym; A.plane[1] = L.plane_ym;
        for (int l = 0; l < SKIP_LEVELS; ++l) A.cap[l] = P.skip_cap[l];
        for (int ax = 0; ax < 2; ++ax) {
            const dim3 grid((A.nq[ax] + 127) / 128, A.np[ax]);
            rz_kernel<<<grid, 128, 0, st>>>(A.pairs[ax], A.np[ax], A.nq[ax], vol->nz, A.pitch[ax], A.plane[ax],
                                            (uint8_t *)(w + L.off_rz[ax]));
            DEXCT_CUDA_OK(cudaGetLastError());
        }
        const size_t smem = sizeof(uint16_t) * 2 * SKIP_LEVELS * (size_t)(max(vol->nx, vol->ny) + 3);
        if (smem > 48 * 1024)
            DEXCT_CUDA_OK(cudaFuncSetAttribute(skip_build3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        skip_build3_kernel<<<dim3(4, vol->nz), SKIP_THREADS, smem, st>>>(A);
        DEXCT_CUDA_OK(cudaGetLastError());
    }
    return DEXCT_OK;
}

extern "C" int dexct_siddon_tables_host(const DexctScan *scan, const void *workspace, double *cb, double *sb, double *cg,
                                        double *sg, double *zd, void *stream)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(scan && workspace, "scan and workspace must be non-null");
    DexctVolume dummy = {1, 1, 1, 0, 1.0, 1.0, 1.0};
    const WsLayout L = ws_layout(*scan, dummy);          // table offsets do not depend on the volume
    const char *w = (const char *)workspace;
    cudaStream_t st = (cudaStream_t)stream;
    DEXCT_CUDA_OK(cudaStreamSynchronize(st));
    if (cb) DEXCT_CUDA_OK(cudaMemcpy(cb, w + L.off_cosb, sizeof(double) * scan->n_views, cudaMemcpyDeviceToHost));
    if (sb) DEXCT_CUDA_OK(cudaMemcpy(sb, w + L.off_sinb, sizeof(double) * scan->n_views, cudaMemcpyDeviceToHost));
    if (cg) DEXCT_CUDA_OK(cudaMemcpy(cg, w + L.off_cosg, sizeof(double) * scan->n_ch, cudaMemcpyDeviceToHost));
    if (sg) DEXCT_CUDA_OK(cudaMemcpy(sg, w + L.off_sing, sizeof(double) * scan->n_ch, cudaMemcpyDeviceToHost));
    if (zd) DEXCT_CUDA_OK(cudaMemcpy(zd, w + L.off_zdet, sizeof(double) * scan->n_rows, cudaMemcpyDeviceToHost));
    return DEXCT_OK;
}

extern "C" int dexct_siddon_poly(const DexctScan *scan, const DexctVolume *vol, const void *workspace,
                                 int32_t view_begin, int32_t view_count, const void *mu, int32_t K, int32_t nE,
                                 const void *W, int32_t nSpec, void *counts, void *logs, void *L_out, int32_t dtype,
                                 void *stream)
{
    if (int e = check_device()) return e;
    SiddonParams P;
    if (int e = make_params(scan, vol, workspace, P)) return e;
    DEXCT_REQUIRE(view_begin >= 0 && view_count >= 0 && view_begin + view_count <= scan->n_views,
                  "view range [%d, %d) outside the scan's %d views", view_begin, view_begin + view_count, scan->n_views);
    DEXCT_REQUIRE(K >= 1 && K <= 256 && nE >= 1, "need 1 <= K <= 256 and nE >= 1 (K=%d nE=%d)", K, nE);
    DEXCT_REQUIRE(nSpec >= 1 && nSpec <= MAX_SPEC, "nSpec must be in 1..%d", MAX_SPEC);
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    if (view_count == 0) return DEXCT_OK;
    DEXCT_REQUIRE(mu && W && counts, "mu, W and counts must be non-null");
    const long long n_rays = (long long)view_count * scan->n_rows * scan->n_ch;
    DetectArgs d{mu, W, K, nE, nSpec, counts, logs, L_out};
    cudaStream_t st = (cudaStream_t)stream;
    const bool is3d = scan_is_3d(*scan, *vol);
    if (dtype == DEXCT_F32)
        return is3d ? dispatch_k<float, true>(P, view_begin, n_rays, d, st) : dispatch_k<float, false>(P, view_begin, n_rays, d, st);
    return is3d ? dispatch_k<double, true>(P, view_begin, n_rays, d, st) : dispatch_k<double, false>(P, view_begin, n_rays, d, st);
}

extern "C" int dexct_poly_detect(const void *L, int64_t n_rays, const void *mu, int32_t K, int32_t nE, const void *W,
                                 int32_t nSpec, void *counts, void *logs, int32_t dtype, void *stream)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(K >= 1 && K <= 256 && nE >= 1, "need 1 <= K <= 256 and nE >= 1 (K=%d nE=%d)", K, nE);
    DEXCT_REQUIRE(nSpec >= 1 && nSpec <= MAX_SPEC, "nSpec must be in 1..%d", MAX_SPEC);
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    if (n_rays <= 0) return DEXCT_OK;
    DEXCT_REQUIRE(L && mu && W && counts, "L, mu, W and counts must be non-null");
    DetectArgs d{mu, W, K, nE, nSpec, counts, logs, nullptr};
    cudaStream_t st = (cudaStream_t)stream;
    const size_t esz = dtype == DEXCT_F32 ? 4 : 8;
    const size_t tab = esz * ((size_t)K * nE + (size_t)nSpec * nE + MAX_SPEC);
    const int block = cta_threads(esz * (size_t)K, tab);
    const size_t smem = esz * (size_t)K * block + tab;
    if (smem > 200 * 1024) {
        set_error("material/energy tables need %zu bytes of shared memory (limit 200 KiB): K=%d nE=%d", smem, K, nE);
        return DEXCT_ECAPACITY;
    }
    const long long grid = (n_rays + block - 1) / block;
    if (dtype == DEXCT_F32) {
        if (smem > 48 * 1024) DEXCT_CUDA_OK(cudaFuncSetAttribute(poly_detect_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        poly_detect_kernel<float><<<(unsigned)grid, block, smem, st>>>((const float *)L, n_rays, d);
    } else {
        if (smem > 48 * 1024) DEXCT_CUDA_OK(cudaFuncSetAttribute(poly_detect_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        poly_detect_kernel<double><<<(unsigned)grid, block, smem, st>>>((const double *)L, n_rays, d);
    }
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}

extern "C" int dexct_siddon_trace(const DexctScan *scan, const DexctVolume *vol, const void *workspace,
                                  const int64_t *ray_ids, int32_t n_rays, int32_t max_seg, int64_t *vox, uint8_t *lab,
                                  void *len, int32_t *n_seg, int32_t dtype, void *stream)
{
    if (int e = check_device()) return e;
    SiddonParams P;
    if (int e = make_params(scan, vol, workspace, P)) return e;
    DEXCT_REQUIRE(ray_ids && n_seg, "ray_ids and n_seg must be non-null");
    DEXCT_REQUIRE((vox && lab && len) || (!vox && !lab && !len), "vox, lab and len must be all given or all NULL");
    DEXCT_REQUIRE(n_rays >= 0 && (max_seg > 0 || !vox), "n_rays >= 0 and max_seg > 0 required");
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    if (n_rays == 0) return DEXCT_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const bool is3d = scan_is_3d(*scan, *vol);
    const int block = 64, grid = (n_rays + block - 1) / block;
    if (dtype == DEXCT_F32) {
        if (is3d) siddon_trace_kernel<float, true><<<grid, block, 0, st>>>(P, ray_ids, n_rays, max_seg, vox, lab, (float *)len, n_seg);
        else siddon_trace_kernel<float, false><<<grid, block, 0, st>>>(P, ray_ids, n_rays, max_seg, vox, lab, (float *)len, n_seg);
    } else {
        if (is3d) siddon_trace_kernel<double, true><<<grid, block, 0, st>>>(P, ray_ids, n_rays, max_seg, vox, lab, (double *)len, n_seg);
        else siddon_trace_kernel<double, false><<<grid, block, 0, st>>>(P, ray_ids, n_rays, max_seg, vox, lab, (double *)len, n_seg);
    }
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}

extern "C" int dexct_siddon_walk_trace(const DexctScan *scan, const DexctVolume *vol, const void *workspace,
                                       const int64_t *ray_ids, int32_t n_rays, int32_t max_steps, int32_t K, int32_t *col,
                                       int32_t *run, int64_t *vox, uint8_t *lab, void *len, int32_t *n_steps, int32_t *meta,
                                       void *L_out, int32_t dtype, void *stream)
{
    if (int e = check_device()) return e;
    SiddonParams P;
    if (int e = make_params(scan, vol, workspace, P)) return e;
    DEXCT_REQUIRE(ray_ids && n_steps && meta && L_out, "ray_ids, n_steps, meta and L_out must be non-null");
    DEXCT_REQUIRE((col && run && vox && lab && len) || (!col && !run && !vox && !lab && !len),
                  "col, run, vox, lab and len must be all given or all NULL");
    DEXCT_REQUIRE(n_rays >= 0 && (max_steps > 0 || !col), "n_rays >= 0 and max_steps > 0 required");
    DEXCT_REQUIRE(K >= 1 && K <= 128, "the skip layouts carry labels below 128: need 1 <= K <= 128 (K=%d)", K);
    DEXCT_REQUIRE(P.skip[0] != nullptr, "this volume has no skip layouts (a side exceeds %d voxels)", SKIP_MAX_ENTRIES - 1);
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    if (n_rays == 0) return DEXCT_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const bool is3d = scan_is_3d(*scan, *vol);
    const int block = 64, grid = (n_rays + block - 1) / block;
    if (dtype == DEXCT_F32) {
        if (is3d) siddon_walk_trace_kernel<float, true><<<grid, block, 0, st>>>(P, ray_ids, n_rays, max_steps, K, col, run, vox, lab, (float *)len, n_steps, meta, (float *)L_out);
        else siddon_walk_trace_kernel<float, false><<<grid, block, 0, st>>>(P, ray_ids, n_rays, max_steps, K, col, run, vox, lab, (float *)len, n_steps, meta, (float *)L_out);
    } else {
        if (is3d) siddon_walk_trace_kernel<double, true><<<grid, block, 0, st>>>(P, ray_ids, n_rays, max_steps, K, col, run, vox, lab, (double *)len, n_steps, meta, (double *)L_out);
        else siddon_walk_trace_kernel<double, false><<<grid, block, 0, st>>>(P, ray_ids, n_rays, max_steps, K, col, run, vox, lab, (double *)len, n_steps, meta, (double *)L_out);
    }
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}
