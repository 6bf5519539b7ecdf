/*
 * ORACLE -- test infrastructure, never the product path.
 *
 * CPU float64 restatement of the forward-projection half of the hot path of gjadick/dex-ct-sim:
 *   get_sino(ct, phantom, spec)  (call site /root/reference/main.py:120; contract matdecomp.py:132-136,178-179)
 *     = Siddon ray traversal of a uint8 material-label volume  ->  per-material path lengths L_m [cm]
 *     + polyenergetic Beer-Lambert detection  counts = sum_E W(E) exp(-sum_m L_m mu_m(E))
 *       (the same forward model the reference's decomposition inverts, matdecomp.py:116-118,146-151).
 *
 * PARITY UNPINNED.  The reference's tracer lives in the git submodule gjadick/x-tomo-sim pinned at commit
 * f21a7e7ed20e491c474829fa5aebee6047b5427a (/root/reference/.gitmodules:1-3, .SUBMODULES.json); that commit is
 * not in the mounted tree (xtomosim/ is empty) and there is no network, and the reference holds no tests or
 * golden vectors for this path.  What is restated here is therefore the PUBLISHED algorithm the reference cites
 * (README.md:28,41): R. L. Siddon, "Fast calculation of the exact radiological path for a three-dimensional CT
 * array", Med. Phys. 12(2) 252-255 (1985), anchored on the reference's call sites and on analytic
 * known-answer tests (tests/test_oracle_siddon.py).  Geometry conventions the reference does not reveal are
 * project-defined and documented in DESIGN.md section 3.
 *
 * Three independent pieces:
 *   (A) dexo_classic_*  : Siddon 1985 -- parametric plane crossings alpha_x(i), alpha_y(j), alpha_z(k), merged in
 *                         increasing alpha; segment length = |T-S| (alpha_{n+1}-alpha_n); voxel from the
 *                         segment midpoint (paper eqs. 4-13).  float64.  This is the TRUTH for line integrals.
 *   (B) dexo_slab_*     : operation-for-operation CPU twin of the device traversal ("slab" form: one major-axis
 *                         column at a time, at most one minor-axis crossing per column).  It exists so that the
 *                         voxel-index / material-id SEQUENCE of the CUDA kernel can be compared bit-exactly in
 *                         both float32 and float64 device modes; it is itself validated against (A).
 *   (C) dexo_poly_detect: float64 Beer-Lambert sum over energy bins.
 *
 * Build: oracle/Makefile (gcc -O2 -ffp-contract=off -fopenmp).  -ffp-contract=off matters: every fused
 * multiply-add in (B) is an explicit fma()/fmaf(), everything else rounds once per operation, exactly like the
 * device code which uses __fma_rn/__fmul_rn/__fadd_rn intrinsics.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
    /* scan */
    int32_t n_views, n_rows, n_ch;
    double sid, sdd;
    const double *cos_beta, *sin_beta;   /* [n_views]  source angle table (host libm, shared with the device) */
    const double *cos_gamma, *sin_gamma; /* [n_ch]     fan angle table */
    const double *zdet;                  /* [n_rows]   z of the detector-row centre at the detector [cm] */
    /* volume, centred on the origin, labels[(k*ny + j)*nx + i] */
    int32_t nx, ny, nz;
    double dx, dy, dz;
} dexo_setup;

/* ------------------------------------------------------------------------------------------------------ */
/* Ray end points.  Equiangular fan: source S = SID (cos b, sin b, 0); the ray of channel c leaves S along
 * -(cos(b+g_c), sin(b+g_c)) and meets the detector arc (radius SDD about S) at T; row r lifts T by zdet[r].
 * cos/sin of the sum come from the angle-addition formulas on the two tables, so that host and device get
 * bit-identical end points without depending on either libm.                                               */
static void ray_endpoints(const dexo_setup *g, int v, int r, int c, double S[3], double D[3])
{
    const double cb = g->cos_beta[v], sb = g->sin_beta[v];
    const double cg = g->cos_gamma[c], sg = g->sin_gamma[c];
    const double ex = cb * cg - sb * sg;
    const double ey = sb * cg + cb * sg;
    S[0] = g->sid * cb; S[1] = g->sid * sb; S[2] = 0.0;
    D[0] = -(g->sdd * ex); D[1] = -(g->sdd * ey); D[2] = g->zdet[r];
}

void dexo_ray_endpoints(const dexo_setup *g, int v, int r, int c, double *S, double *T)
{
    double D[3];
    ray_endpoints(g, v, r, c, S, D);
    for (int a = 0; a < 3; ++a) T[a] = S[a] + D[a];
}

/* ------------------------------------------------------------------------------------------------------ */
/* (A) Siddon 1985.                                                                                        */

typedef void (*seg_sink)(void *ctx, int64_t voxel, int label, double len);

/* alpha of plane number i on one axis (paper eq. 4): planes at min + i*d, i = 0..n */
static inline double plane_alpha(double vmin, double d, int i, double s, double D)
{
    return ((vmin + (double)i * d) - s) / D;
}

/* Walk the ray S + alpha D, alpha in [0,1], through the volume; call sink for every segment of positive
 * length, in order of increasing alpha.  Returns the number of segments emitted.                           */
static int classic_walk(const double S[3], const double D[3], const uint8_t *labels,
                        const int n[3], const double d[3], seg_sink sink, void *ctx)
{
    double vmin[3];
    double amin = 0.0, amax = 1.0;
    for (int a = 0; a < 3; ++a) {
        vmin[a] = -0.5 * (double)n[a] * d[a];
        if (D[a] != 0.0) {
            double a0 = plane_alpha(vmin[a], d[a], 0, S[a], D[a]);
            double a1 = plane_alpha(vmin[a], d[a], n[a], S[a], D[a]);
            double lo = a0 < a1 ? a0 : a1, hi = a0 < a1 ? a1 : a0;      /* paper eq. 5 */
            if (lo > amin) amin = lo;
            if (hi < amax) amax = hi;
        } else {
            /* ray parallel to this axis' planes: it is inside the slab [vmin, vmin+n d) or misses */
            double rel = (S[a] - vmin[a]) / d[a];
            if (!(rel >= 0.0 && rel < (double)n[a])) return 0;
        }
    }
    if (!(amin < amax)) return 0;

    const double len = sqrt(D[0] * D[0] + D[1] * D[1] + D[2] * D[2]);
    /* per axis: index of the next plane to be crossed and the step (paper eq. 6, written incrementally) */
    int next[3], step[3], last[3];
    double anext[3];
    for (int a = 0; a < 3; ++a) {
        if (D[a] == 0.0) { anext[a] = INFINITY; next[a] = step[a] = last[a] = 0; continue; }
        double pin = (S[a] + amin * D[a] - vmin[a]) / d[a];    /* plane coordinate at entry */
        if (D[a] > 0.0) {
            step[a] = 1;
            next[a] = (int)floor(pin) + 1;
            /* guard against pin sitting a rounding error below an integer plane we are actually on */
            while (next[a] <= n[a] && plane_alpha(vmin[a], d[a], next[a], S[a], D[a]) <= amin) next[a]++;
            while (next[a] - 1 >= 0 && plane_alpha(vmin[a], d[a], next[a] - 1, S[a], D[a]) > amin) next[a]--;
            last[a] = n[a];
        } else {
            step[a] = -1;
            next[a] = (int)ceil(pin) - 1;
            while (next[a] >= 0 && plane_alpha(vmin[a], d[a], next[a], S[a], D[a]) <= amin) next[a]--;
            while (next[a] + 1 <= n[a] && plane_alpha(vmin[a], d[a], next[a] + 1, S[a], D[a]) > amin) next[a]++;
            last[a] = 0;
        }
        int in_range = step[a] > 0 ? (next[a] <= last[a]) : (next[a] >= last[a]);
        anext[a] = in_range ? plane_alpha(vmin[a], d[a], next[a], S[a], D[a]) : INFINITY;
    }

    int count = 0;
    double acur = amin;
    while (acur < amax) {
        /* smallest next crossing (paper: merge of the three sorted alpha sets, eq. 8) */
        int ax = 0;
        if (anext[1] < anext[ax]) ax = 1;
        if (anext[2] < anext[ax]) ax = 2;
        double anew = anext[ax] < amax ? anext[ax] : amax;
        if (anew > acur) {
            double amid = 0.5 * (anew + acur);                       /* paper eq. 12 */
            int idx[3];
            for (int a = 0; a < 3; ++a) {
                double p = (S[a] + amid * D[a] - vmin[a]) / d[a];
                int ii = (int)floor(p);
                if (ii < 0) ii = 0;
                if (ii > n[a] - 1) ii = n[a] - 1;
                idx[a] = ii;
            }
            int64_t vox = ((int64_t)idx[2] * n[1] + idx[1]) * n[0] + idx[0];
            sink(ctx, vox, labels[vox], len * (anew - acur));        /* paper eq. 10 */
            ++count;
            acur = anew;
        }
        if (anext[ax] >= amax) break;
        /* advance that axis; planes crossed at exactly the same alpha are consumed one after another and
         * produce zero-length segments, which are skipped by the test above                              */
        next[ax] += step[ax];
        int in_range = step[ax] > 0 ? (next[ax] <= last[ax]) : (next[ax] >= last[ax]);
        anext[ax] = in_range ? plane_alpha(vmin[ax], d[ax], next[ax], S[ax], D[ax]) : INFINITY;
    }
    return count;
}

typedef struct { double *L; } len_ctx;
static void len_sink(void *ctx, int64_t voxel, int label, double len)
{
    (void)voxel;
    ((len_ctx *)ctx)->L[label] += len;
}

typedef struct { int64_t *vox; uint8_t *lab; double *len; int cap, n; } trace_ctx;
static void trace_sink(void *ctx, int64_t voxel, int label, double len)
{
    trace_ctx *t = (trace_ctx *)ctx;
    if (t->n < t->cap) { t->vox[t->n] = voxel; t->lab[t->n] = (uint8_t)label; t->len[t->n] = len; }
    t->n++;
}

/* Per-material lengths for an explicit ray S->T (used by the analytic known-answer tests). */
int dexo_classic_lengths_ray(const double *S, const double *T, const uint8_t *labels,
                             int nx, int ny, int nz, double dx, double dy, double dz, int K, double *L)
{
    const int n[3] = { nx, ny, nz };
    const double d[3] = { dx, dy, dz };
    double D[3] = { T[0] - S[0], T[1] - S[1], T[2] - S[2] };
    for (int m = 0; m < K; ++m) L[m] = 0.0;
    len_ctx c = { L };
    return classic_walk(S, D, labels, n, d, len_sink, &c);
}

/* Ordered (voxel, label, length) list of one scanner ray.  Returns the segment count (may exceed cap). */
int dexo_classic_trace(const dexo_setup *g, const uint8_t *labels, int v, int r, int c,
                       int64_t *vox, uint8_t *lab, double *len, int cap)
{
    const int n[3] = { g->nx, g->ny, g->nz };
    const double d[3] = { g->dx, g->dy, g->dz };
    double S[3], D[3];
    ray_endpoints(g, v, r, c, S, D);
    trace_ctx t = { vox, lab, len, cap, 0 };
    classic_walk(S, D, labels, n, d, trace_sink, &t);
    return t.n;
}

/* L[(view - v0), row, ch, K] for views [v0, v0+nv).  OpenMP over (view,row).  seg_total (nullable) receives
 * the total number of voxel segments, which is the unit of the bytes-per-ray roofline.          */
void dexo_classic_lengths(const dexo_setup *g, const uint8_t *labels, int K, int v0, int nv,
                          double *L, int64_t *seg_total)
{
    const int n[3] = { g->nx, g->ny, g->nz };
    const double d[3] = { g->dx, g->dy, g->dz };
    int64_t total = 0;
    const int64_t lines = (int64_t)nv * g->n_rows;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : total)
    for (int64_t ln = 0; ln < lines; ++ln) {
        int v = v0 + (int)(ln / g->n_rows), r = (int)(ln % g->n_rows);
        for (int c = 0; c < g->n_ch; ++c) {
            double S[3], D[3];
            ray_endpoints(g, v, r, c, S, D);
            double *Lr = L + ((ln * g->n_ch) + c) * (int64_t)K;
            for (int m = 0; m < K; ++m) Lr[m] = 0.0;
            len_ctx ctx = { Lr };
            total += classic_walk(S, D, labels, n, d, len_sink, &ctx);
        }
    }
    if (seg_total) *seg_total = total;
}

/* ------------------------------------------------------------------------------------------------------ */
/* (C) polyenergetic Beer-Lambert detection (forward model of matdecomp.py:116-118 with K materials).
 * counts[s, ray] = sum_e W[s,e] exp(-sum_m L[ray,m] mu[m,e]);  logs[s, ray] = -ln(counts / sum_e W[s,e]).   */
void dexo_poly_detect(const double *L, int64_t n_rays, int K, const double *mu, int nE,
                      const double *W, int nSpec, double *counts, double *logs)
{
    double *air = (double *)malloc(sizeof(double) * (size_t)nSpec);
    for (int s = 0; s < nSpec; ++s) {
        double a = 0.0;
        for (int e = 0; e < nE; ++e) a += W[(size_t)s * nE + e];
        air[s] = a;
    }
#pragma omp parallel for schedule(static)
    for (int64_t ray = 0; ray < n_rays; ++ray) {
        const double *Lr = L + ray * K;
        double acc[8];
        double *accp = nSpec <= 8 ? acc : (double *)malloc(sizeof(double) * (size_t)nSpec);
        for (int s = 0; s < nSpec; ++s) accp[s] = 0.0;
        for (int e = 0; e < nE; ++e) {
            double t = 0.0;
            for (int m = 0; m < K; ++m) t += Lr[m] * mu[(size_t)m * nE + e];
            double att = exp(-t);
            for (int s = 0; s < nSpec; ++s) accp[s] += W[(size_t)s * nE + e] * att;
        }
        for (int s = 0; s < nSpec; ++s) {
            counts[(size_t)s * n_rays + ray] = accp[s];
            if (logs) logs[(size_t)s * n_rays + ray] = -log(accp[s] / air[s]);
        }
        if (accp != acc) free(accp);
    }
    free(air);
}

/* Whole forward projection of a view range on the host cores: the timed CPU baseline of bench.py.          */
void dexo_project_poly(const dexo_setup *g, const uint8_t *labels, int K, const double *mu, int nE,
                       const double *W, int nSpec, int v0, int nv, double *counts, double *logs,
                       int64_t *seg_total)
{
    const int64_t n_rays = (int64_t)nv * g->n_rows * g->n_ch;
    double *L = (double *)malloc(sizeof(double) * (size_t)n_rays * (size_t)K);
    dexo_classic_lengths(g, labels, K, v0, nv, L, seg_total);
    dexo_poly_detect(L, n_rays, K, mu, nE, W, nSpec, counts, logs);
    free(L);
}

/* ------------------------------------------------------------------------------------------------------ */
/* (B) CPU twin of the device "slab" traversal.  See DESIGN.md section 4 for the derivation; the CUDA source is
 * dex-ct-sim_b200/csrc/siddon_core.cuh.  Everything up to the per-column loop is float64 with one rounding per
 * operation; the loop runs in REAL = float or double.                                                       */

/* 3-D column routine whenever anything can leave the z = 0 plane (decided per scan, as on the device) */
static int scan_is_3d(const dexo_setup *g) { return g->nz > 1 || g->n_rows > 1 || g->zdet[0] != 0.0; }

typedef struct {
    int hit;                 /* 0 = ray misses the volume */
    int xmajor;              /* 1: columns are x-planes slabs, minor axis y; 0: columns along y, minor x */
    int np, nq;              /* voxel counts along major / minor axis */
    int i0, i1;              /* first / last column index touched */
    int fwd;                 /* 1 if the ray travels towards increasing column index */
    int i_c, j_c, k_c;       /* per-ray origin: column boundary in the middle of the range, voxel indices there */
    double m, mz;            /* minor / z slope per column, voxel units */
    double cq, cw;           /* q'(p') = cq + m p', w'(p') = cw + mz p', p' = column coordinate - i_c; cq, cw in [0,1) */
    double ua, ub;           /* clipped column-space interval, relative to i_c */
    double inv_am, inv_amz;  /* 1/|m|, 1/|mz| (capped at DEXO_BIG) */
    double seglen;           /* ray length per unit column [cm] */
} slab_ray;

#define DEXO_BIG 1e30

/* Few divisions on purpose (the device pays ~15 instructions for each): one reciprocal of the major step, one of each
 * slope, everything else multiplies by a reciprocal -- operation for operation as ray_setup() in siddon_core.cuh.     */
static void slab_setup(const dexo_setup *g, int v, int r, int c, slab_ray *s)
{
    double S[3], D[3];
    ray_endpoints(g, v, r, c, S, D);
    const int is3d = scan_is_3d(g);
    const double inv_dx = 1.0 / g->dx, inv_dy = 1.0 / g->dy, inv_dz = 1.0 / g->dz;
    if (!is3d) D[2] = 0.0;
    const double gx = D[0] * inv_dx, gy = D[1] * inv_dy, gz = is3d ? D[2] * inv_dz : 0.0;
    double gp, gq, pS, qS;
    s->hit = 0;
    s->xmajor = fabs(gx) >= fabs(gy);
    if (s->xmajor) { gp = gx; gq = gy; pS = S[0] * inv_dx; qS = S[1] * inv_dy; s->np = g->nx; s->nq = g->ny; }
    else           { gp = gy; gq = gx; pS = S[1] * inv_dy; qS = S[0] * inv_dx; s->np = g->ny; s->nq = g->nx; }
    const double inv_gp = 1.0 / gp;
    s->m = gq * inv_gp;
    s->mz = is3d ? gz * inv_gp : 0.0;
    const double q0 = qS - s->m * pS;
    const double w0 = is3d ? 0.0 - s->mz * pS : 0.0;
    const double hp = 0.5 * (double)s->np, hq = 0.5 * (double)s->nq, hw = 0.5 * (double)g->nz;
    double pe = pS + gp;
    double pa = pS < pe ? pS : pe, pb = pS < pe ? pe : pS;
    if (pa < -hp) pa = -hp;
    if (pb > hp) pb = hp;
    s->inv_am = DEXO_BIG;
    s->inv_amz = DEXO_BIG;
    if (s->m != 0.0) {
        const double inv_m = 1.0 / s->m;
        double t1 = (-hq - q0) * inv_m, t2 = (hq - q0) * inv_m;
        double lo = t1 < t2 ? t1 : t2, hi = t1 < t2 ? t2 : t1;
        if (lo > pa) pa = lo;
        if (hi < pb) pb = hi;
        s->inv_am = fabs(inv_m);
        if (s->inv_am > DEXO_BIG) s->inv_am = DEXO_BIG;
    } else if (!(q0 >= -hq && q0 < hq)) return;
    if (is3d) {
        if (s->mz != 0.0) {
            const double inv_mz = 1.0 / s->mz;
            double t1 = (-hw - w0) * inv_mz, t2 = (hw - w0) * inv_mz;
            double lo = t1 < t2 ? t1 : t2, hi = t1 < t2 ? t2 : t1;
            if (lo > pa) pa = lo;
            if (hi < pb) pb = hi;
            s->inv_amz = fabs(inv_mz);
            if (s->inv_amz > DEXO_BIG) s->inv_amz = DEXO_BIG;
        } else if (!(w0 >= -hw && w0 < hw)) return;
    }
    if (!(pa < pb)) return;
    const double ua = pa + hp, ub = pb + hp;
    int i0 = (int)floor(ua), i1 = (int)ceil(ub) - 1;
    if (i0 < 0) i0 = 0;
    if (i1 > s->np - 1) i1 = s->np - 1;
    if (i1 < i0) return;
    s->i0 = i0; s->i1 = i1;
    s->fwd = gp > 0.0;
    s->i_c = (i0 + i1 + 1) / 2;
    const double pc = (double)s->i_c - hp;
    const double qfull = s->m * pc + (q0 + hq);
    const double qfl = floor(qfull);
    s->j_c = (int)qfl;
    s->cq = qfull - qfl;
    if (is3d) {
        const double wfull = s->mz * pc + (w0 + hw);
        const double wfl = floor(wfull);
        s->k_c = (int)wfl;
        s->cw = wfull - wfl;
    } else {
        s->k_c = 0;
        s->cw = 0.5;
    }
    s->ua = ua - (double)s->i_c;
    s->ub = ub - (double)s->i_c;
    s->seglen = sqrt(D[0] * D[0] + D[1] * D[1] + D[2] * D[2]) * fabs(inv_gp);
    s->hit = 1;
}

static inline int64_t slab_voxel(const dexo_setup *g, const slab_ray *s, int i, int j, int k)
{
    return s->xmajor ? ((int64_t)k * g->ny + j) * g->nx + i : ((int64_t)k * g->ny + i) * g->nx + j;
}

/* One column visits up to three voxels (2-D: two).  piece order = order along increasing column coordinate;
 * the caller reverses it for rays travelling towards decreasing column index.  j, k are ABSOLUTE indices here. */
typedef struct { int j, k; double w; } slab_piece;

static inline int iclamp(int x, int lo, int hi) { return x < lo ? lo : (x > hi ? hi : x); }

#define SLAB_COLUMN(REAL, FMA, FLOOR, FMIN, FMAX)                                                           \
    const REAL fi = (REAL)(i - s->i_c);                                                                     \
    const REAL a = FMAX(fi, ua), b = FMIN(fi + (REAL)1, ub);                                                \
    const REAL wtot = b - a;                                                                                \
    int np_ = 0;                                                                                            \
    if (wtot > (REAL)0) {                                                                                   \
        const REAL qa = FMA(m, a, cq), qb = FMA(m, b, cq);                                                  \
        const REAL qmin = FMIN(qa, qb), qmax = FMAX(qa, qb);                                                \
        int jr = (int)FLOOR(qmin);                                                                          \
        const REAL qhi = FMIN(FMAX((qmax - (REAL)(jr + 1)) * inv_am, (REAL)0), wtot);                       \
        jr = iclamp(jr, -1 - s->j_c, s->nq - 1 - s->j_c);                                                   \
        const int j = jr + s->j_c;                                                                          \
        /* p-length before the minor crossing, and which voxel comes first along +p */                     \
        const REAL tq = (m >= (REAL)0) ? wtot - qhi : qhi;                                                  \
        const int jfirst = (m >= (REAL)0) ? j : j + 1, jsecond = (m >= (REAL)0) ? j + 1 : j;                \
        if (!is3d) {                                                                                        \
            const REAL w1 = (m >= (REAL)0) ? wtot - qhi : qhi, w2 = (m >= (REAL)0) ? qhi : wtot - qhi;      \
            pc[0].j = jfirst; pc[0].k = s->k_c; pc[0].w = (double)w1;                                       \
            pc[1].j = jsecond; pc[1].k = s->k_c; pc[1].w = (double)w2;                                      \
            np_ = 2;                                                                                        \
        } else {                                                                                            \
            const REAL wa = FMA(mz, a, cw), wb = FMA(mz, b, cw);                                            \
            const REAL wmin = FMIN(wa, wb), wmax = FMAX(wa, wb);                                            \
            int kr = (int)FLOOR(wmin);                                                                      \
            const REAL zhi = FMIN(FMAX((wmax - (REAL)(kr + 1)) * inv_amz, (REAL)0), wtot);                  \
            kr = iclamp(kr, -1 - s->k_c, nz_ - 1 - s->k_c);                                                 \
            const int k = kr + s->k_c;                                                                      \
            const REAL tz = (mz >= (REAL)0) ? wtot - zhi : zhi;                                             \
            const int kfirst = (mz >= (REAL)0) ? k : k + 1, ksecond = (mz >= (REAL)0) ? k + 1 : k;          \
            const REAL t1 = FMIN(tq, tz), t2 = FMAX(tq, tz);                                                \
            const REAL w1 = t1, w3 = wtot - t2, w2 = t2 - t1;                                               \
            pc[0].j = jfirst; pc[0].k = kfirst; pc[0].w = (double)w1;                                       \
            pc[1].j = (tq < tz) ? jsecond : jfirst; pc[1].k = (tq < tz) ? kfirst : ksecond;                 \
            pc[1].w = (double)w2;                                                                           \
            pc[2].j = jsecond; pc[2].k = ksecond; pc[2].w = (double)w3;                                     \
            np_ = 3;                                                                                        \
        }                                                                                                   \
    }                                                                                                       \
    return np_;

static int slab_column_f32(const slab_ray *s, int is3d, int nz_, int i, slab_piece pc[3])
{
    const float m = (float)s->m, mz = (float)s->mz, cq = (float)s->cq, cw = (float)s->cw;
    const float ua = (float)s->ua, ub = (float)s->ub;
    const float inv_am = (float)s->inv_am, inv_amz = (float)s->inv_amz;
    SLAB_COLUMN(float, fmaf, floorf, fminf, fmaxf)
}

static int slab_column_f64(const slab_ray *s, int is3d, int nz_, int i, slab_piece pc[3])
{
    const double m = s->m, mz = s->mz, cq = s->cq, cw = s->cw;
    const double ua = s->ua, ub = s->ub;
    const double inv_am = s->inv_am, inv_amz = s->inv_amz;
    SLAB_COLUMN(double, fma, floor, fmin, fmax)
}

/* Ordered trace of one scanner ray with the slab traversal.  use_f64: arithmetic of the column loop.
 * Zero-weight pieces and pieces that fall in the one-voxel safety apron (index -1 or n) are dropped, as on the
 * device.  len = weight * seglen.  Returns the number of segments (may exceed cap).                         */
int dexo_slab_trace(const dexo_setup *g, const uint8_t *labels, int v, int r, int c, int use_f64,
                    int64_t *vox, uint8_t *lab, double *len, int cap)
{
    slab_ray s;
    slab_setup(g, v, r, c, &s);
    if (!s.hit) return 0;
    const int is3d = scan_is_3d(g);
    int n = 0;
    const int step = s.fwd ? 1 : -1;
    for (int i = s.fwd ? s.i0 : s.i1; s.fwd ? (i <= s.i1) : (i >= s.i0); i += step) {
        slab_piece pc[3];
        int np_ = use_f64 ? slab_column_f64(&s, is3d, g->nz, i, pc) : slab_column_f32(&s, is3d, g->nz, i, pc);
        for (int t = 0; t < np_; ++t) {
            const slab_piece *p = &pc[s.fwd ? t : np_ - 1 - t];
            if (!(p->w > 0.0)) continue;
            if (p->j < 0 || p->j >= s.nq || p->k < 0 || p->k >= g->nz) continue;
            int64_t vx = slab_voxel(g, &s, i, p->j, p->k);
            double l = use_f64 ? p->w * s.seglen : (double)((float)p->w * (float)s.seglen);
            if (n < cap) { vox[n] = vx; lab[n] = labels[vx]; len[n] = l; }
            ++n;
        }
    }
    return n;
}

/* Per-material lengths with the slab traversal (float64 loop) -- used to cross-check (B) against (A). */
void dexo_slab_lengths(const dexo_setup *g, const uint8_t *labels, int K, int v0, int nv, int use_f64, double *L)
{
    const int64_t lines = (int64_t)nv * g->n_rows;
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t ln = 0; ln < lines; ++ln) {
        int v = v0 + (int)(ln / g->n_rows), r = (int)(ln % g->n_rows);
        const int is3d = scan_is_3d(g);
        (void)r;
        for (int c = 0; c < g->n_ch; ++c) {
            double *Lr = L + ((ln * g->n_ch) + c) * (int64_t)K;
            for (int m = 0; m < K; ++m) Lr[m] = 0.0;
            slab_ray s;
            slab_setup(g, v, r, c, &s);
            if (!s.hit) continue;
            for (int i = s.i0; i <= s.i1; ++i) {
                slab_piece pc[3];
                int np_ = use_f64 ? slab_column_f64(&s, is3d, g->nz, i, pc) : slab_column_f32(&s, is3d, g->nz, i, pc);
                for (int t = 0; t < np_; ++t) {
                    if (!(pc[t].w > 0.0)) continue;
                    if (pc[t].j < 0 || pc[t].j >= s.nq || pc[t].k < 0 || pc[t].k >= g->nz) continue;
                    Lr[labels[slab_voxel(g, &s, i, pc[t].j, pc[t].k)]] += pc[t].w;
                }
            }
            for (int m = 0; m < K; ++m) Lr[m] *= s.seglen;
        }
    }
}

/* A launcher may export OMP_NUM_THREADS=1 (torchrun does); the timed CPU baseline asks for its threads explicitly. */
void dexo_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int dexo_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
