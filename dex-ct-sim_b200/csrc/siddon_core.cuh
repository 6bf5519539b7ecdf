// Ray set-up and the "slab" form of Siddon's traversal, shared by the production and the trace kernel.
//
// Replaces the tracer inside get_sino (/root/reference/main.py:120; source absent from the reference tree, the
// algorithm is Siddon 1985 as cited at /root/reference/README.md:28,41).  DESIGN.md section 4 derives the form
// used here; oracle/siddon_oracle.c section (B) is its operation-for-operation CPU twin and section (A) there
// (the paper's alpha-merge) is the independent truth both are tested against.
//
// Idea: pick the in-plane axis along which the ray advances fastest (the MAJOR axis, coordinate p, "columns").
// Inside one column the ray crosses at most one plane of the other in-plane axis (MINOR, coordinate q) and at
// most one z plane, because |dq/dp| <= 1 and CT cone angles keep |dz/dp| <= 1.  So a column contributes to at
// most two (2-D) or three (3-D) voxels, found with one floor() per axis -- no alpha lists, no sorting, and every
// column is independent of the previous one (no serial dependency, unlike the incremental alpha update).
// The lengths are exactly Siddon's: column share w in [0,1] times the ray length per column.
#pragma once
#include "common.cuh"

namespace dexct {

struct WsLayout {
    size_t off_cosb, off_sinb, off_cosg, off_sing, off_zdet;   // double tables
    size_t off_xm, off_ym;                                     // uint16 pair layouts (bytes from workspace start)
    size_t off_skip[4];                                        // uint16 skip layouts (2-D scans only): x+, x-, y+, y-
    bool has_skip, skip_3d;
    size_t off_rz[2];                                          // z-homogeneity radius volumes (3-D scans only)
    size_t total;
    int pitch_xm, pitch_ym;                                    // entries per column (nq + 2, rounded up)
    long long plane_xm, plane_ym;                              // entries per z plane
};

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// 3-D column routine whenever anything can leave the z = 0 plane (decided per scan)
inline bool scan_is_3d(const DexctScan &s, const DexctVolume &v)
{
    return v.nz > 1 || s.n_rows > 1 || s.z0 != 0.0;
}

inline WsLayout ws_layout(const DexctScan &s, const DexctVolume &v)
{
    WsLayout L;
    size_t o = 0;
    L.off_cosb = o; o = align_up(o + sizeof(double) * (size_t)s.n_views, 256);
    L.off_sinb = o; o = align_up(o + sizeof(double) * (size_t)s.n_views, 256);
    L.off_cosg = o; o = align_up(o + sizeof(double) * (size_t)s.n_ch, 256);
    L.off_sing = o; o = align_up(o + sizeof(double) * (size_t)s.n_ch, 256);
    L.off_zdet = o; o = align_up(o + sizeof(double) * (size_t)s.n_rows, 256);
    L.pitch_xm = (int)align_up((size_t)v.ny + 2, 8);
    L.pitch_ym = (int)align_up((size_t)v.nx + 2, 8);
    L.plane_xm = (long long)v.nx * L.pitch_xm;
    L.plane_ym = (long long)v.ny * L.pitch_ym;
    L.off_xm = o; o = align_up(o + sizeof(uint16_t) * (size_t)L.plane_xm * (size_t)(v.nz + 2), 256);
    L.off_ym = o; o = align_up(o + sizeof(uint16_t) * (size_t)L.plane_ym * (size_t)(v.nz + 2), 256);
    // skip layouts, one per (major axis, sign of the minor slope).  2-D scans: one plane, indexed like one plane of the pair
    // layouts.  3-D scans: indexed like the whole pair layout (z aprons included), plus two byte volumes of the per-voxel
    // z-homogeneity radius that the sweep reads (rz [nz][ny][nx] and its in-plane transpose [nz][nx][ny]).
    L.has_skip = true;
    L.skip_3d = scan_is_3d(s, v);
    const size_t planes = L.skip_3d ? (size_t)v.nz + 2 : 1;
    for (int a = 0; a < 4; ++a) {
        L.off_skip[a] = o;
        o = align_up(o + sizeof(uint16_t) * (size_t)(a < 2 ? L.plane_xm : L.plane_ym) * planes, 256);
    }
    L.off_rz[0] = L.off_rz[1] = o;
    if (L.skip_3d) {
        for (int a = 0; a < 2; ++a) {
            L.off_rz[a] = o;
            o = align_up(o + (size_t)v.nx * v.ny * v.nz, 256);
        }
    }
    L.total = o;
    return L;
}

// Everything the kernels need about scan + volume + workspace, passed by value.
struct SiddonParams {
    DexctScan scan;
    DexctVolume vol;
    const double *cosb, *sinb, *cosg, *sing, *zdet;
    const uint16_t *xm, *ym;          // pair layouts; element (k, i, j) at ((k+1)*plane + i*pitch + (j+1))
    const uint16_t *skip[4];          // skip layouts (null when unavailable); element (k, i, j) at ((k+1)*plane + i*pitch + (j+1))
                                      // for 3-D scans, (i, j) at (i*pitch + (j+1)) for 2-D scans
    int skip_cap[8];                  // 3-D: longest run allowed at z-homogeneity level l (radius l + 1 planes); SKIP_LEVELS used
    int pitch_xm, pitch_ym;
    long long plane_xm, plane_ym;
    double inv_dx, inv_dy, inv_dz;    // 1/voxel size, rounded once on the host
    int row_band;                     // 3-D production kernel: detector rows per band of its launch order (set per launch)
};

// Per-ray constants, computed in double with one rounding per operation (the CPU twin does the same).
//
// Coordinates of the column loop are RE-CENTRED PER RAY: the column coordinate is measured from the integer column
// boundary i_c in the middle of the ray's clipped column range, the minor / z coordinates from the integer voxel
// indices j_c / k_c the ray occupies there.  The loop therefore evaluates q'(p') = m p' + cq with |p'| <= np/2 and
// cq in [0,1): |q'| <= 1 + |m| np/2, so its float32 rounding error scales with |m| and the crossing position
// (q' - integer) / |m| keeps an absolute accuracy of about 2^-24 np columns for EVERY slope, shallow ones included.
struct RaySetup {
    bool hit, xmajor, fwd;
    int np, nq, i0, i1, i_c, j_c, k_c;
    double m, mz, cq, cw, ua, ub, inv_am, inv_amz, seglen;     // ua, ub relative to i_c
};

#define DEXCT_BIG 1e30

// Divisions are few on purpose (one reciprocal of the major step, one of the slope, one for the length): everything else
// multiplies by a reciprocal, in the same order as oracle/siddon_oracle.c slab_setup().  IS3D = false compiles the z part
// out; the values it would produce for a single-slice scan (mz = 0, k_c = 0, cw = 1/2) are set directly.
template <bool IS3D>
__device__ __forceinline__ void ray_setup(const SiddonParams &P, int v, int r, int c, RaySetup &s)
{
    typedef Rn<double> D;
    const double cb = P.cosb[v], sb = P.sinb[v], cg = P.cosg[c], sg = P.sing[c];
    const double ex = D::sub(D::mul(cb, cg), D::mul(sb, sg));
    const double ey = D::add(D::mul(sb, cg), D::mul(cb, sg));
    const double Sx = D::mul(P.scan.sid, cb), Sy = D::mul(P.scan.sid, sb);
    const double Dx = -D::mul(P.scan.sdd, ex), Dy = -D::mul(P.scan.sdd, ey), Dz = IS3D ? P.zdet[r] : 0.0;
    const double gx = D::mul(Dx, P.inv_dx), gy = D::mul(Dy, P.inv_dy), gz = IS3D ? D::mul(Dz, P.inv_dz) : 0.0;
    double gp, gq, pS, qS;
    s.hit = false;
    s.xmajor = fabs(gx) >= fabs(gy);
    if (s.xmajor) { gp = gx; gq = gy; pS = D::mul(Sx, P.inv_dx); qS = D::mul(Sy, P.inv_dy); s.np = P.vol.nx; s.nq = P.vol.ny; }
    else          { gp = gy; gq = gx; pS = D::mul(Sy, P.inv_dy); qS = D::mul(Sx, P.inv_dx); s.np = P.vol.ny; s.nq = P.vol.nx; }
    const double inv_gp = D::div(1.0, gp);
    s.m = D::mul(gq, inv_gp);
    s.mz = IS3D ? D::mul(gz, inv_gp) : 0.0;
    const double q0 = D::sub(qS, D::mul(s.m, pS));
    const double w0 = IS3D ? D::sub(0.0, D::mul(s.mz, pS)) : 0.0;         // the source sits in the z = 0 plane
    const double hp = D::mul(0.5, (double)s.np), hq = D::mul(0.5, (double)s.nq), hw = D::mul(0.5, (double)P.vol.nz);
    const double pe = D::add(pS, gp);
    double pa = pS < pe ? pS : pe, pb = pS < pe ? pe : pS;
    if (pa < -hp) pa = -hp;
    if (pb > hp) pb = hp;
    s.inv_am = DEXCT_BIG;
    s.inv_amz = DEXCT_BIG;
    if (s.m != 0.0) {
        const double inv_m = D::div(1.0, s.m);
        const double t1 = D::mul(D::sub(-hq, q0), inv_m), t2 = D::mul(D::sub(hq, q0), inv_m);
        const double lo = t1 < t2 ? t1 : t2, hi = t1 < t2 ? t2 : t1;
        if (lo > pa) pa = lo;
        if (hi < pb) pb = hi;
        s.inv_am = fabs(inv_m);
        if (s.inv_am > DEXCT_BIG) s.inv_am = DEXCT_BIG;
    } else if (!(q0 >= -hq && q0 < hq)) return;
    if (IS3D) {
        if (s.mz != 0.0) {
            const double inv_mz = D::div(1.0, s.mz);
            const double t1 = D::mul(D::sub(-hw, w0), inv_mz), t2 = D::mul(D::sub(hw, w0), inv_mz);
            const double lo = t1 < t2 ? t1 : t2, hi = t1 < t2 ? t2 : t1;
            if (lo > pa) pa = lo;
            if (hi < pb) pb = hi;
            s.inv_amz = fabs(inv_mz);
            if (s.inv_amz > DEXCT_BIG) s.inv_amz = DEXCT_BIG;
        } else if (!(w0 >= -hw && w0 < hw)) return;
    }
    if (!(pa < pb)) return;
    const double ua = D::add(pa, hp), ub = D::add(pb, hp);          // column coordinates in [0, np]
    int i0 = (int)floor(ua), i1 = (int)ceil(ub) - 1;
    if (i0 < 0) i0 = 0;
    if (i1 > s.np - 1) i1 = s.np - 1;
    if (i1 < i0) return;
    s.i0 = i0; s.i1 = i1;
    s.fwd = gp > 0.0;
    s.i_c = (i0 + i1 + 1) / 2;
    const double pc = D::sub((double)s.i_c, hp);
    const double qfull = D::add(D::mul(s.m, pc), D::add(q0, hq));   // minor index coordinate at column boundary i_c
    const double qfl = floor(qfull);
    s.j_c = (int)qfl;
    s.cq = D::sub(qfull, qfl);
    if (IS3D) {
        const double wfull = D::add(D::mul(s.mz, pc), D::add(w0, hw));
        const double wfl = floor(wfull);
        s.k_c = (int)wfl;
        s.cw = D::sub(wfull, wfl);
    } else {
        s.k_c = 0;
        s.cw = 0.5;
    }
    s.ua = D::sub(ua, (double)s.i_c);
    s.ub = D::sub(ub, (double)s.i_c);
    const double len2 = D::add(D::add(D::mul(Dx, Dx), D::mul(Dy, Dy)), D::mul(Dz, Dz));
    s.seglen = D::mul(__dsqrt_rn(len2), fabs(inv_gp));
    s.hit = true;
}

template <typename R> struct FloorInt;
template <> struct FloorInt<float> { static __device__ __forceinline__ int of(float x) { return __float2int_rd(x); } };
template <> struct FloorInt<double> { static __device__ __forceinline__ int of(double x) { return __double2int_rd(x); } };

// Loop constants in the compute precision.
template <typename R> struct RayLoop {
    R m, mz, cq, cw, ua, ub, inv_am, inv_amz;
    bool mpos, mzpos;
    int i_c, jmin, jmax, kmin, kmax;      // clamp range of the RELATIVE minor / z index (memory safety only)
    const uint16_t *pairs;                // this ray's layout, offset to (column 0, row j_c, plane k_c)
    int pitch;
    long long plane;
    __device__ __forceinline__ void init(const SiddonParams &P, const RaySetup &s)
    {
        m = (R)s.m; mz = (R)s.mz; cq = (R)s.cq; cw = (R)s.cw; ua = (R)s.ua; ub = (R)s.ub;
        inv_am = (R)s.inv_am; inv_amz = (R)s.inv_amz;
        mpos = m >= (R)0; mzpos = mz >= (R)0;
        i_c = s.i_c;
        jmin = -1 - s.j_c; jmax = s.nq - 1 - s.j_c;
        kmin = -1 - s.k_c; kmax = P.vol.nz - 1 - s.k_c;
        pitch = s.xmajor ? P.pitch_xm : P.pitch_ym;
        plane = s.xmajor ? P.plane_xm : P.plane_ym;
        pairs = (s.xmajor ? P.xm : P.ym) + (long long)(s.k_c + 1) * plane + (s.j_c + 1);
    }
};

// One column, general form (partial first / last columns included).  Calls sink(slot, jrel, krel, label, w) for
// slot = 0,1(,2) in order of INCREASING column coordinate; jrel / krel are relative to (j_c, k_c).  Weights may be
// zero; indices may sit in the one-voxel apron (-1 or n) with a weight at rounding level.
template <typename R, bool IS3D, typename Sink>
__device__ __forceinline__ void slab_column(const RayLoop<R> &L, int i, Sink &sink)
{
    typedef Rn<R> A;
    const R fi = (R)(i - L.i_c);
    const R a = A::max(fi, L.ua), b = A::min(A::add(fi, (R)1), L.ub);
    const R wtot = A::sub(b, a);
    if (!(wtot > (R)0)) return;
    const R qa = A::fma(L.m, a, L.cq), qb = A::fma(L.m, b, L.cq);
    const R qmin = A::min(qa, qb), qmax = A::max(qa, qb);
    int j = FloorInt<R>::of(qmin);
    const R qhi = A::min(A::max(A::mul(A::sub(qmax, (R)(j + 1)), L.inv_am), (R)0), wtot);
    j = min(max(j, L.jmin), L.jmax);                     // memory safety only; a no-op for finite geometry
    const uint16_t *col = L.pairs + (long long)i * L.pitch + j;
    if (!IS3D) {
        const unsigned pr = *col;
        const int lo = pr & 0xff, hi = pr >> 8;
        const R wlo = A::sub(wtot, qhi);
        if (L.mpos) { sink(0, j, 0, lo, wlo); sink(1, j + 1, 0, hi, qhi); }
        else        { sink(0, j + 1, 0, hi, qhi); sink(1, j, 0, lo, wlo); }
    } else {
        const R wa = A::fma(L.mz, a, L.cw), wb = A::fma(L.mz, b, L.cw);
        const R wmin = A::min(wa, wb), wmax = A::max(wa, wb);
        int k = FloorInt<R>::of(wmin);
        const R zhi = A::min(A::max(A::mul(A::sub(wmax, (R)(k + 1)), L.inv_amz), (R)0), wtot);
        k = min(max(k, L.kmin), L.kmax);
        const unsigned p0 = col[(long long)k * L.plane];
        const unsigned p1 = col[(long long)(k + 1) * L.plane];
        // p-length spent before the q / z crossing, walking towards +p
        const R tq = L.mpos ? A::sub(wtot, qhi) : qhi;
        const R tz = L.mzpos ? A::sub(wtot, zhi) : zhi;
        const int jA = L.mpos ? j : j + 1, jB = L.mpos ? j + 1 : j;        // first / second along +p
        const int kA = L.mzpos ? k : k + 1, kB = L.mzpos ? k + 1 : k;
        const unsigned pA = L.mzpos ? p0 : p1, pB = L.mzpos ? p1 : p0;     // pair of plane kA / kB
        const int shA = L.mpos ? 0 : 8, shB = L.mpos ? 8 : 0;              // byte of jA / jB inside a pair
        const R t1 = A::min(tq, tz), t2 = A::max(tq, tz);
        const R w1 = t1, w3 = A::sub(wtot, t2), w2 = A::sub(t2, t1);
        const bool qfirst = tq < tz;                                       // q plane is crossed before z plane
        sink(0, jA, kA, (pA >> shA) & 0xff, w1);
        if (qfirst) sink(1, jB, kA, (pA >> shB) & 0xff, w2);
        else        sink(1, jA, kB, (pB >> shA) & 0xff, w2);
        sink(2, jB, kB, (pB >> shB) & 0xff, w3);
    }
}

}  // namespace dexct
