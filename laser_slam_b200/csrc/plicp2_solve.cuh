// laser_slam_b200/csrc/plicp2_solve.cuh -- the closed-form pieces of one PL-ICP iteration, executed by ONE WARP
// (all 32 lanes run the same scalar code redundantly; the polynomial root search uses the lanes as 32 probes):
//   gpc_solve          SURVEY 8a-7: constrained 4x4 least squares of CSM lib/gpc/gpc.c (Censi ICRA'08 eq. 31)
//   pose_diff          SURVEY 8a-8: icp_loop's termination test algebra (math_utils.c pose_diff_d)
// FP64, compiled with -fmad=false.  gsl_poly_complex_solve (companion-matrix eigenvalues) is replaced by an
// exact search on the monotone pieces of the quartic: parity is on lambda's value (few ulp), not on GSL's bits.
#pragma once
#include <math.h>

#ifndef PL_NAMESPACE
#define PL_NAMESPACE plicp2  // (plicp2_common.cuh: one namespace per instance file)
#endif

namespace PL_NAMESPACE {

__device__ __forceinline__ double poly_eval(const double *c, int deg, double x) {
    double v = c[deg];
#pragma unroll
    for (int i = deg - 1; i >= 0; i--) v = v * x + c[i];
    return v;
}

// Root of f = c[0..DEG] on [lo,hi], where f is monotone and f(lo), f(hi) have opposite signs (or one is 0).
// Warp-wide 33-section (each lane probes one point, the ballot locates the sign change): 4 rounds shrink the bracket
// 1.2e6 times, then Newton (quadratic) reaches rounding in 2-3 steps; if Newton leaves the bracket or stalls
// (ill-conditioned root) four more section rounds finish the job.
template <int DEG>
__device__ __forceinline__ void section_rounds(const double *c, double &lo, double &hi, bool lo_neg, int rounds) {
    const int lane = threadIdx.x & 31;
#pragma unroll 1
    for (int round = 0; round < rounds; round++) {
        const double w = (hi - lo) * (1.0 / 33.0);
        const double x = lo + w * (double)(lane + 1);
        const double fx = poly_eval(c, DEG, x);
        const bool same = (fx != 0) && ((fx < 0) == lo_neg);
        const unsigned m = __ballot_sync(0xffffffffu, same);
        const int cnt = (m == 0xffffffffu) ? 32 : (__ffs(~m) - 1);  // length of the same-sign prefix
        const double nlo = cnt == 0 ? lo : lo + w * (double)cnt;
        const double nhi = cnt == 32 ? hi : lo + w * (double)(cnt + 1);
        if (!(nhi > nlo)) break;
        lo = nlo;
        hi = nhi;
    }
}

// 1/x to ~1e-9 relative (hardware seed + one Newton step): the root iteration below only needs a good step, its fixed
// point does not depend on how the quotient was rounded
__device__ __forceinline__ double rcp_fast(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    const double e = 1.0 - x * r;
    return r + r * e;
}

template <int DEG>
__device__ __forceinline__ double root_monotone_warp(const double *c, double lo, double hi, double flo) {
    const bool lo_neg = flo < 0;
    section_rounds<DEG>(c, lo, hi, lo_neg, 4);
    double dc[DEG];
#pragma unroll
    for (int i = 1; i <= DEG; i++) dc[i - 1] = c[i] * i;
    double x = 0.5 * (lo + hi);
    bool converged = false;
#pragma unroll 1
    for (int it = 0; it < 5; it++) {
        const double fx = poly_eval(c, DEG, x);
        if (fx == 0) { converged = true; break; }
        const double dfx = poly_eval(dc, DEG - 1, x);
        if (dfx == 0) break;
        const double xn = x - fx * rcp_fast(dfx);
        if (!(xn >= lo && xn <= hi)) break;
        const double step = fabs(xn - x);
        x = xn;
        if (step <= 4e-16 * fabs(x)) { converged = true; break; }
    }
    if (converged) return x;
    section_rounds<DEG>(c, lo, hi, lo_neg, 6);
    return 0.5 * (lo + hi);
}

// upper bound on |root| for c[0..deg] (Fujiwara): 2*max_k |c[deg-k]/c[deg]|^(1/k)
__device__ __forceinline__ double root_bound(const double *c, int deg) {
    double b = 0;
    for (int k = 1; k <= deg; k++) {
        double t = fabs(c[deg - k] / c[deg]);
        if (k == deg) t *= 0.5;
        double v = k == 1 ? t : k == 2 ? sqrt(t) : k == 3 ? cbrt(t) : sqrt(sqrt(t));
        b = v > b ? v : b;
    }
    return 2.0 * b * (1.0 + 1e-9) + 1e-300;
}

// Real roots of the cubic c[0..3], produced lazily in DESCENDING order from its monotone pieces (split at the
// roots r2[0..n2) of its derivative).  *piece walks from the right-most piece down; returns false when exhausted.
__device__ __forceinline__ bool next_cubic_root_desc(const double *c, const double *pts, int np, int *piece,
                                                     double *last, bool *have_last, double *root) {
    while (*piece >= 0) {
        const int i = *piece;
        (*piece)--;
        const double a = pts[i], b = pts[i + 1];
        const double fa = poly_eval(c, 3, a), fb = poly_eval(c, 3, b);
        if (fb == 0) {
            if (!*have_last || *last != b) { *last = b; *have_last = true; *root = b; (*piece)++; 
                // revisit the same piece for a possible interior root / fa == 0 after reporting b
                // (handled by the branch below on the next call through the have_last guard)
                return true; }
        }
        if (fa != 0 && fb != 0 && ((fa < 0) != (fb < 0))) {
            const double r = root_monotone_warp<3>(c, a, b, fa);
            *last = r; *have_last = true; *root = r;
            return true;
        }
        if (fa == 0 && i == 0 && (!*have_last || *last != a)) { *last = a; *have_last = true; *root = a; return true; }
    }
    return false;
}

// greatest real root of the quartic a[0..4] (poly_greatest_real_root of gpc.c): walk the monotone pieces of the
// quartic from the right (the pieces are bounded by the real roots of its derivative, taken in descending order)
// and solve in the first piece whose ends differ in sign.  Typical cost: one cubic root + one quartic root.
static __device__ __noinline__ bool greatest_real_root4_warp(const double *a, double *root) {
    if (a[4] == 0 || isnan(a[4])) return false;
    double d3[4], d2[3];
#pragma unroll
    for (int i = 1; i <= 4; i++) d3[i - 1] = a[i] * i;
#pragma unroll
    for (int i = 1; i <= 3; i++) d2[i - 1] = d3[i] * i;
    // Every root of the derivative lies in the convex hull of the quartic's roots (Gauss-Lucas), so the quartic's
    // root bound also bounds the cubic's.  Monotone pieces of the cubic: [-B, roots of the quadratic d2, B]
    const double B4 = root_bound(a, 4);
    const double B3 = B4;
    double pts[4];
    int np = 0;
    pts[np++] = -B3;
    {
        const double A = d2[2], Bq = d2[1], Cq = d2[0];
        const double disc = Bq * Bq - 4 * A * Cq;
        if (disc > 0) {
            const double sq = sqrt(disc);
            const double qq = -0.5 * (Bq + (Bq >= 0 ? sq : -sq));
            const double x1 = qq / A, x2 = (qq != 0) ? Cq / qq : x1;
            const double lo = x1 < x2 ? x1 : x2, hi = x1 < x2 ? x2 : x1;
            if (lo > -B3 && lo < B3) pts[np++] = lo;
            if (hi > -B3 && hi < B3 && hi != lo) pts[np++] = hi;
        } else if (disc == 0) {
            const double x = -Bq / (2 * A);
            if (x > -B3 && x < B3) pts[np++] = x;
        }
    }
    pts[np++] = B3;
    int piece = np - 2;
    double last = 0;
    bool have_last = false;
    double hi = B4, fhi = poly_eval(a, 4, hi);
    if (fhi == 0) { *root = hi; return true; }
    if (isnan(fhi)) return false;
    // Usual case, without refining the critical point: the right-most piece of the cubic changes sign, a few section
    // rounds bracket that greatest critical point from above by cu, and the quartic changes sign on [cu, B] -- it is
    // monotone there, so that sign change is the greatest root.  Anything else takes the general path below.
    {
        const double ca = pts[np - 2], cb = pts[np - 1];
        const double fa = poly_eval(d3, 3, ca), fb = poly_eval(d3, 3, cb);
        if (fa != 0 && fb != 0 && ((fa < 0) != (fb < 0))) {
            double cl = ca, cu = cb;
            section_rounds<3>(d3, cl, cu, fa < 0, 4);
            const double fcu = poly_eval(a, 4, cu);
            if (cu < hi && fcu != 0 && ((fcu < 0) != (fhi < 0))) {
                *root = root_monotone_warp<4>(a, cu, hi, fcu);
                return true;
            }
        }
    }
#pragma unroll 1
    for (int guard = 0; guard < 8; guard++) {
        double c;
        const bool more = next_cubic_root_desc(d3, pts, np, &piece, &last, &have_last, &c);
        double lo = more ? c : -B4;
        if (lo >= hi) { if (!more) return false; continue; }
        if (lo < -B4) lo = -B4;
        const double flo = poly_eval(a, 4, lo);
        if (flo == 0) { *root = lo; return true; }
        if ((flo < 0) != (fhi < 0)) { *root = root_monotone_warp<4>(a, lo, hi, flo); return true; }
        if (!more || lo <= -B4) return false;
        hi = lo;
        fhi = flo;
    }
    return false;
}

__device__ __forceinline__ bool inv2(const double A[2][2], double inv[2][2]) {
    double det = A[0][0] * A[1][1] - A[0][1] * A[1][0];
    if (det == 0 || isnan(det)) return false;
    inv[0][0] = A[1][1] / det;  inv[0][1] = -A[0][1] / det;
    inv[1][0] = -A[1][0] / det; inv[1][1] = A[0][0] / det;
    return true;
}

__device__ __forceinline__ bool solve4(double A[4][4], double b[4], double x[4]) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
        int piv = k;
#pragma unroll
        for (int r = k + 1; r < 4; r++) if (fabs(A[r][k]) > fabs(A[piv][k])) piv = r;
        if (A[piv][k] == 0 || isnan(A[piv][k])) return false;
#pragma unroll
        for (int r = k + 1; r < 4; r++) {
            if (r == piv) {
#pragma unroll
                for (int c2 = 0; c2 < 4; c2++) { double t = A[k][c2]; A[k][c2] = A[r][c2]; A[r][c2] = t; }
                double t = b[k]; b[k] = b[r]; b[r] = t;
            }
        }
#pragma unroll
        for (int r = k + 1; r < 4; r++) {
            double f = A[r][k] / A[k][k];
#pragma unroll
            for (int c2 = k; c2 < 4; c2++) A[r][c2] -= f * A[k][c2];
            b[r] -= f * b[k];
        }
    }
#pragma unroll
    for (int k = 3; k >= 0; k--) {
        double s = b[k];
#pragma unroll
        for (int c2 = k + 1; c2 < 4; c2++) s -= A[k][c2] * x[c2];
        x[k] = s / A[k][k];
    }
    return true;
}

// S[] = the 14 accumulated sums: M00 M01 M02 M03 M11 M12 M13 M22 M23 M33 g0 g1 g2 g3 (M symmetric: every
// mirrored pair in gpc.c's d_bigM is the same floating-point expression up to commutation).  Whole-warp call.
template <class Clock>
static __device__ __noinline__ bool gpc_solve_warp(const double *S, double x_out[3], Clock *pc) {
    double dM[4][4] = {{S[0], S[1], S[2], S[3]}, {S[1], S[4], S[5], S[6]}, {S[2], S[5], S[7], S[8]}, {S[3], S[6], S[8], S[9]}};
    double bigM[4][4], g[4];
#pragma unroll
    for (int a = 0; a < 4; a++) g[a] = -2 * S[10 + a];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) bigM[a][b] = 2 * dM[a][b];
    double mA[2][2] = {{bigM[0][0], bigM[0][1]}, {bigM[1][0], bigM[1][1]}};
    double mB[2][2] = {{bigM[0][2], bigM[0][3]}, {bigM[1][2], bigM[1][3]}};
    double mD[2][2] = {{bigM[2][2], bigM[2][3]}, {bigM[3][2], bigM[3][3]}};
    double mAi[2][2];
    if (!inv2(mA, mAi)) return false;
    double AiB[2][2], mS[2][2];
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 2; j++) AiB[i][j] = mAi[i][0] * mB[0][j] + mAi[i][1] * mB[1][j];
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 2; j++) mS[i][j] = mD[i][j] + (-1.0) * (mB[0][i] * AiB[0][j] + mB[1][i] * AiB[1][j]);
    double mSi[2][2], mSa[2][2];
    if (!inv2(mS, mSi)) return false;
    double detS = mS[0][0] * mS[1][1] - mS[0][1] * mS[1][0];
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 2; j++) mSa[i][j] = mSi[i][j] * detS;
    double g1[2] = {g[0], g[1]}, g2[2] = {g[2], g[3]};
    double t12[2] = {g1[0] * mAi[0][0] + g1[1] * mAi[1][0], g1[0] * mAi[0][1] + g1[1] * mAi[1][1]};
    double m1[2] = {t12[0] * mB[0][0] + t12[1] * mB[1][0], t12[0] * mB[0][1] + t12[1] * mB[1][1]};
    double m2[2] = {m1[0] * mSa[0][0] + m1[1] * mSa[1][0], m1[0] * mSa[0][1] + m1[1] * mSa[1][1]};
    double m3[2] = {g2[0] * mSa[0][0] + g2[1] * mSa[1][0], g2[0] * mSa[0][1] + g2[1] * mSa[1][1]};
#define PLICP_DOT(a, b) ((a)[0] * (b)[0] + (a)[1] * (b)[1])
    double p[3] = {PLICP_DOT(m2, m2) - 2 * PLICP_DOT(m2, m3) + PLICP_DOT(m3, m3),
                   4 * PLICP_DOT(m2, m1) - 8 * PLICP_DOT(m2, g2) + 4 * PLICP_DOT(g2, m3),
                   4 * PLICP_DOT(m1, m1) - 8 * PLICP_DOT(m1, g2) + 4 * PLICP_DOT(g2, g2)};
#undef PLICP_DOT
    double l[3] = {detS, 2 * mS[0][0] + 2 * mS[1][1], 4};
    double q[5] = {p[0] - (l[0] * l[0]), p[1] - (2 * l[1] * l[0]), p[2] - (l[1] * l[1] + 2 * l[0] * l[2]),
                   -(2 * l[2] * l[1]), -(l[2] * l[2])};
    double lambda;
    pc->mark(12);
    if (!greatest_real_root4_warp(q, &lambda)) return false;
    pc->mark(13);
    double A[4][4], b[4], x[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
#pragma unroll
        for (int j = 0; j < 4; j++) A[i][j] = bigM[i][j];
        b[i] = g[i];
    }
    A[2][2] += 2 * lambda;
    A[3][3] += 2 * lambda;
    if (!solve4(A, b, x)) return false;
    pc->mark(14);
    x_out[0] = -x[0];
    x_out[1] = -x[1];
    x_out[2] = atan2(-x[3], -x[2]);
    if (isnan(x_out[0]) || isnan(x_out[1]) || isnan(x_out[2])) return false;
    return true;
}

// pose_diff_d(pose2, pose1) = ominus(pose1) (+) pose2, angle wrapped (math_utils.c); c,s = cos/sin(pose1[2])
__device__ __forceinline__ void pose_diff(const double pose2[3], const double pose1[3], double c, double s, double res[3]) {
    const double PI = 3.14159265358979323846;
    double t0 = -c * pose1[0] - s * pose1[1], t1 = s * pose1[0] - c * pose1[1], t2 = -pose1[2];
    double c2 = c, s2 = -s;  // cos(-a), sin(-a)
    res[0] = t0 + c2 * pose2[0] - s2 * pose2[1];
    res[1] = t1 + s2 * pose2[0] + c2 * pose2[1];
    res[2] = t2 + pose2[2];
    while (res[2] > +PI) res[2] -= 2 * PI;
    while (res[2] < -PI) res[2] += 2 * PI;
}

}  // namespace PL_NAMESPACE
