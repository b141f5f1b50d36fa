// laser_slam_b200/csrc/plicp_solve.cuh -- single-thread closed-form pieces of PL-ICP:
// the constrained 4x4 least-squares solve of gpc_solve (Censi ICRA'08 eq. 31) and the pose algebra
// of icp_loop's termination test.  All FP64, compiled with -fmad=false so each operation
// rounds like the x86-64 (no FMA) build of CSM the reference links (src/zhcsm/CMakeLists.txt:2-9).
#pragma once
#include <math.h>

namespace plicp {

__device__ __forceinline__ double poly_eval(const double *c, int deg, double x) {
    double v = c[deg];
    for (int i = deg - 1; i >= 0; i--) v = v * x + c[i];
    return v;
}

// root of a polynomial that is monotone on [lo,hi] with a sign change: bisection-safeguarded Newton
__device__ __noinline__ double root_bracketed(const double *c, int deg, double lo, double hi) {
    double flo = poly_eval(c, deg, lo), fhi = poly_eval(c, deg, hi);
    if (flo == 0) return lo;
    if (fhi == 0) return hi;
    double dc[5];
    for (int i = 1; i <= deg; i++) dc[i - 1] = c[i] * i;
    double x = 0.5 * (lo + hi);
    for (int it = 0; it < 200; it++) {
        double fx = poly_eval(c, deg, x);
        if (fx == 0) return x;
        if ((fx < 0) == (flo < 0)) { lo = x; flo = fx; } else { hi = x; fhi = fx; }
        double dfx = poly_eval(dc, deg - 1, x);
        double xn = (dfx != 0) ? x - fx / dfx : lo - 1;
        if (!(xn > lo && xn < hi)) xn = 0.5 * (lo + hi);
        if (xn == x || hi - lo <= 0) break;
        if (fabs(xn - x) <= 1e-17 * fabs(x)) { x = xn; break; }
        x = xn;
    }
    return x;
}

// real roots (ascending) of c[0..deg], deg in 1..4, given the real roots of its derivative
__device__ __noinline__ int roots_from_crit(const double *c, int deg, const double *crit, int nc, double *roots) {
    double B = 0;
    for (int i = 0; i < deg; i++) { double t = fabs(c[i] / c[deg]); if (t > B) B = t; }
    B += 1.0;
    double pts[6];
    int np = 0;
    pts[np++] = -B;
    for (int i = 0; i < nc; i++) if (crit[i] > -B && crit[i] < B) pts[np++] = crit[i];
    pts[np++] = B;
    int nr = 0;
    for (int i = 0; i + 1 < np; i++) {
        double a = pts[i], b = pts[i + 1];
        double fa = poly_eval(c, deg, a), fb = poly_eval(c, deg, b);
        if (fa == 0) { if (nr == 0 || roots[nr - 1] != a) roots[nr++] = a; continue; }
        if ((fa < 0) != (fb < 0) && fb != 0) roots[nr++] = root_bracketed(c, deg, a, b);
        else if (fb == 0 && i + 2 == np) roots[nr++] = b;
    }
    return nr;
}

// greatest real root of the quartic a[0..4] (poly_greatest_real_root of gpc.c; GSL's eigenvalue solver is
// replaced by a bracketing solver that needs no libm call, so CPU oracle and GPU agree to the last bit)
__device__ inline bool greatest_real_root4(const double *a, double *root) {
    if (a[4] == 0) return false;
    double d3[4], d2[3], d1[2];
    for (int i = 1; i <= 4; i++) d3[i - 1] = a[i] * i;
    for (int i = 1; i <= 3; i++) d2[i - 1] = d3[i] * i;
    for (int i = 1; i <= 2; i++) d1[i - 1] = d2[i] * i;
    double r1[1], r2[2], r3[3], r4[4];
    r1[0] = -d1[0] / d1[1];
    int n2 = roots_from_crit(d2, 2, r1, 1, r2);
    int n3 = roots_from_crit(d3, 3, r2, n2, r3);
    int n4 = roots_from_crit(a, 4, r3, n3, r4);
    if (n4 == 0) return false;
    *root = r4[n4 - 1];
    return true;
}

__device__ inline bool inv2(const double A[2][2], double inv[2][2]) {
    double det = A[0][0] * A[1][1] - A[0][1] * A[1][0];
    if (det == 0 || isnan(det)) return false;
    inv[0][0] = A[1][1] / det;  inv[0][1] = -A[0][1] / det;
    inv[1][0] = -A[1][0] / det; inv[1][1] = A[0][0] / det;
    return true;
}

__device__ inline bool solve4(double A[4][4], double b[4], double x[4]) {
    for (int k = 0; k < 4; k++) {
        int piv = k;
        for (int r = k + 1; r < 4; r++) if (fabs(A[r][k]) > fabs(A[piv][k])) piv = r;
        if (A[piv][k] == 0 || isnan(A[piv][k])) return false;
        if (piv != k) {
            for (int c2 = 0; c2 < 4; c2++) { double t = A[k][c2]; A[k][c2] = A[piv][c2]; A[piv][c2] = t; }
            double t = b[k]; b[k] = b[piv]; b[piv] = t;
        }
        for (int r = k + 1; r < 4; r++) {
            double f = A[r][k] / A[k][k];
            for (int c2 = k; c2 < 4; c2++) A[r][c2] -= f * A[k][c2];
            b[r] -= f * b[k];
        }
    }
    for (int k = 3; k >= 0; k--) {
        double s = b[k];
        for (int c2 = k + 1; c2 < 4; c2++) s -= A[k][c2] * x[c2];
        x[k] = s / A[k][k];
    }
    return true;
}

// S[] = the 14 accumulated sums: M00 M01 M02 M03 M11 M12 M13 M22 M23 M33 g0 g1 g2 g3 (M symmetric: every
// mirrored pair in gpc.c's d_bigM is the same floating-point expression up to commutation)
__device__ __noinline__ bool gpc_solve_sums(const double *S, double x_out[3]) {
    double dM[4][4] = {{S[0], S[1], S[2], S[3]}, {S[1], S[4], S[5], S[6]}, {S[2], S[5], S[7], S[8]}, {S[3], S[6], S[8], S[9]}};
    double bigM[4][4], g[4];
    for (int a = 0; a < 4; a++) g[a] = -2 * S[10 + a];
    for (int a = 0; a < 4; a++) for (int b = 0; b < 4; b++) bigM[a][b] = 2 * dM[a][b];
    double mA[2][2] = {{bigM[0][0], bigM[0][1]}, {bigM[1][0], bigM[1][1]}};
    double mB[2][2] = {{bigM[0][2], bigM[0][3]}, {bigM[1][2], bigM[1][3]}};
    double mD[2][2] = {{bigM[2][2], bigM[2][3]}, {bigM[3][2], bigM[3][3]}};
    double mAi[2][2];
    if (!inv2(mA, mAi)) return false;
    double AiB[2][2], mS[2][2];
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) AiB[i][j] = mAi[i][0] * mB[0][j] + mAi[i][1] * mB[1][j];
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++)
        mS[i][j] = mD[i][j] + (-1.0) * (mB[0][i] * AiB[0][j] + mB[1][i] * AiB[1][j]);
    double mSi[2][2], mSa[2][2];
    if (!inv2(mS, mSi)) return false;
    double detS = mS[0][0] * mS[1][1] - mS[0][1] * mS[1][0];
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) mSa[i][j] = mSi[i][j] * detS;
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
    if (!greatest_real_root4(q, &lambda)) return false;
    double A[4][4], b[4], x[4];
    for (int i = 0; i < 4; i++) { for (int j = 0; j < 4; j++) A[i][j] = bigM[i][j]; b[i] = g[i]; }
    A[2][2] += 2 * lambda;
    A[3][3] += 2 * lambda;
    if (!solve4(A, b, x)) return false;
    for (int i = 0; i < 4; i++) x[i] = -x[i];
    x_out[0] = x[0];
    x_out[1] = x[1];
    x_out[2] = atan2(x[3], x[2]);
    if (isnan(x_out[0]) || isnan(x_out[1]) || isnan(x_out[2])) return false;
    return true;
}

// pose_diff_d(pose2, pose1) = ominus(pose1) (+) pose2, angle wrapped (math_utils.c)
__device__ inline void pose_diff(const double pose2[3], const double pose1[3], double res[3]) {
    const double PI = 3.14159265358979323846;
    double c = cos(pose1[2]), s = sin(pose1[2]);
    double t0 = -c * pose1[0] - s * pose1[1], t1 = s * pose1[0] - c * pose1[1], t2 = -pose1[2];
    double c2 = cos(t2), s2 = sin(t2);
    res[0] = t0 + c2 * pose2[0] - s2 * pose2[1];
    res[1] = t1 + s2 * pose2[0] + c2 * pose2[1];
    res[2] = t2 + pose2[2];
    while (res[2] > +PI) res[2] -= 2 * PI;
    while (res[2] < -PI) res[2] += 2 * PI;
}

}  // namespace plicp
