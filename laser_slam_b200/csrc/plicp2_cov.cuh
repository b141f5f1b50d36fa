// laser_slam_b200/csrc/plicp2_cov.cuh -- compute_covariance_exact of CSM (icp_covariance.c; Censi ICRA'07), the
// closed-form covariance sm_icp returns when do_compute_covariance is set (CCanonicalMatcher::setCov,
// src/zhcsm/ZHCanonical_Matcher.cpp:79-81; read back at :292-296).  SURVEY 8a-8 / 8f-2.
//
//   cov0 = H^-1 (G1 G1' + G2 G2') H^-T,   H = d2J/dx2 (3x3),  G1 = d2J/dx dy1 (3 x n_ref),  G2 = d2J/dx dy2 (3 x n_sens)
//   cov_x = sigma^2 cov0
//
// One pass over the kept correspondences accumulates H (6 sums) and G2 G2' (6 sums; column i of G2 belongs to sens
// ray i alone) and scatter-adds the columns j1, j2 of G1 into shared memory (FP64 atomicAdd; three arrays that
// reuse rr2 / d2j / the jump tables, all dead by now); a second pass reduces G1 G1'.  The reference's
// cos/sin/atan2 calls are evaluated algebraically (unit normal from the segment, angle sums from the cos/sin table):
// equal up to rounding, parity tolerance on the covariance is relative 1e-6.
#pragma once
#include "plicp2_common.cuh"

namespace PL_NAMESPACE {

// C_k = n n' with n = (cos a, sin a), a = pi/2 + atan2(j1y - j2y, j1x - j2x)
__device__ __forceinline__ void compute_C_k(double j1x, double j1y, double j2x, double j2y, double C[3]) {
    const double dx = j1x - j2x, dy = j1y - j2y;
    const double inv = 1.0 / sqrt(dx * dx + dy * dy);
    const double c = -dy * inv, s = dx * inv;
    C[0] = c * c; C[1] = c * s; C[2] = s * s;
}

// (1/eps) (C(p1 + eps p1/|p1|, p2) - C(p1, p2)), eps = 0.001
__device__ __forceinline__ void dC_drho(double p1x, double p1y, double p2x, double p2y, double dC[3]) {
    const double eps = 0.001;
    double Ck[3], Ce[3];
    compute_C_k(p1x, p1y, p2x, p2y, Ck);
    const double nrm = sqrt(p1x * p1x + p1y * p1y);
    compute_C_k(p1x + (eps / nrm) * p1x, p1y + (eps / nrm) * p1y, p2x, p2y, Ce);
#pragma unroll
    for (int a = 0; a < 3; a++) dC[a] = (1 / eps) * (Ce[a] - Ck[a]);
}

// u' C v for symmetric C = [C0 C1; C1 C2]
__device__ __forceinline__ double quad2(double ux, double uy, const double C[3], double vx, double vy) {
    return ux * (C[0] * vx + C[1] * vy) + uy * (C[1] * vx + C[2] * vy);
}

// Block-wide sum of NV doubles per thread; result in out[] on every thread.  Two barriers.
template <int T, int NV>
__device__ __forceinline__ void block_sum_f64(double (&v)[NV], double *red, double (&out)[NV]) {
    constexpr int NW = T / 32;
    const int lane = threadIdx.x & 31, warp = pl_tid<T>() >> 5;
#pragma unroll
    for (int q = 0; q < NV; q++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_down_sync(0xffffffffu, v[q], o);
    }
    pl_sync<T>();  // red may still be read by a previous user
    if (lane == 0) {
#pragma unroll
        for (int q = 0; q < NV; q++) red[warp * NSUM + q] = v[q];
    }
    pl_sync<T>();
#pragma unroll
    for (int q = 0; q < NV; q++) {
        double s = 0;
        for (int w = 0; w < NW; w++) s += red[w * NSUM + q];
        out[q] = s;
    }
}

__device__ __forceinline__ bool inv3(const double A[3][3], double inv[3][3]) {
    const double det = A[0][0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) - A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
                       A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
    if (det == 0 || isnan(det)) return false;
    inv[0][0] = (A[1][1] * A[2][2] - A[1][2] * A[2][1]) / det;
    inv[0][1] = (A[0][2] * A[2][1] - A[0][1] * A[2][2]) / det;
    inv[0][2] = (A[0][1] * A[1][2] - A[0][2] * A[1][1]) / det;
    inv[1][0] = (A[1][2] * A[2][0] - A[1][0] * A[2][2]) / det;
    inv[1][1] = (A[0][0] * A[2][2] - A[0][2] * A[2][0]) / det;
    inv[1][2] = (A[0][2] * A[1][0] - A[0][0] * A[1][2]) / det;
    inv[2][0] = (A[1][0] * A[2][1] - A[1][1] * A[2][0]) / det;
    inv[2][1] = (A[0][1] * A[2][0] - A[0][0] * A[2][1]) / det;
    inv[2][2] = (A[0][0] * A[1][1] - A[0][1] * A[1][0]) / det;
    return true;
}

// Called by the whole CTA after the final correspondences are in cj1/cj2.  x = best pose, (ct, st) = cos/sin(x[2]).
// Writes sigma^2 * cov0 as the upper triangle xx xy xt yy yt tt into cov[6] (thread 0's copy is the one used).
template <int T>
__device__ __forceinline__ void covariance_exact(const Lay &l, int n, int C, const double2 *cs_table, double x0,
                                                 double x1, double ct, double st, double sigma, double cov[6]) {
    const int t = pl_tid<T>();
    const double2 *refp = PL_SH(const double2, l.refp);
    const float *sensr = PL_SH(const float, l.sensr);
    const short *cj1 = PL_SH(const short, l.cj1), *cj2 = PL_SH(const short, l.cj2);
    double *g1x = PL_SH(double, l.rr2), *g1y = PL_SH(double, l.d2j), *g1t = PL_SH(double, l.jt);
    double *red = PL_SH(double, l.red);
    pl_sync<T>();  // everybody is done with rr2 / d2j / jt
    for (int j = t; j < n; j += T) { g1x[j] = 0; g1y[j] = 0; g1t[j] = 0; }
    pl_sync<T>();
    double acc[12];
#pragma unroll
    for (int q = 0; q < 12; q++) acc[q] = 0;
#pragma unroll 1
    for (int k = 0; k < C; k++) {
        const int i = k * T + t;
        if (i >= n) break;
        const int j1 = cj1[i];
        if (j1 < 0) continue;
        const int j2 = cj2[i];
        const double2 p = sens_point(sensr, cs_table, i), q1 = refp[j1], q2 = refp[j2];
        // rotations of p by theta + pi/2 and theta + pi: (-st, ct) and (-ct, -st)
        const double v1x = -st * p.x - ct * p.y, v1y = ct * p.x - st * p.y;
        const double v2x = ct * p.x - st * p.y + x0 - q1.x, v2y = st * p.x + ct * p.y + x1 - q1.y;
        const double2 ci = cs_table[i];
        const double v3x = ct * ci.x - st * ci.y, v3y = st * ci.x + ct * ci.y;  // cos/sin(theta + theta_i)
        const double v4x = -v3y, v4y = v3x;                                      // ... + pi/2
        const double vnx = -(ct * p.x - st * p.y), vny = -(st * p.x + ct * p.y);
        double Ck[3];
        compute_C_k(q1.x, q1.y, q2.x, q2.y, Ck);
        acc[0] += 2.0 * Ck[0];
        acc[1] += 2.0 * Ck[1];
        acc[2] += 2.0 * Ck[2];
        acc[3] += 2.0 * (Ck[0] * v1x + Ck[1] * v1y);
        acc[4] += 2.0 * (Ck[1] * v1x + Ck[2] * v1y);
        acc[5] += 2.0 * (quad2(v2x, v2y, Ck, vnx, vny) + quad2(v1x, v1y, Ck, v1x, v1y));
        // measurement rho_i of the second scan: column i of G2
        const double a = 2.0 * (Ck[0] * v3x + Ck[1] * v3y);
        const double b = 2.0 * (Ck[1] * v3x + Ck[2] * v3y);
        const double c = 2.0 * (quad2(v2x, v2y, Ck, v4x, v4y) + quad2(v3x, v3y, Ck, v1x, v1y));
        acc[6] += a * a; acc[7] += a * b; acc[8] += a * c; acc[9] += b * b; acc[10] += b * c; acc[11] += c * c;
        // measurements rho_j1, rho_j2 of the first scan: columns j1, j2 of G1
        double dC1[3], dC2[3];
        dC_drho(q1.x, q1.y, q2.x, q2.y, dC1);
        dC_drho(q2.x, q2.y, q1.x, q1.y, dC2);
        const double2 cj = cs_table[j1];
        atomicAdd(&g1x[j1], -2.0 * (Ck[0] * cj.x + Ck[1] * cj.y) + 2.0 * (dC1[0] * v2x + dC1[1] * v2y));
        atomicAdd(&g1y[j1], -2.0 * (Ck[1] * cj.x + Ck[2] * cj.y) + 2.0 * (dC1[1] * v2x + dC1[2] * v2y));
        atomicAdd(&g1t[j1], -2.0 * quad2(cj.x, cj.y, Ck, v1x, v1y) + quad2(v2x, v2y, dC1, v1x, v1y));
        atomicAdd(&g1x[j2], 2.0 * (dC2[0] * v2x + dC2[1] * v2y));
        atomicAdd(&g1y[j2], 2.0 * (dC2[1] * v2x + dC2[2] * v2y));
        atomicAdd(&g1t[j2], 2.0 * quad2(v2x, v2y, dC2, v1x, v1y));
    }
    double tot[12];
    block_sum_f64<T, 12>(acc, red, tot);  // its barriers also complete the scatter into g1*
    double s1[6] = {0, 0, 0, 0, 0, 0};
    for (int j = t; j < n; j += T) {
        const double a = g1x[j], b = g1y[j], c = g1t[j];
        s1[0] += a * a; s1[1] += a * b; s1[2] += a * c; s1[3] += b * b; s1[4] += b * c; s1[5] += c * c;
    }
    double tot1[6];
    block_sum_f64<T, 6>(s1, red, tot1);
    const double H[3][3] = {{tot[0], tot[1], tot[3]}, {tot[1], tot[2], tot[4]}, {tot[3], tot[4], tot[5]}};
    double Hi[3][3];
    if (!inv3(H, Hi)) {
#pragma unroll
        for (int q = 0; q < 6; q++) cov[q] = nan("");
        return;
    }
    const double S[3][3] = {{tot[6] + tot1[0], tot[7] + tot1[1], tot[8] + tot1[2]},
                            {tot[7] + tot1[1], tot[9] + tot1[3], tot[10] + tot1[4]},
                            {tot[8] + tot1[2], tot[10] + tot1[4], tot[11] + tot1[5]}};
    double Tm[3][3];
#pragma unroll
    for (int a = 0; a < 3; a++)
#pragma unroll
        for (int b = 0; b < 3; b++) Tm[a][b] = Hi[a][0] * S[0][b] + Hi[a][1] * S[1][b] + Hi[a][2] * S[2][b];
    const double s2 = sigma * sigma;
    int q = 0;
#pragma unroll
    for (int a = 0; a < 3; a++)
#pragma unroll
        for (int b = a; b < 3; b++) cov[q++] = s2 * (Tm[a][0] * Hi[b][0] + Tm[a][1] * Hi[b][1] + Tm[a][2] * Hi[b][2]);
}

}  // namespace PL_NAMESPACE
