/*
 * oracle/icp_oracle.c -- TEST INFRASTRUCTURE ONLY (parity oracle, CPU baseline).
 *
 * CPU restatement of the point-to-point ICP the reference reaches through CICPWarpper::ICPMatch
 * (src/zhicp/ZHIcp_Warpper.cpp:125-176) -> CICPWarpper::ICP (:57-123) -> mrpt::slam::CICP::Align with
 * ICP_algorithm = icpClassic and the options of initICP (:43-53).
 *
 * PARITY UNPINNED: MRPT 0.9.6 (pinned by the reference as the directory name mrpt-0.9.6, src/CMakeLists.txt:31-56)
 * is not vendored and cannot be built here.  The arithmetic restates the published algorithm of
 *   libs/slam/src/slam/CICP.cpp            ICP_Method_Classic: loop control, threshold schedule, covariance, quality
 *   libs/maps/src/maps/CPointsMap.cpp      computeMatchingWith2D: nearest neighbour + distance/angle gate +
 *                                          "only keep the closest" filter + correspondencesRatio
 *   libs/scanmatching/src/scan_matching.cpp leastSquareErrorRigidTransformation (closed-form 2-D rigid LSQ)
 * Every option value the wrapper sets is authoritative (in-tree); MRPT's remaining defaults are from its public
 * documentation and are parameters here (corresponding_points_decimation in particular: see icp_oracle.h).
 * A KD-tree and a brute-force search return the same nearest neighbour except for exact distance ties.
 */
#include "icp_oracle.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#define SQF(x) ((x) * (x))

void orc_icp_default_options(orc_icp_options *o) {
    memset(o, 0, sizeof(*o));
    /* CICPWarpper::initICP (src/zhicp/ZHIcp_Warpper.cpp:43-53) */
    o->maxIterations = 100;
    o->thresholdAng = (float)(10.0f * M_PI / 180.0); /* DEG2RAD(10.0f) */
    o->thresholdDist = 0.75f;
    o->ALFA = 0.5f;
    o->smallestThresholdDist = 0.05f;
    /* MRPT defaults [UPSTREAM] */
    o->onlyClosestCorrespondences = 1;
    o->minAbsStep_trans = 1e-6f;
    o->minAbsStep_rot = 1e-6f;
    o->corresponding_points_decimation = 5;
    o->rotate_decimation_offset = 2;
    o->kernel_rho = 0.07f;
    o->use_kernel = 1;
    o->Axy_aprox_derivatives = 0.05f;
}

static double wrap_to_pi(double a) {
    /* mrpt::math::wrapToPi: wrapTo2Pi(a + pi) - pi */
    a += M_PI;
    int was_neg = a < 0;
    a = fmod(a, 2.0 * M_PI);
    if (was_neg) a += 2.0 * M_PI;
    return a - M_PI;
}

typedef struct {
    int this_idx, other_idx;
    float this_x, this_y, other_x, other_y;
    float err2;
} pair_t;

/* CPointsMap::computeMatchingWith2D.  xl/yl: scratch [N2].  Returns the number of correspondences kept. */
static int compute_matching(const float *m1x, const float *m1y, int N1, const float *m2x, const float *m2y, int N2,
                            const double pose[3], float maxDist, float maxAng, int keep_closest, int decimation,
                            int offset, pair_t *tmp, pair_t *out, float *xl, float *yl, float *best, float *ratio) {
    *ratio = 0;
    if (!N1 || !N2) return 0;
    const double sin_phi = sin(pose[2]), cos_phi = cos(pose[2]);
    float lxmin = FLT_MAX, lxmax = -FLT_MAX, lymin = FLT_MAX, lymax = -FLT_MAX;
    for (int i = 0; i < N2; i++) {
        xl[i] = (float)(pose[0] + cos_phi * m2x[i] - sin_phi * m2y[i]);
        yl[i] = (float)(pose[1] + sin_phi * m2x[i] + cos_phi * m2y[i]);
        if (xl[i] < lxmin) lxmin = xl[i];
        if (xl[i] > lxmax) lxmax = xl[i];
        if (yl[i] < lymin) lymin = yl[i];
        if (yl[i] > lymax) lymax = yl[i];
    }
    float gxmin = FLT_MAX, gxmax = -FLT_MAX, gymin = FLT_MAX, gymax = -FLT_MAX;
    for (int j = 0; j < N1; j++) {
        if (m1x[j] < gxmin) gxmin = m1x[j];
        if (m1x[j] > gxmax) gxmax = m1x[j];
        if (m1y[j] < gymin) gymin = m1y[j];
        if (m1y[j] > gymax) gymax = m1y[j];
    }
    if (lxmin > gxmax || lxmax < gxmin || lymin > gymax || lymax < gymin) return 0;
    int nt = 0, nOther = 0;
    for (int i = offset; i < N2; i += decimation) {
        const float x = xl[i], y = yl[i];
        float bd = FLT_MAX;
        int bj = 0;
        for (int j = 0; j < N1; j++) {
            const float dx = m1x[j] - x, dy = m1y[j] - y;
            const float d = dx * dx + dy * dy;
            if (d < bd) { bd = d; bj = j; }
        }
        const double lim = maxAng * sqrt(SQF(pose[0] - x) + SQF(pose[1] - y)) + maxDist;
        if (bd < lim * lim) {
            pair_t *p = &tmp[nt++];
            p->this_idx = bj; p->other_idx = i;
            p->this_x = m1x[bj]; p->this_y = m1y[bj];
            p->other_x = m2x[i]; p->other_y = m2y[i];
            p->err2 = bd;
            nOther++;
        }
    }
    int nk = 0;
    if (keep_closest) {
        for (int j = 0; j < N1; j++) best[j] = FLT_MAX;
        for (int k = 0; k < nt; k++)
            if (tmp[k].err2 < best[tmp[k].this_idx]) best[tmp[k].this_idx] = tmp[k].err2;
        for (int k = 0; k < nt; k++)
            if (best[tmp[k].this_idx] == tmp[k].err2) out[nk++] = tmp[k];
    } else {
        memcpy(out, tmp, sizeof(pair_t) * nt);
        nk = nt;
    }
    *ratio = nOther / (float)(N2 / decimation);
    return nk;
}

/* scanmatching::leastSquareErrorRigidTransformation */
static int lsq_rigid(const pair_t *c, int N, double pose[3]) {
    if (N < 2) return 0;
    const float N_inv = 1.0f / N;
    float SumXa = 0, SumXb = 0, SumYa = 0, SumYb = 0, Sxx = 0, Sxy = 0, Syx = 0, Syy = 0;
    for (int k = 0; k < N; k++) {
        const float xa = c[k].this_x, ya = c[k].this_y, xb = c[k].other_x, yb = c[k].other_y;
        SumXa += xa; SumYa += ya; SumXb += xb; SumYb += yb;
        Sxx += xa * xb; Sxy += xa * yb; Syx += ya * xb; Syy += ya * yb;
    }
    const float mean_x_a = SumXa * N_inv, mean_y_a = SumYa * N_inv, mean_x_b = SumXb * N_inv, mean_y_b = SumYb * N_inv;
    const float Ax = N * (Sxx + Syy) - SumXa * SumXb - SumYa * SumYb;
    const float Ay = SumXa * SumYb + N * (Syx - Sxy) - SumXb * SumYa;
    pose[2] = (Ax != 0 || Ay != 0) ? atan2((double)Ay, (double)Ax) : 0;
    const double ccos = cos(pose[2]), csin = sin(pose[2]);
    pose[0] = mean_x_a - mean_x_b * ccos + mean_y_b * csin;
    pose[1] = mean_y_a - mean_x_b * csin - mean_y_b * ccos;
    return 1;
}

static double kernelf(const orc_icp_options *o, double x2, double rho2) { return o->use_kernel ? x2 / (x2 + rho2) : x2; }

static int inv3(const double A[9], double inv[9]) {
    const double det = A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) + A[2] * (A[3] * A[7] - A[4] * A[6]);
    if (det == 0 || isnan(det)) return 0;
    inv[0] = (A[4] * A[8] - A[5] * A[7]) / det; inv[1] = (A[2] * A[7] - A[1] * A[8]) / det; inv[2] = (A[1] * A[5] - A[2] * A[4]) / det;
    inv[3] = (A[5] * A[6] - A[3] * A[8]) / det; inv[4] = (A[0] * A[8] - A[2] * A[6]) / det; inv[5] = (A[2] * A[3] - A[0] * A[5]) / det;
    inv[6] = (A[3] * A[7] - A[4] * A[6]) / det; inv[7] = (A[1] * A[6] - A[0] * A[7]) / det; inv[8] = (A[0] * A[4] - A[1] * A[3]) / det;
    return 1;
}

/* three-point parabola through (w1,q1),(w2,q2),(w3,q3): derivative at x0 = 2*A*x0 + B */
static double parab_deriv(double w1, double q1, double w2, double q2, double w3, double q3, double x0) {
    const double A = ((q3 - q2) / ((w3 - w2) * (w3 - w1))) - ((q1 - q2) / ((w1 - w2) * (w3 - w1)));
    const double B = ((q1 - q2) + (A * ((w2 * w2) - (w1 * w1)))) / (w1 - w2);
    return (2 * A * x0) + B;
}

void orc_mrpt_icp(const float *m1x, const float *m1y, int N1, const float *m2x, const float *m2y, int N2,
                  const double init[3], const orc_icp_options *opt_or_null, orc_icp_result *res) {
    orc_icp_options od;
    if (!opt_or_null) { orc_icp_default_options(&od); opt_or_null = &od; }
    const orc_icp_options *o = opt_or_null;
    memset(res, 0, sizeof(*res));
    res->goodness = 1;
    double mean[3] = {init[0], init[1], init[2]}, last[3] = {0, 0, 0};
    float umbral_dist = o->thresholdDist, umbral_ang = o->thresholdAng, ratio = 0;
    int nCorr = 0, keep, offset = 0;
    const int dec = o->corresponding_points_decimation < 1 ? 1 : o->corresponding_points_decimation;
    pair_t *tmp = (pair_t *)malloc(sizeof(pair_t) * (N2 + 1)), *corr = (pair_t *)malloc(sizeof(pair_t) * (N2 + 1));
    float *xl = (float *)malloc(sizeof(float) * (N2 + 1)), *yl = (float *)malloc(sizeof(float) * (N2 + 1));
    float *best = (float *)malloc(sizeof(float) * (N1 + 1));
    if (N2 > 0) {
        do {
            nCorr = compute_matching(m1x, m1y, N1, m2x, m2y, N2, mean, umbral_dist, umbral_ang,
                                     o->onlyClosestCorrespondences, dec, offset, tmp, corr, xl, yl, best, &ratio);
            if (!nCorr) keep = 0;
            else {
                lsq_rigid(corr, nCorr, mean);
                keep = 1;
                if (!(fabs(last[0] - mean[0]) > o->minAbsStep_trans || fabs(last[1] - mean[1]) > o->minAbsStep_trans ||
                      fabs(wrap_to_pi(last[2] - mean[2])) > o->minAbsStep_rot)) {
                    umbral_dist *= o->ALFA;
                    umbral_ang *= o->ALFA;
                    if (umbral_dist < o->smallestThresholdDist) keep = 0;
                    if (o->rotate_decimation_offset == 2) offset = (offset + 1) % dec;
                }
                last[0] = mean[0]; last[1] = mean[1]; last[2] = mean[2];
            }
            res->nIterations++;
            if (o->rotate_decimation_offset == 1) offset = (offset + 1) % dec;
            if (res->nIterations >= o->maxIterations && umbral_dist > o->smallestThresholdDist) umbral_dist *= o->ALFA;
        } while ((keep && res->nIterations < o->maxIterations) ||
                 (res->nIterations >= o->maxIterations && umbral_dist > o->smallestThresholdDist));
        if (nCorr) {
            /* covariance: (D D^T + 1e-6 I)^-1 with D = numerical Jacobian of the (kernelised) squared residuals */
            const double ccos = cos(mean[2]), csin = sin(mean[2]);
            const double Axy = o->Axy_aprox_derivatives, rho2 = (double)o->kernel_rho * o->kernel_rho;
            double DDt[9] = {0};
            for (int k = 0; k < nCorr; k++) {
                const pair_t *c = &corr[k];
                const float oxt = (float)(mean[0] + ccos * c->other_x - csin * c->other_y);
                const float oyt = (float)(mean[1] + csin * c->other_x + ccos * c->other_y);
                double D[3], w1, w2, w3, q1, q2, q3;
                w1 = oxt - Axy; w2 = oxt; w3 = oxt + Axy;
                q1 = kernelf(o, SQF(c->this_x - w1) + SQF(c->this_y - oyt), rho2);
                q2 = kernelf(o, SQF(c->this_x - w2) + SQF(c->this_y - oyt), rho2);
                q3 = kernelf(o, SQF(c->this_x - w3) + SQF(c->this_y - oyt), rho2);
                D[0] = parab_deriv(w1, q1, w2, q2, w3, q3, oxt);
                w1 = oyt - Axy; w2 = oyt; w3 = oyt + Axy;
                q1 = kernelf(o, SQF(c->this_x - oxt) + SQF(c->this_y - w1), rho2);
                q2 = kernelf(o, SQF(c->this_x - oxt) + SQF(c->this_y - w2), rho2);
                q3 = kernelf(o, SQF(c->this_x - oxt) + SQF(c->this_y - w3), rho2);
                D[1] = parab_deriv(w1, q1, w2, q2, w3, q3, oyt);
                w1 = mean[2] - Axy; w2 = mean[2]; w3 = mean[2] + Axy;
                double q[3], w[3] = {w1, w2, w3};
                for (int a = 0; a < 3; a++) {
                    const double cx = mean[0] + cos(w[a]) * c->other_x - sin(w[a]) * c->other_y;
                    const double cy = mean[1] + sin(w[a]) * c->other_x + cos(w[a]) * c->other_y;
                    q[a] = kernelf(o, SQF(c->this_x - cx) + SQF(c->this_y - cy), rho2);
                }
                D[2] = parab_deriv(w1, q[0], w2, q[1], w3, q[2], mean[2]);
                for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) DDt[3 * a + b] += D[a] * D[b];
            }
            for (int a = 0; a < 3; a++) DDt[4 * a] += 1e-6;
            if (!inv3(DDt, res->cov)) for (int a = 0; a < 9; a++) res->cov[a] = NAN;
            res->goodness = ratio;
            /* quality: relative growth of the mean squared matching error under +-Axy shifts of the pose */
            {
                const float A2 = umbral_dist * 0.9f;
                double E[5];
                const double shifts[5][2] = {{0, 0}, {-A2, 0}, {A2, 0}, {0, -A2}, {0, A2}};
                for (int s = 0; s < 5; s++) {
                    double P[3] = {mean[0] + shifts[s][0], mean[1] + shifts[s][1], mean[2]};
                    float r2;
                    int nk = compute_matching(m1x, m1y, N1, m2x, m2y, N2, P, umbral_dist, umbral_ang,
                                              o->onlyClosestCorrespondences, dec, 0, tmp, corr, xl, yl, best, &r2);
                    double sum = 0;
                    for (int k = 0; k < nk; k++) sum += corr[k].err2;
                    E[s] = nk ? sum / nk : 0;
                }
                double m = E[1] - E[0];
                for (int s = 2; s < 5; s++) if (E[s] - E[0] > m) m = E[s] - E[0];
                res->quality = (float)(-m / (E[0] + 1e-1));
            }
        }
    }
    res->pose[0] = mean[0]; res->pose[1] = mean[1]; res->pose[2] = mean[2];
    res->nCorrespondences = nCorr;
    free(tmp); free(corr); free(xl); free(yl); free(best);
}

typedef struct {
    const float *pts; const int *off; const int *cnt; const int *ri; const int *ai; const double *init;
    const orc_icp_options *opt; orc_icp_result *out; int lo, hi;
} icp_job_t;

static void *icp_worker(void *a) {
    icp_job_t *j = (icp_job_t *)a;
    for (int k = j->lo; k < j->hi; k++) {
        const int r = j->ri[k], c = j->ai[k];
        const float *p1 = j->pts + 2 * (size_t)j->off[r], *p2 = j->pts + 2 * (size_t)j->off[c];
        /* cloud layout: x[0..N) then y[0..N) */
        orc_mrpt_icp(p1, p1 + j->cnt[r], j->cnt[r], p2, p2 + j->cnt[c], j->cnt[c], j->init + 3 * (size_t)k, j->opt, &j->out[k]);
    }
    return 0;
}

void orc_mrpt_icp_batch(const float *pts, const int *off, const int *cnt, const int *ref_idx, const int *cur_idx,
                        const double *init, int njobs, const orc_icp_options *opt_or_null, orc_icp_result *out,
                        int nthreads) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    icp_job_t jobs[256];
    for (int t = 0; t < nthreads; t++) {
        icp_job_t jb = {pts, off, cnt, ref_idx, cur_idx, init, opt_or_null, out, (int)((long long)njobs * t / nthreads),
                        (int)((long long)njobs * (t + 1) / nthreads)};
        jobs[t] = jb;
    }
    if (nthreads == 1) { icp_worker(&jobs[0]); return; }
    for (int t = 0; t < nthreads; t++) pthread_create(&th[t], 0, icp_worker, &jobs[t]);
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], 0);
}
