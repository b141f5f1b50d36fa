/*
 * oracle/graph_oracle.c -- TEST INFRASTRUCTURE ONLY (parity oracle for SURVEY 8f-1).
 *
 * CPU restatement of the loop-closure candidate gating and consensus arithmetic of src/mapGraph/MapGraph.cpp, the code
 * that produces the candidate batches for the ICP matcher and consumes its results:
 *   isOverlappedArea :957-994   isLoopyArea :996-1032   isAvaliableArea :918-944
 *   deleteMisMatch :169-209 (+ statis :48-61, disPoeses :45-47)   calculateWeightedPose :146-167
 *   calculateFinalPose :110-123 (the choice of the representative; its g2o edge bookkeeping is out of scope)
 * MapGraph.cpp itself needs g2o/Eigen/Qt and cannot be compiled here; OrientedPoint2D::oplus (src/zhpsm/point.cpp:63-69),
 * which these functions call, IS compiled from the reference (oracle/_ref) and tests pin this file's copy against it.
 * float / double are used exactly where the reference declares them.
 */
#include "graph_oracle.h"

#include <math.h>
#include <stdlib.h>

static void oplus_point(const double p[3], double x, double y, double out[2]) { /* point.cpp:63-69 */
    const double c = cos(p[2]), s = sin(p[2]);
    out[0] = p[0] + x * c - y * s;
    out[1] = p[1] + x * s + y * c;
}

static void cal_bounding(double ld[2], double ru[2]) { /* MapGraph.cpp:946-955: through float */
    const float x_min = ld[0] < ru[0] ? ld[0] : ru[0], x_max = ld[0] > ru[0] ? ld[0] : ru[0];
    const float y_min = ld[1] < ru[1] ? ld[1] : ru[1], y_max = ld[1] > ru[1] ? ld[1] : ru[1];
    ld[0] = x_min; ld[1] = y_min; ru[0] = x_max; ru[1] = y_max;
}

static int is_overlapped(const double *pc, const double *ldc, const double *ruc, const double *pr, const double *ldr,
                         const double *rur) {
    double ld_cur[2], ru_cur[2], ld_ref[2], ru_ref[2];
    oplus_point(pc, ldc[0], ldc[1], ld_cur); oplus_point(pc, ruc[0], ruc[1], ru_cur);
    oplus_point(pr, ldr[0], ldr[1], ld_ref); oplus_point(pr, rur[0], rur[1], ru_ref);
    cal_bounding(ld_cur, ru_cur);
    cal_bounding(ld_ref, ru_ref);
    if (ru_cur[0] <= ld_cur[0] || ru_cur[1] <= ld_cur[1] || ru_ref[0] <= ld_ref[0] || ru_ref[1] <= ld_ref[1]) return 0;
    if (ru_cur[0] <= ld_ref[0] || ru_cur[1] <= ld_ref[1] || ru_ref[0] <= ld_cur[0] || ru_ref[1] <= ld_cur[1]) return 0;
    const float x_min = ld_cur[0] > ld_ref[0] ? ld_cur[0] : ld_ref[0], y_min = ld_cur[1] > ld_ref[1] ? ld_cur[1] : ld_ref[1];
    const float x_max = ru_cur[0] < ru_ref[0] ? ru_cur[0] : ru_ref[0], y_max = ru_cur[1] < ru_ref[1] ? ru_cur[1] : ru_ref[1];
    const float overlapped = (x_max - x_min) * (y_max - y_min);
    const float curArea = (ru_cur[0] - ld_cur[0]) * (ru_cur[1] - ld_cur[1]);
    const float ratio = overlapped / curArea;
    if (overlapped <= 0) return 0;
    return ratio >= 0.4;
}

static float cov_radius2(const double *cov_xy, float constant, float beta) {
    if (!cov_xy) return constant;
    const float dis_x = cov_xy[0], dis_y = cov_xy[1];
    return (fabsf(dis_x) + fabsf(dis_y)) * beta;
}

static int is_loopy(const double *pc, const double *pr, const double *cov_cur, float beta) {
    const float dis = cov_radius2(cov_cur, 2 * 2, beta);
    const float cur_dis = (pc[0] - pr[0]) * (pc[0] - pr[0]) + (pc[1] - pr[1]) * (pc[1] - pr[1]);
    return dis >= cur_dis;
}

void orc_gate_candidates(const double *pose, const double *ld, const double *ru, const double *cov_xy, int N,
                         const int32_t *cur_idx, int M, float beta, uint8_t *flags) {
    for (int m = 0; m < M; m++) {
        const int c = cur_idx[m];
        for (int j = 0; j < N; j++) {
            int f = 0;
            if (is_overlapped(pose + 3 * c, ld + 2 * c, ru + 2 * c, pose + 3 * j, ld + 2 * j, ru + 2 * j)) f |= ORC_GATE_OVERLAPPED;
            if (is_loopy(pose + 3 * c, pose + 3 * j, cov_xy ? cov_xy + 2 * c : 0, beta)) f |= ORC_GATE_LOOPY;
            flags[(size_t)m * N + j] = (uint8_t)f;
        }
    }
}

void orc_available_area(const double *trans, const double *cov_xy, int M, uint8_t *ok) {
    const float beta = 25, angle_thre = M_PI / 3.0;
    for (int m = 0; m < M; m++) {
        const double *t = trans + 3 * m;
        if (fabs(t[2]) >= angle_thre) { ok[m] = 0; continue; }
        const float dis = cov_radius2(cov_xy ? cov_xy + 2 * m : 0, 2.5 * 2.5, beta);
        const float cur_dis = t[0] * t[0] + t[1] * t[1];
        ok[m] = dis >= cur_dis;
    }
}

static float dis_poses(const double *a, const double *b) { /* :45-47 */
    return sqrt((a[0] - b[0]) * (a[0] - b[0]) + (a[1] - b[1]) * (a[1] - b[1]));
}

void orc_consensus(const double *poses, const float *good, const int32_t *set_off, int S, uint8_t *keep,
                   double *final_pose, int32_t *final_idx) {
    for (int s = 0; s < S; s++) {
        const int lo = set_off[s], n = set_off[s + 1] - lo;
        for (int i = 0; i < n; i++) keep[lo + i] = 1;
        final_idx[s] = -1;
        final_pose[3 * s] = final_pose[3 * s + 1] = final_pose[3 * s + 2] = 0;
        if (n <= 0) continue;
        /* deleteMisMatch: mean distance of each pose to the others; drop those above mean + 1.731 * "sigma"
         * (statis returns the variance, not its root; fewer than 3 entries: nothing is dropped) */
        if (n >= 3) {
            float *dis = (float *)malloc(sizeof(float) * n);
            for (int i = 0; i < n; i++) {
                float mean_dis = 0;
                for (int j = 0; j < n; j++) {
                    if (j == i) continue;
                    mean_dis += dis_poses(poses + 3 * (lo + i), poses + 3 * (lo + j));
                }
                dis[i] = mean_dis / (n - 1);
            }
            float sum = 0;
            for (int i = 0; i < n; i++) sum += dis[i];
            const float mean = sum / n;
            float sq = 0;
            for (int i = 0; i < n; i++) sq += (dis[i] - mean) * (dis[i] - mean);
            const float sigma = sq / (n - 1);
            const float scale = 1.731;
            for (int i = 0; i < n; i++)
                if (dis[i] > mean + scale * sigma) keep[lo + i] = 0;
            free(dis);
        }
        /* calculateWeightedPose over what is left */
        float total = 0;
        for (int i = 0; i < n; i++) if (keep[lo + i]) total += good[lo + i];
        float px = 0, py = 0, pth = 0;
        for (int i = 0; i < n; i++) {
            if (!keep[lo + i]) continue;
            const float w = (float)good[lo + i] / (float)total;
            px += w * poses[3 * (lo + i)];
            py += w * poses[3 * (lo + i) + 1];
            pth += w * poses[3 * (lo + i) + 2];
        }
        final_pose[3 * s] = px; final_pose[3 * s + 1] = py; final_pose[3 * s + 2] = pth;
        /* calculateFinalPose: the kept entry closest to the weighted pose */
        float min_dis = 1000000;
        for (int i = 0; i < n; i++) {
            if (!keep[lo + i]) continue;
            const float d = dis_poses(final_pose + 3 * s, poses + 3 * (lo + i));
            if (d < min_dis) { min_dis = d; final_idx[s] = lo + i; }
        }
    }
}

/* Eigen/src/LU/InverseImpl.h, compute_inverse<MatrixType, ResultType, 3>: cofactors of column 0, det = their dot
 * product with column 0 (redux order a0 + (a1 + a2)), inverse(i,j) = cofactor(j,i) / det. */
static double cof3(const double m[9], int i, int j) {
    int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return m[3 * i1 + j1] * m[3 * i2 + j2] - m[3 * i1 + j2] * m[3 * i2 + j1];
}
void orc_cov_to_information(const double *cov6, int M, double scale, const double *const_diag, double *info) {
    for (int k = 0; k < M; k++) {
        const double *c = cov6 + 6 * (size_t)k;
        double *o = info + 9 * (size_t)k;
        if (const_diag) {
            for (int q = 0; q < 9; q++) o[q] = 0.0;
            o[0] = const_diag[0]; o[4] = const_diag[1]; o[8] = const_diag[2];
            continue;
        }
        double m[9] = {c[0] * scale, c[1], c[2], c[1], c[3] * scale, c[4], c[2], c[4], c[5]};
        double c00 = cof3(m, 0, 0), c10 = cof3(m, 1, 0), c20 = cof3(m, 2, 0);
        double det = c00 * m[0] + (c10 * m[3] + c20 * m[6]);
        double invdet = 1.0 / det;
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) o[3 * i + j] = cof3(m, j, i) * invdet;
    }
}
