/*
 * oracle/plicp_oracle.c -- TEST INFRASTRUCTURE ONLY.  See plicp_oracle.h.
 *
 * PARITY UNPINNED (CSM is not vendored under /root/reference).  Each function cites
 * (a) the in-tree call site / parameter source it serves and (b) the upstream CSM file
 * whose published algorithm it restates.
 */
#include "plicp_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define SQ(x) ((x) * (x))

/* ------------------------------------------------------------------ laser_data */

/* ld_alloc_new: called at src/zhcsm/ZHCanonical_Matcher.cpp:196. [UPSTREAM sm/csm/laser_data.c] */
orc_laser_data *orc_ld_alloc_new(int n) {
    orc_laser_data *ld = (orc_laser_data *)calloc(1, sizeof(*ld));
    ld->nrays = n;
    ld->theta = (double *)malloc(sizeof(double) * n);
    ld->valid = (int *)malloc(sizeof(int) * n);
    ld->readings = (double *)malloc(sizeof(double) * n);
    ld->cluster = (int *)malloc(sizeof(int) * n);
    ld->px = (double *)malloc(sizeof(double) * n);
    ld->py = (double *)malloc(sizeof(double) * n);
    ld->wx = (double *)malloc(sizeof(double) * n);
    ld->wy = (double *)malloc(sizeof(double) * n);
    ld->wrho = (double *)malloc(sizeof(double) * n);
    ld->wphi = (double *)malloc(sizeof(double) * n);
    ld->corr_valid = (int *)malloc(sizeof(int) * n);
    ld->corr_j1 = (int *)malloc(sizeof(int) * n);
    ld->corr_j2 = (int *)malloc(sizeof(int) * n);
    ld->corr_dist2_j1 = (double *)malloc(sizeof(double) * n);
    ld->up_bigger = (int *)malloc(sizeof(int) * n);
    ld->up_smaller = (int *)malloc(sizeof(int) * n);
    ld->down_bigger = (int *)malloc(sizeof(int) * n);
    ld->down_smaller = (int *)malloc(sizeof(int) * n);
    for (int i = 0; i < n; i++) {
        ld->theta[i] = NAN;
        ld->valid[i] = 0;
        ld->readings[i] = NAN;
        ld->cluster[i] = -1;
        ld->px[i] = ld->py[i] = ld->wx[i] = ld->wy[i] = ld->wrho[i] = ld->wphi[i] = NAN;
        ld->corr_valid[i] = 0;
        ld->corr_j1[i] = ld->corr_j2[i] = -1;
        ld->corr_dist2_j1[i] = NAN;
    }
    ld->min_theta = ld->max_theta = NAN;
    return ld;
}

/* ld_free: called at ZHCanonical_Matcher.cpp:246-247 */
void orc_ld_free(orc_laser_data *ld) {
    if (!ld) return;
    free(ld->theta); free(ld->valid); free(ld->readings); free(ld->cluster);
    free(ld->px); free(ld->py); free(ld->wx); free(ld->wy); free(ld->wrho); free(ld->wphi);
    free(ld->corr_valid); free(ld->corr_j1); free(ld->corr_j2); free(ld->corr_dist2_j1);
    free(ld->up_bigger); free(ld->up_smaller); free(ld->down_bigger); free(ld->down_smaller);
    free(ld);
}

static int ld_valid_ray(const orc_laser_data *ld, int i) {
    return i >= 0 && i < ld->nrays && ld->valid[i];
}
static int ld_valid_corr(const orc_laser_data *ld, int i) { return ld->corr_valid[i]; }
static void ld_set_null_corr(orc_laser_data *ld, int i) {
    ld->corr_valid[i] = 0;
    ld->corr_j1[i] = -1;
    ld->corr_j2[i] = -1;
}
static int ld_next_valid_up(const orc_laser_data *ld, int i) {
    for (int j = i + 1; j < ld->nrays; j++) if (ld->valid[j]) return j;
    return -1;
}
static int ld_next_valid_down(const orc_laser_data *ld, int i) {
    for (int j = i - 1; j >= 0; j--) if (ld->valid[j]) return j;
    return -1;
}
static int ld_num_valid_corr(const orc_laser_data *ld) {
    int c = 0;
    for (int i = 0; i < ld->nrays; i++) c += ld->corr_valid[i] ? 1 : 0;
    return c;
}

/* ld_valid_fields [UPSTREAM laser_data.c]: sanity checks sm_icp runs first.  The wrapper relies on it
 * accepting readings = -1 on invalid rays (ZHCanonical_Matcher.cpp:203-204). */
static int ld_valid_fields(const orc_laser_data *ld) {
    if (!ld) return 0;
    if (ld->nrays < 10 || ld->nrays > 10000) return 0;
    if (isnan(ld->min_theta) || isnan(ld->max_theta)) return 0;
    double fov = ld->max_theta - ld->min_theta;
    if (fov < 20.0 * M_PI / 180.0 || fov > 2.01 * M_PI) return 0;
    if (fabs(ld->min_theta - ld->theta[0]) > 1e-8) return 0;
    if (fabs(ld->max_theta - ld->theta[ld->nrays - 1]) > 1e-8) return 0;
    int num_valid = 0, num_invalid = 0;
    for (int i = 0; i < ld->nrays; i++) {
        double th = ld->theta[i];
        if (ld->valid[i]) {
            double r = ld->readings[i];
            if (isnan(r) || isnan(th)) return 0;
            if (!(0.0 < r && r < 100.0)) return 0;
        } else {
            if (isnan(th)) return 0;
            if (ld->cluster[i] != -1) return 0;
        }
        if (ld->cluster[i] < -1) return 0;
        if (ld->valid[i] == 1) num_valid++;
        if (ld->valid[i] == 0) num_invalid++;
    }
    if (num_valid < ld->nrays * 0.10) return 0;
    if (num_valid + num_invalid != ld->nrays) return 0;
    return 1;
}

/* ld_invalid_if_outside [UPSTREAM laser_data.c]: SURVEY 8a-3 step (2) */
static void ld_invalid_if_outside(orc_laser_data *ld, double min_reading, double max_reading) {
    for (int i = 0; i < ld->nrays; i++) {
        if (!ld_valid_ray(ld, i)) continue;
        double r = ld->readings[i];
        if (r <= min_reading || r > max_reading) ld->valid[i] = 0;
    }
}

/* ld_create_jump_tables [UPSTREAM laser_data.c]: SURVEY 8a-3 step (3) */
static void ld_create_jump_tables(orc_laser_data *ld) {
    for (int i = 0; i < ld->nrays; i++) {
        int j;
        j = i + 1;
        while (j < ld->nrays && ld->valid[j] && ld->readings[j] <= ld->readings[i]) j++;
        ld->up_bigger[i] = j - i;
        j = i + 1;
        while (j < ld->nrays && ld->valid[j] && ld->readings[j] >= ld->readings[i]) j++;
        ld->up_smaller[i] = j - i;
        j = i - 1;
        while (j >= 0 && ld->valid[j] && ld->readings[j] >= ld->readings[i]) j--;
        ld->down_smaller[i] = j - i;
        j = i - 1;
        while (j >= 0 && ld->valid[j] && ld->readings[j] <= ld->readings[i]) j--;
        ld->down_bigger[i] = j - i;
    }
}

/* ld_compute_cartesian [UPSTREAM laser_data.c]: SURVEY 8a-3 step (4) */
static void ld_compute_cartesian(orc_laser_data *ld) {
    for (int i = 0; i < ld->nrays; i++) {
        if (!ld_valid_ray(ld, i)) continue;
        ld->px[i] = cos(ld->theta[i]) * ld->readings[i];
        ld->py[i] = sin(ld->theta[i]) * ld->readings[i];
    }
}

/* ld_compute_world_coords [UPSTREAM laser_data.c]: SURVEY 8a-4 */
static void ld_compute_world_coords(orc_laser_data *ld, const double *pose) {
    double cos_theta = cos(pose[2]);
    double sin_theta = sin(pose[2]);
    for (int i = 0; i < ld->nrays; i++) {
        if (!ld_valid_ray(ld, i)) continue;
        double x0 = ld->px[i], y0 = ld->py[i];
        ld->wx[i] = cos_theta * x0 - sin_theta * y0 + pose[0];
        ld->wy[i] = sin_theta * x0 + cos_theta * y0 + pose[1];
        double x = ld->wx[i], y = ld->wy[i];
        ld->wrho[i] = sqrt(x * x + y * y);
        ld->wphi[i] = atan2(y, x);
    }
}

/* ld_corr_hash [UPSTREAM laser_data.c]: SURVEY 8a-8 */
unsigned int orc_ld_corr_hash(const orc_laser_data *ld) {
    unsigned int hash = 0;
    for (unsigned int i = 0; i < (unsigned)ld->nrays; i++) {
        int str = ld_valid_corr(ld, (int)i) ? (ld->corr_j1[i] + 1000 * ld->corr_j2[i]) : -1;
        hash ^= ((i & 1) == 0) ? ((hash << 7) ^ (unsigned)str ^ (hash >> 3))
                               : (~((hash << 11) ^ (unsigned)str ^ (hash >> 5)));
    }
    return hash & 0x7FFFFFFF;
}

/* ------------------------------------------------------------------ math utils */

static double dist2_d(double ax, double ay, double bx, double by) {
    double dx = ax - bx, dy = ay - by;
    return dx * dx + dy * dy;
}

/* dist_to_segment_d [UPSTREAM sm/csm/math_utils.c]: SURVEY 8a-6 */
static double dist_to_segment(double ax, double ay, double bx, double by, double x, double y) {
    double t0 = ax - bx;
    double t1 = ay - by;
    double one_on_r = 1 / sqrt(t0 * t0 + t1 * t1);
    double nx = t1 * one_on_r;
    double ny = -t0 * one_on_r;
    double c = nx, s = ny;
    double rho = c * ax + s * ay;
    double projx = c * rho + s * s * x - c * s * y;
    double projy = s * rho - c * s * x + c * c * y;
    double distance = fabs(rho - (c * x + s * y));
    if ((projx - ax) * (projx - bx) + (projy - ay) * (projy - by) < 0) {
        return distance; /* projection inside the segment */
    }
    double da = dist2_d(ax, ay, x, y), db = dist2_d(bx, by, x, y);
    return sqrt(da < db ? da : db);
}

/* pose_diff_d = oplus(ominus(pose1), pose2), angle wrapped [UPSTREAM math_utils.c]: SURVEY 8a-8 */
static void pose_diff(const double pose2[3], const double pose1[3], double res[3]) {
    double c = cos(pose1[2]), s = sin(pose1[2]);
    double t[3] = {-c * pose1[0] - s * pose1[1], s * pose1[0] - c * pose1[1], -pose1[2]};
    double c2 = cos(t[2]), s2 = sin(t[2]);
    res[0] = t[0] + c2 * pose2[0] - s2 * pose2[1];
    res[1] = t[1] + s2 * pose2[0] + c2 * pose2[1];
    res[2] = t[2] + pose2[2];
    while (res[2] > +M_PI) res[2] -= 2 * M_PI;
    while (res[2] < -M_PI) res[2] += 2 * M_PI;
}

/* ------------------------------------------------------------------ correspondences */

static void finish_corr(const orc_sm_params *params, orc_laser_data *ref, orc_laser_data *sens,
                        int i, int j1, double best_dist) {
    (void)params;
    /* We ignore matching the first or the last point in the scan */
    if (0 == j1 || j1 == ref->nrays - 1) { ld_set_null_corr(sens, i); return; }
    int j2;
    int j2up = ld_next_valid_up(ref, j1);
    int j2down = ld_next_valid_down(ref, j1);
    if (j2up == -1 && j2down == -1) { ld_set_null_corr(sens, i); return; }
    if (j2up == -1) j2 = j2down;
    else if (j2down == -1) j2 = j2up;
    else {
        double dist_up = dist2_d(sens->wx[i], sens->wy[i], ref->px[j2up], ref->py[j2up]);
        double dist_down = dist2_d(sens->wx[i], sens->wy[i], ref->px[j2down], ref->py[j2down]);
        j2 = dist_up < dist_down ? j2up : j2down;
    }
    sens->corr_valid[i] = 1;
    sens->corr_j1[i] = j1;
    sens->corr_j2[i] = j2;
    sens->corr_dist2_j1[i] = best_dist;
}

/* find_correspondences_tricks [UPSTREAM sm/csm/icp/icp_corr_tricks.c]: SURVEY 8a-4.
 * Selected by use_corr_tricks=1 (ZHCanonical_Matcher.cpp:106). */
/* analysis counters (single-thread runs only): distance evaluations, rays searched, and the sum over groups of 32
 * consecutive rays of the slowest ray's evaluations -- the cost model of a 32-lane lock-step walk */
static __thread long long g_walk_visits, g_walk_rays, g_walk_warpmax, g_walk_calls; /* thread-local: no sharing between the batch workers */
void orc_debug_walk_stats(long long out[4], int reset) {
    out[0] = g_walk_visits; out[1] = g_walk_rays; out[2] = g_walk_warpmax; out[3] = g_walk_calls;
    if (reset) g_walk_visits = g_walk_rays = g_walk_warpmax = g_walk_calls = 0;
}

static void find_correspondences_tricks(orc_sm_params *params) {
    long long visits_this = 0, group_max = 0;
    g_walk_calls++;
    orc_laser_data *laser_ref = params->laser_ref;
    orc_laser_data *laser_sens = params->laser_sens;
    double C1 = (double)laser_ref->nrays / (laser_ref->max_theta - laser_ref->min_theta);
    double max_correspondence_dist2 = SQ(params->max_correspondence_dist);
    int last_best = -1;
    for (int i = 0; i < laser_sens->nrays; i++) {
        if ((i & 31) == 0) { g_walk_warpmax += group_max; group_max = 0; }
        if (visits_this > group_max) group_max = visits_this;
        visits_this = 0;
        if (!ld_valid_ray(laser_sens, i)) { ld_set_null_corr(laser_sens, i); continue; }
        g_walk_rays++;
        double pwx = laser_sens->wx[i], pwy = laser_sens->wy[i];
        double p_i_w_nrm2 = laser_sens->wrho[i];
        double p_i_w_phi = laser_sens->wphi[i];
        int from = 0;
        int to = laser_ref->nrays - 1;
        int start_cell = (int)((p_i_w_phi - laser_ref->min_theta) * C1);
        int j1 = -1;
        double best_dist = 42;
        int we_start_at = (last_best == -1) ? start_cell : last_best + 1;
        if (we_start_at > to) we_start_at = to;
        if (we_start_at < from) we_start_at = from;
        int up = we_start_at + 1;
        int down = we_start_at;
        double last_dist_up = 0; /* first is down */
        double last_dist_down = -1;
        int up_stopped = 0, down_stopped = 0;
        while (!up_stopped || !down_stopped) {
            int now_up = up_stopped ? 0 : down_stopped ? 1 : (last_dist_up < last_dist_down);
            if (now_up) {
                if (up > to) { up_stopped = 1; continue; }
                if (!laser_ref->valid[up]) { ++up; continue; }
                last_dist_up = dist2_d(pwx, pwy, laser_ref->px[up], laser_ref->py[up]);
                g_walk_visits++; visits_this++;
                if (last_dist_up < best_dist || j1 == -1) { j1 = up; best_dist = last_dist_up; }
                if (up > start_cell) {
                    double delta_theta = laser_ref->theta[up] - p_i_w_phi;
                    double min_dist_up = p_i_w_nrm2 * ((delta_theta > M_PI * 0.5) ? 1 : sin(delta_theta));
                    if (SQ(min_dist_up) > best_dist) { up_stopped = 1; continue; }
                    up += (laser_ref->readings[up] < p_i_w_nrm2) ? laser_ref->up_bigger[up]
                                                                  : laser_ref->up_smaller[up];
                } else {
                    ++up;
                }
            }
            if (!now_up) {
                if (down < from) { down_stopped = 1; continue; }
                if (!laser_ref->valid[down]) { --down; continue; }
                last_dist_down = dist2_d(pwx, pwy, laser_ref->px[down], laser_ref->py[down]);
                g_walk_visits++; visits_this++;
                if (last_dist_down < best_dist || j1 == -1) { j1 = down; best_dist = last_dist_down; }
                if (down < start_cell) {
                    double delta_theta = p_i_w_phi - laser_ref->theta[down];
                    double min_dist_down = p_i_w_nrm2 * ((delta_theta > M_PI * 0.5) ? 1 : sin(delta_theta));
                    if (SQ(min_dist_down) > best_dist) { down_stopped = 1; continue; }
                    down += (laser_ref->readings[down] < p_i_w_nrm2) ? laser_ref->down_bigger[down]
                                                                      : laser_ref->down_smaller[down];
                } else {
                    --down;
                }
            }
        }
        if (-1 == j1 || best_dist > max_correspondence_dist2) { ld_set_null_corr(laser_sens, i); continue; }
        if (0 == j1 || j1 == laser_ref->nrays - 1) { ld_set_null_corr(laser_sens, i); continue; }
        finish_corr(params, laser_ref, laser_sens, i, j1, best_dist);
        if (laser_sens->corr_valid[i]) last_best = j1;
    }
}

/* find_correspondences + possible_interval [UPSTREAM sm/csm/icp/icp_corr_dumb.c]: SURVEY 8a-4 (naive twin).
 * mode 0: restricted to possible_interval (upstream "dumb").  mode 2: all rays (analysis only). */
static void find_correspondences_dumb(orc_sm_params *params, int exhaustive) {
    orc_laser_data *ref = params->laser_ref;
    orc_laser_data *sens = params->laser_sens;
    double maxd2 = SQ(params->max_correspondence_dist);
    for (int i = 0; i < sens->nrays; i++) {
        if (!ld_valid_ray(sens, i)) { ld_set_null_corr(sens, i); continue; }
        double pwx = sens->wx[i], pwy = sens->wy[i];
        int from = 0, to = ref->nrays - 1;
        if (!exhaustive) {
            double angle_res = (ref->max_theta - ref->min_theta) / ref->nrays;
            double delta = fabs(params->max_angular_correction_deg * M_PI / 180.0) +
                           fabs(atan(params->max_linear_correction / sqrt(pwx * pwx + pwy * pwy)));
            int range = (int)ceil(delta / angle_res);
            double start_theta = atan2(pwy, pwx);
            if (start_theta < ref->min_theta) start_theta += 2 * M_PI;
            if (start_theta > ref->max_theta) start_theta -= 2 * M_PI;
            int start_cell = (int)((start_theta - ref->min_theta) / (ref->max_theta - ref->min_theta) * ref->nrays);
            from = start_cell - range; if (from < 0) from = 0; if (from > ref->nrays - 1) from = ref->nrays - 1;
            to = start_cell + range;   if (to < 0) to = 0;     if (to > ref->nrays - 1) to = ref->nrays - 1;
        }
        int j1 = -1;
        double best_dist = 10000;
        for (int j = from; j <= to; j++) {
            if (!ld_valid_ray(ref, j)) continue;
            double dist = dist2_d(pwx, pwy, ref->px[j], ref->py[j]);
            if (dist > maxd2) continue;
            if (-1 == j1 || dist < best_dist) { j1 = j; best_dist = dist; }
        }
        if (j1 == -1) { ld_set_null_corr(sens, i); continue; }
        finish_corr(params, ref, sens, i, j1, best_dist);
    }
}

/* ------------------------------------------------------------------ outliers */

/* kill_outliers_double [UPSTREAM sm/csm/icp/icp_outliers.c]: SURVEY 8a-5 (outliers_remove_doubles=1, :139) */
static void kill_outliers_double(orc_sm_params *params) {
    double threshold = 3;
    orc_laser_data *ref = params->laser_ref, *sens = params->laser_sens;
    double *dist2_i = (double *)malloc(sizeof(double) * sens->nrays);
    double *dist2_j = (double *)malloc(sizeof(double) * ref->nrays);
    for (int j = 0; j < ref->nrays; j++) dist2_j[j] = 1000000;
    for (int i = 0; i < sens->nrays; i++) {
        if (!ld_valid_corr(sens, i)) continue;
        int j1 = sens->corr_j1[i];
        dist2_i[i] = sens->corr_dist2_j1[i];
        dist2_j[j1] = dist2_j[j1] < dist2_i[i] ? dist2_j[j1] : dist2_i[i];
    }
    for (int i = 0; i < sens->nrays; i++) {
        if (!ld_valid_corr(sens, i)) continue;
        int j1 = sens->corr_j1[i];
        if (dist2_i[i] > (threshold * threshold) * dist2_j[j1]) sens->corr_valid[i] = 0;
    }
    free(dist2_i);
    free(dist2_j);
}

static int cmp_double(const void *a, const void *b) {
    double x = *(const double *)a, y = *(const double *)b;
    return (x > y) - (x < y);
}

/* kill_outliers_trim [UPSTREAM icp_outliers.c]: SURVEY 8a-6 (maxPerc .95, adaptive .7 x 2.0, :118-130) */
static void kill_outliers_trim(orc_sm_params *params, double *total_error) {
    orc_laser_data *ref = params->laser_ref, *sens = params->laser_sens;
    int n = sens->nrays;
    double *dist2 = (double *)malloc(sizeof(double) * n);
    double *dist = (double *)malloc(sizeof(double) * n);
    int k = 0;
    for (int i = 0; i < n; i++) {
        if (!ld_valid_corr(sens, i)) { dist[i] = NAN; continue; }
        int j1 = sens->corr_j1[i], j2 = sens->corr_j2[i];
        dist[i] = dist_to_segment(ref->px[j1], ref->py[j1], ref->px[j2], ref->py[j2], sens->wx[i], sens->wy[i]);
        dist2[k] = dist[i];
        k++;
    }
    int order = (int)floor(k * params->outliers_maxPerc);
    order = order < k - 1 ? order : k - 1; order = order > 0 ? order : 0;
    qsort(dist2, k, sizeof(double), cmp_double);
    double error_limit1 = k > 0 ? dist2[order] : 0;
    int order2 = (int)floor(k * params->outliers_adaptive_order);
    order2 = order2 < k - 1 ? order2 : k - 1; order2 = order2 > 0 ? order2 : 0;
    double error_limit2 = k > 0 ? params->outliers_adaptive_mult * dist2[order2] : 0;
    double error_limit = error_limit1 < error_limit2 ? error_limit1 : error_limit2;
    *total_error = 0;
    for (int i = 0; i < n; i++) {
        if (!ld_valid_corr(sens, i)) continue;
        if (dist[i] > error_limit) {
            sens->corr_valid[i] = 0;
            sens->corr_j1[i] = -1;
            sens->corr_j2[i] = -1;
        } else {
            *total_error += dist[i];
        }
    }
    free(dist2);
    free(dist);
}

/* ------------------------------------------------------------------ gpc */

/* Greatest real root of a[0]+a[1]x+..+a[4]x^4.  Upstream uses gsl_poly_complex_solve (companion-matrix
 * eigenvalues) and keeps the largest root whose imaginary part is exactly 0 [UPSTREAM lib/gpc/gpc.c].
 * Restated as a bracketing solver: real critical points of the monic quartic split the axis into
 * monotone pieces; the right-most piece with a sign change is bisected + Newton-polished.  Parity is on
 * the root's value to double rounding, not on GSL's bits. */
static double poly_eval(const double *c, int deg, double x) {
    double v = c[deg];
    for (int i = deg - 1; i >= 0; i--) v = v * x + c[i];
    return v;
}

/* root of monotone f on [lo,hi] with f(lo)<=0<=f(hi) (increasing) or reversed: safeguarded Newton */
static double root_bracketed(const double *c, int deg, double lo, double hi) {
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

/* all real roots (ascending) of a polynomial of degree deg<=4 with c[deg]!=0, via its derivative's roots */
static int real_roots(const double *c, int deg, double *roots) {
    if (deg == 1) { roots[0] = -c[0] / c[1]; return 1; }
    double dc[5], crit[4];
    for (int i = 1; i <= deg; i++) dc[i - 1] = c[i] * i;
    int nc = real_roots(dc, deg - 1, crit);
    /* Cauchy bound */
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

int orc_poly_greatest_real_root(const double a[5], double *root) {
    if (a[4] == 0) return 0;
    double roots[4];
    int nr = real_roots(a, 4, roots);
    if (nr == 0) return 0;
    *root = roots[nr - 1];
    return 1;
}

static int inv2(const double A[2][2], double inv[2][2]) {
    double det = A[0][0] * A[1][1] - A[0][1] * A[1][0];
    if (det == 0 || isnan(det)) return 0;
    inv[0][0] = A[1][1] / det;  inv[0][1] = -A[0][1] / det;
    inv[1][0] = -A[1][0] / det; inv[1][1] = A[0][0] / det;
    return 1;
}

/* solve A x = b, 4x4, Gaussian elimination with partial pivoting (stands for m_inv + m_mult) */
static int solve4(double A[4][4], double b[4], double x[4]) {
    int n = 4;
    for (int k = 0; k < n; k++) {
        int piv = k;
        for (int r = k + 1; r < n; r++) if (fabs(A[r][k]) > fabs(A[piv][k])) piv = r;
        if (A[piv][k] == 0 || isnan(A[piv][k])) return 0;
        if (piv != k) {
            for (int c2 = 0; c2 < n; c2++) { double t = A[k][c2]; A[k][c2] = A[piv][c2]; A[piv][c2] = t; }
            double t = b[k]; b[k] = b[piv]; b[piv] = t;
        }
        for (int r = k + 1; r < n; r++) {
            double f = A[r][k] / A[k][k];
            for (int c2 = k; c2 < n; c2++) A[r][c2] -= f * A[k][c2];
            b[r] -= f * b[k];
        }
    }
    for (int k = n - 1; k >= 0; k--) {
        double s = b[k];
        for (int c2 = k + 1; c2 < n; c2++) s -= A[k][c2] * x[c2];
        x[k] = s / A[k][k];
    }
    return 1;
}

/* gpc_solve given the accumulated sums d_bigM, d_g [UPSTREAM sm/lib/gpc/gpc.c]: SURVEY 8a-7.
 * Censi, "An ICP variant using a point-to-line metric", ICRA 2008, eq. (31). */
static int gpc_solve_sums(double d_bigM[4][4], double d_g[4], double x_out[3]) {
    double bigM[4][4], g[4];
    for (int a = 0; a < 4; a++) g[a] = -2 * d_g[a];
    for (int a = 0; a < 4; a++) for (int b = 0; b < 4; b++) bigM[a][b] = 2 * d_bigM[a][b];

    double mA[2][2] = {{bigM[0][0], bigM[0][1]}, {bigM[1][0], bigM[1][1]}};
    double mB[2][2] = {{bigM[0][2], bigM[0][3]}, {bigM[1][2], bigM[1][3]}};
    double mD[2][2] = {{bigM[2][2], bigM[2][3]}, {bigM[3][2], bigM[3][3]}};
    double mAi[2][2];
    if (!inv2(mA, mAi)) return 0;
    /* mS = mD - mB' * inv(A) * mB */
    double AiB[2][2], mS[2][2];
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) AiB[i][j] = mAi[i][0] * mB[0][j] + mAi[i][1] * mB[1][j];
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++)
        mS[i][j] = mD[i][j] + (-1.0) * (mB[0][i] * AiB[0][j] + mB[1][i] * AiB[1][j]);
    /* mSa = inv(mS) * det(mS) */
    double mSi[2][2], mSa[2][2];
    if (!inv2(mS, mSi)) return 0;
    double detS = mS[0][0] * mS[1][1] - mS[0][1] * mS[1][0];
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) mSa[i][j] = mSi[i][j] * detS;

    double g1[2] = {g[0], g[1]}, g2[2] = {g[2], g[3]};
    /* m1t = g1t*mAi*mB ; m2t = m1t*mSa ; m3t = g2t*mSa */
    double t12[2] = {g1[0] * mAi[0][0] + g1[1] * mAi[1][0], g1[0] * mAi[0][1] + g1[1] * mAi[1][1]};
    double m1[2] = {t12[0] * mB[0][0] + t12[1] * mB[1][0], t12[0] * mB[0][1] + t12[1] * mB[1][1]};
    double m2[2] = {m1[0] * mSa[0][0] + m1[1] * mSa[1][0], m1[0] * mSa[0][1] + m1[1] * mSa[1][1]};
    double m3[2] = {g2[0] * mSa[0][0] + g2[1] * mSa[1][0], g2[0] * mSa[0][1] + g2[1] * mSa[1][1]};
#define DOT(a, b) ((a)[0] * (b)[0] + (a)[1] * (b)[1])
    double p[3] = {DOT(m2, m2) - 2 * DOT(m2, m3) + DOT(m3, m3),
                   4 * DOT(m2, m1) - 8 * DOT(m2, g2) + 4 * DOT(g2, m3),
                   4 * DOT(m1, m1) - 8 * DOT(m1, g2) + 4 * DOT(g2, g2)};
#undef DOT
    double l[3] = {detS, 2 * mS[0][0] + 2 * mS[1][1], 4};
    double q[5] = {p[0] - (l[0] * l[0]), p[1] - (2 * l[1] * l[0]), p[2] - (l[1] * l[1] + 2 * l[0] * l[2]),
                   -(2 * l[2] * l[1]), -(l[2] * l[2])};
    double lambda;
    if (!orc_poly_greatest_real_root(q, &lambda)) return 0;
    /* x = -(bigM + 2 lambda W)^-1 g, W = diag(0,0,1,1) */
    double A[4][4], b[4], x[4];
    for (int i = 0; i < 4; i++) { for (int j = 0; j < 4; j++) A[i][j] = bigM[i][j]; b[i] = g[i]; }
    A[2][2] += 2 * lambda;
    A[3][3] += 2 * lambda;
    if (!solve4(A, b, x)) return 0;
    for (int i = 0; i < 4; i++) x[i] = -x[i];
    x_out[0] = x[0];
    x_out[1] = x[1];
    x_out[2] = atan2(x[3], x[2]);
    if (isnan(x_out[0]) || isnan(x_out[1]) || isnan(x_out[2])) return 0;
    return 1;
}

/* compute_next_estimate [UPSTREAM sm/csm/icp/icp_loop.c] + the accumulation half of gpc_solve: SURVEY 8a-7 */
static int compute_next_estimate(orc_sm_params *params, const double x_old[3], double x_new[3]) {
    (void)x_old;
    orc_laser_data *ref = params->laser_ref, *sens = params->laser_sens;
    double M[4][4] = {{0}}, g[4] = {0};
    for (int i = 0; i < sens->nrays; i++) {
        if (!sens->valid[i]) continue;
        if (!ld_valid_corr(sens, i)) continue;
        int j1 = sens->corr_j1[i], j2 = sens->corr_j2[i];
        double px = sens->px[i], py = sens->py[i];
        double qx = ref->px[j1], qy = ref->py[j1];
        double diff0 = ref->px[j1] - ref->px[j2];
        double diff1 = ref->py[j1] - ref->py[j2];
        double one_on_norm = 1 / sqrt(diff0 * diff0 + diff1 * diff1);
        double cos_alpha = +diff1 * one_on_norm;
        double sin_alpha = -diff0 * one_on_norm;
        double C00 = cos_alpha * cos_alpha, C01 = cos_alpha * sin_alpha, C10 = C01, C11 = sin_alpha * sin_alpha;

        M[0][0] += C00;
        M[0][1] += C01;
        M[0][2] += +px * C00 + py * C01;
        M[0][3] += -py * C00 + px * C01;
        M[1][0] += C10;
        M[1][1] += C11;
        M[1][2] += +px * C10 + py * C11;
        M[1][3] += +px * C11 - py * C10;
        M[2][0] += px * C00 + py * C10;
        M[2][1] += px * C01 + py * C11;
        M[2][2] += (px * px) * (+C00) + (px * py) * (+C10 + C01) + (py * py) * (+C11);
        M[2][3] += (px * px) * (+C01) + (px * py) * (-C00 + C11) + (py * py) * (-C10);
        M[3][0] += -py * C00 + px * C10;
        M[3][1] += -py * C01 + px * C11;
        M[3][2] += (px * px) * (+C10) + (px * py) * (-C00 + C11) + (py * py) * (-C01);
        M[3][3] += (px * px) * (+C11) + (px * py) * (-C10 - C01) + (py * py) * (+C00);
        g[0] += C00 * qx + C10 * qy;
        g[1] += C01 * qx + C11 * qy;
        g[2] += qx * (C00 * px + C01 * py) + qy * (C10 * px + C11 * py);
        g[3] += qx * (C00 * (-py) + C01 * px) + qy * (C10 * (-py) + C11 * px);
    }
    return gpc_solve_sums(M, g, x_new);
}

/* ------------------------------------------------------------------ icp_loop / sm_icp */

/* icp_loop [UPSTREAM sm/csm/icp/icp_loop.c]: SURVEY 8a-8 */
static int icp_loop(orc_sm_params *params, const double *q0, double *x_new, double *total_error, int *valid,
                    int *iterations) {
    if (isnan(q0[0]) || isnan(q0[1]) || isnan(q0[2])) return 0;
    orc_laser_data *laser_sens = params->laser_sens;
    double x_old[3] = {q0[0], q0[1], q0[2]}, delta[3];
    unsigned int *hashes = (unsigned int *)malloc(sizeof(unsigned int) * (params->max_iterations + 1));
    int iteration;
    int all_is_okay = 1;
    for (iteration = 0; iteration < params->max_iterations; iteration++) {
        ld_compute_world_coords(laser_sens, x_old);
        if (params->use_corr_tricks == 1) find_correspondences_tricks(params);
        else find_correspondences_dumb(params, params->use_corr_tricks == 2);

        int num_corr = ld_num_valid_corr(laser_sens);
        double fail_perc = 0.05;
        if (num_corr < fail_perc * laser_sens->nrays) { all_is_okay = 0; break; }
        if (params->outliers_remove_doubles) kill_outliers_double(params);
        double error = 0;
        kill_outliers_trim(params, &error);
        int num_corr_after = ld_num_valid_corr(laser_sens);
        *total_error = error;
        *valid = num_corr_after;
        if (num_corr_after < fail_perc * laser_sens->nrays) { all_is_okay = 0; break; }
        if (!compute_next_estimate(params, x_old, x_new)) { all_is_okay = 0; break; }
        pose_diff(x_new, x_old, delta);
        hashes[iteration] = orc_ld_corr_hash(laser_sens);
        if (params->use_point_to_line_distance) {
            int loop_detected = 0;
            for (int a = iteration - 1; a >= 0; a--) {
                if (hashes[a] == hashes[iteration]) { loop_detected = 1; break; }
            }
            if (loop_detected) break;
        }
        {
            double a = sqrt(delta[0] * delta[0] + delta[1] * delta[1]);
            double b = fabs(delta[2]);
            if (a < params->epsilon_xy && b < params->epsilon_theta) break;
        }
        x_old[0] = x_new[0]; x_old[1] = x_new[1]; x_old[2] = x_new[2];
    }
    free(hashes);
    *iterations = iteration + 1;
    return all_is_okay;
}

/* compute_covariance_exact [UPSTREAM sm/csm/icp/icp_covariance.c]: SURVEY 8a-8, Censi ICRA'07.
 * Reached when setCov(true) (ZHCanonical_Matcher.cpp:79-81); read back at :292-296. */
static void compute_C_k(double j1x, double j1y, double j2x, double j2y, double C[2][2]) {
    double alpha = M_PI / 2 + atan2(j1y - j2y, j1x - j2x);
    double c = cos(alpha), s = sin(alpha);
    C[0][0] = c * c; C[0][1] = c * s; C[1][0] = c * s; C[1][1] = s * s;
}
static void dC_drho(double p1x, double p1y, double p2x, double p2y, double dC[2][2]) {
    double eps = 0.001;
    double C_k[2][2], C_eps[2][2];
    compute_C_k(p1x, p1y, p2x, p2y, C_k);
    double nrm = sqrt(p1x * p1x + p1y * p1y);
    double p1bx = p1x + (eps / nrm) * p1x, p1by = p1y + (eps / nrm) * p1y;
    compute_C_k(p1bx, p1by, p2x, p2y, C_eps);
    for (int a = 0; a < 2; a++) for (int b = 0; b < 2; b++) dC[a][b] = (1 / eps) * (C_eps[a][b] - C_k[a][b]);
}
static double quad2(const double u[2], const double C[2][2], const double v[2]) {
    return u[0] * (C[0][0] * v[0] + C[0][1] * v[1]) + u[1] * (C[1][0] * v[0] + C[1][1] * v[1]);
}
static int inv3(const double A[3][3], double inv[3][3]) {
    double det = A[0][0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) - A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
                 A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
    if (det == 0 || isnan(det)) return 0;
    inv[0][0] = (A[1][1] * A[2][2] - A[1][2] * A[2][1]) / det;
    inv[0][1] = (A[0][2] * A[2][1] - A[0][1] * A[2][2]) / det;
    inv[0][2] = (A[0][1] * A[1][2] - A[0][2] * A[1][1]) / det;
    inv[1][0] = (A[1][2] * A[2][0] - A[1][0] * A[2][2]) / det;
    inv[1][1] = (A[0][0] * A[2][2] - A[0][2] * A[2][0]) / det;
    inv[1][2] = (A[0][2] * A[1][0] - A[0][0] * A[1][2]) / det;
    inv[2][0] = (A[1][0] * A[2][1] - A[1][1] * A[2][0]) / det;
    inv[2][1] = (A[0][1] * A[2][0] - A[0][0] * A[2][1]) / det;
    inv[2][2] = (A[0][0] * A[1][1] - A[0][1] * A[1][0]) / det;
    return 1;
}
static int compute_covariance_exact(orc_laser_data *ref, orc_laser_data *sens, const double x[3], double cov0[9]) {
    int nr = ref->nrays, ns = sens->nrays;
    double *dy1 = (double *)calloc(3 * nr, sizeof(double)); /* d2J_dxdy1 [3][nr] */
    double *dy2 = (double *)calloc(3 * ns, sizeof(double)); /* d2J_dxdy2 [3][ns] */
    double d2J_dt2[2][2] = {{0}}, d2J_dt_dtheta[2] = {0}, d2J_dtheta2 = 0;
    double theta = x[2];
    double t[2] = {x[0], x[1]};
    for (int i = 0; i < ns; i++) {
        if (!ld_valid_corr(sens, i)) continue;
        int j1 = sens->corr_j1[i], j2 = sens->corr_j2[i];
        double pi_[2] = {sens->px[i], sens->py[i]};
        double pj1[2] = {ref->px[j1], ref->py[j1]};
        double pj2[2] = {ref->px[j2], ref->py[j2]};
        double a1 = theta + M_PI / 2;
        double v1[2] = {cos(a1) * pi_[0] - sin(a1) * pi_[1], sin(a1) * pi_[0] + cos(a1) * pi_[1]};
        double v2[2] = {cos(theta) * pi_[0] - sin(theta) * pi_[1] + t[0] - pj1[0],
                        sin(theta) * pi_[0] + cos(theta) * pi_[1] + t[1] - pj1[1]};
        double v3[2] = {cos(theta + sens->theta[i]), sin(theta + sens->theta[i])};
        double v4[2] = {cos(theta + sens->theta[i] + M_PI / 2), sin(theta + sens->theta[i] + M_PI / 2)};
        double C_k[2][2];
        compute_C_k(pj1[0], pj1[1], pj2[0], pj2[1], C_k);
        double a2 = theta + M_PI;
        double v_new[2] = {cos(a2) * pi_[0] - sin(a2) * pi_[1], sin(a2) * pi_[0] + cos(a2) * pi_[1]};
        for (int a = 0; a < 2; a++) for (int b = 0; b < 2; b++) d2J_dt2[a][b] += 2.0 * C_k[a][b];
        for (int a = 0; a < 2; a++) d2J_dt_dtheta[a] += 2.0 * (C_k[a][0] * v1[0] + C_k[a][1] * v1[1]);
        d2J_dtheta2 += 2.0 * (quad2(v2, C_k, v_new) + quad2(v1, C_k, v1));
        /* measurement rho_i in the second scan */
        dy2[0 * ns + i] += 2.0 * (C_k[0][0] * v3[0] + C_k[0][1] * v3[1]);
        dy2[1 * ns + i] += 2.0 * (C_k[1][0] * v3[0] + C_k[1][1] * v3[1]);
        dy2[2 * ns + i] += 2.0 * (quad2(v2, C_k, v4) + quad2(v3, C_k, v1));
        /* measurements rho_j1, rho_j2 in the first scan */
        double dC1[2][2], dC2[2][2];
        dC_drho(pj1[0], pj1[1], pj2[0], pj2[1], dC1);
        dC_drho(pj2[0], pj2[1], pj1[0], pj1[1], dC2);
        double v_j1[2] = {cos(ref->theta[j1]), sin(ref->theta[j1])};
        dy1[0 * nr + j1] += -2.0 * (C_k[0][0] * v_j1[0] + C_k[0][1] * v_j1[1]) + 2.0 * (dC1[0][0] * v2[0] + dC1[0][1] * v2[1]);
        dy1[1 * nr + j1] += -2.0 * (C_k[1][0] * v_j1[0] + C_k[1][1] * v_j1[1]) + 2.0 * (dC1[1][0] * v2[0] + dC1[1][1] * v2[1]);
        dy1[2 * nr + j1] += -2.0 * quad2(v_j1, C_k, v1) + quad2(v2, dC1, v1);
        dy1[0 * nr + j2] += 2.0 * (dC2[0][0] * v2[0] + dC2[0][1] * v2[1]);
        dy1[1 * nr + j2] += 2.0 * (dC2[1][0] * v2[0] + dC2[1][1] * v2[1]);
        dy1[2 * nr + j2] += 2.0 * quad2(v2, dC2, v1);
    }
    double H[3][3] = {{d2J_dt2[0][0], d2J_dt2[0][1], d2J_dt_dtheta[0]},
                      {d2J_dt2[1][0], d2J_dt2[1][1], d2J_dt_dtheta[1]},
                      {d2J_dt_dtheta[0], d2J_dt_dtheta[1], d2J_dtheta2}};
    double Hi[3][3];
    int ok = inv3(H, Hi);
    if (ok) {
        /* cov0 = (Hi*dy1)(Hi*dy1)' + (Hi*dy2)(Hi*dy2)' = Hi (dy1 dy1' + dy2 dy2') Hi' */
        double S[3][3] = {{0}};
        for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) {
            double s = 0;
            for (int k = 0; k < nr; k++) s += dy1[a * nr + k] * dy1[b * nr + k];
            for (int k = 0; k < ns; k++) s += dy2[a * ns + k] * dy2[b * ns + k];
            S[a][b] = s;
        }
        double T[3][3];
        for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) T[a][b] = Hi[a][0] * S[0][b] + Hi[a][1] * S[1][b] + Hi[a][2] * S[2][b];
        for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++)
            cov0[a * 3 + b] = T[a][0] * Hi[b][0] + T[a][1] * Hi[b][1] + T[a][2] * Hi[b][2];
    } else {
        for (int a = 0; a < 9; a++) cov0[a] = NAN;
    }
    free(dy1);
    free(dy2);
    return ok;
}

/* sm_icp [UPSTREAM sm/csm/icp/icp.c]; call sites src/zhcsm/ZHCanonical_Matcher.cpp:274,344: SURVEY 8a-3 */
void orc_sm_icp(orc_sm_params *params, orc_sm_result *res) {
    memset(res, 0, sizeof(*res));
    res->valid = 0;
    orc_laser_data *laser_ref = params->laser_ref;
    orc_laser_data *laser_sens = params->laser_sens;
    if (!ld_valid_fields(laser_ref) || !ld_valid_fields(laser_sens)) return;
    ld_invalid_if_outside(laser_ref, params->min_reading, params->max_reading);
    ld_invalid_if_outside(laser_sens, params->min_reading, params->max_reading);
    if (params->use_corr_tricks == 1 || params->debug_verify_tricks) ld_create_jump_tables(laser_ref);
    ld_compute_cartesian(laser_ref);
    ld_compute_cartesian(laser_sens);

    double x_new[3] = {0, 0, 0};
    double x_old[3] = {params->first_guess[0], params->first_guess[1], params->first_guess[2]};
    double error = 0;
    int iterations = 0, nvalid = 0;
    res->n_loops = 1;
    if (!icp_loop(params, x_old, x_new, &error, &nvalid, &iterations)) {
        res->valid = 0;
        res->iterations = iterations;
        res->nvalid = 0;
        return;
    }
    int restarted = 0;
    double best_error = error;
    double best_x[3] = {x_new[0], x_new[1], x_new[2]};
    if (params->restart && (error / nvalid) > params->restart_threshold_mean_error) {
        restarted = 1;
        double dt = params->restart_dt;
        double dth = params->restart_dtheta;
        double perturb[6][3] = {{dt, 0, 0}, {-dt, 0, 0}, {0, dt, 0}, {0, -dt, 0}, {0, 0, dth}, {0, 0, -dth}};
        for (int a = 0; a < 6; a++) {
            double start[3] = {x_new[0] + perturb[a][0], x_new[1] + perturb[a][1], x_new[2] + perturb[a][2]};
            double x_a[3], my_error;
            int my_valid, my_iterations;
            res->n_loops++;
            if (!icp_loop(params, start, x_a, &my_error, &my_valid, &my_iterations)) break;
            iterations += my_iterations;
            if (my_error < best_error) {
                best_x[0] = x_a[0]; best_x[1] = x_a[1]; best_x[2] = x_a[2];
                best_error = my_error;
            }
        }
    }
    res->valid = 1;
    res->x[0] = best_x[0]; res->x[1] = best_x[1]; res->x[2] = best_x[2];
    if (restarted) {
        ld_compute_world_coords(laser_sens, res->x);
        if (params->use_corr_tricks == 1) find_correspondences_tricks(params);
        else find_correspondences_dumb(params, params->use_corr_tricks == 2);
    }
    if (params->do_compute_covariance) {
        double cov0[9];
        compute_covariance_exact(laser_ref, laser_sens, best_x, cov0);
        for (int a = 0; a < 9; a++) res->cov_x_m[a] = SQ(params->sigma) * cov0[a];
    }
    res->error = best_error;
    res->iterations = iterations;
    res->nvalid = nvalid;
    res->corr_hash = orc_ld_corr_hash(laser_sens);
}

/* ------------------------------------------------------------------ zhcsm wrapper */

/* CCanonicalMatcher::sm_init (src/zhcsm/ZHCanonical_Matcher.cpp:83-157) */
void orc_sm_default_params(orc_sm_params *p, int pm_min_range_cm, int pm_max_range_cm) {
    memset(p, 0, sizeof(*p));
    p->min_reading = pm_min_range_cm / 100.;
    p->max_reading = pm_max_range_cm / 100.;
    p->max_angular_correction_deg = 45;
    p->max_linear_correction = 0.5;
    p->max_iterations = 10;
    p->epsilon_xy = 0.001;
    p->epsilon_theta = 0.001;
    p->max_correspondence_dist = 2;
    p->sigma = 0.04;
    p->use_corr_tricks = 1;
    p->restart = 1;
    p->restart_threshold_mean_error = 0.01;
    p->restart_dt = 0.01;
    p->restart_dtheta = 0.0261799;
    p->clustering_threshold = 0.05;
    p->orientation_neighbourhood = 10;
    p->use_point_to_line_distance = 1;
    p->do_alpha_test = 0;
    p->do_alpha_test_thresholdDeg = 20.0;
    p->outliers_maxPerc = 0.95;
    p->outliers_adaptive_order = 0.7;
    p->outliers_adaptive_mult = 2.0;
    p->do_visibility_test = 0;
    p->outliers_remove_doubles = 1;
    p->do_compute_covariance = 0;
    p->debug_verify_tricks = 0;
    p->use_ml_weights = 0;
    p->use_sigma_weights = 0;
}

/* CCanonicalMatcher::median_filter (src/zhcsm/ZHCanonical_Matcher.cpp:159-189): in place, so taps
 * i-2,i-1 already hold filtered values. */
void orc_median_filter(float *r, int n) {
    float w[5];
    for (int i = 0; i < n; i++) {
        int k = 0;
        for (int j = i - 2; j <= i + 2; j++) {
            int l = j >= 0 ? j : 0;
            w[k++] = r[l < n ? l : n - 1];
        }
        for (int j = 4; j > 0; j--)
            for (k = 0; k < j; k++)
                if (w[k] > w[k + 1]) { float t = w[k]; w[k] = w[k + 1]; w[k + 1] = t; }
        r[i] = w[2];
    }
}

/* CCanonicalMatcher::fromPMScan2LDP (src/zhcsm/ZHCanonical_Matcher.cpp:193-227) */
orc_laser_data *orc_from_pmscan(const float *r_cm, int n, float pm_fi_min, float pm_dfi, int pm_min_range_cm,
                                int pm_max_range_cm) {
    orc_laser_data *ld = orc_ld_alloc_new(n);
    for (int i = 0; i < n; i++) {
        double r = r_cm[i];
        if (r <= pm_min_range_cm || r >= pm_max_range_cm) {
            ld->valid[i] = 0;
            ld->readings[i] = -1;
        } else {
            ld->valid[i] = 1;
            ld->readings[i] = r / 100.0;
        }
        /* float + float*unsigned: evaluated in float (pm_dfi*i converts i to float), then widened */
        ld->theta[i] = pm_fi_min + pm_dfi * (float)(unsigned int)i;
        ld->cluster[i] = -1;
    }
    ld->min_theta = ld->theta[0];
    ld->max_theta = ld->theta[n - 1];
    return ld;
}

int orc_zhcsm_fmatch(const float *ref_r_cm, const float *sens_r_cm, int n, float pm_fi_min, float pm_dfi,
                     int pm_min_range_cm, int pm_max_range_cm, const orc_sm_params *params_or_null,
                     const double *first_guess_or_null, float pose_out[3], orc_sm_result *res, int *j1_out,
                     int *j2_out) {
    orc_sm_params p;
    if (params_or_null) p = *params_or_null;
    else orc_sm_default_params(&p, pm_min_range_cm, pm_max_range_cm);
    orc_laser_data *ldpref = orc_from_pmscan(ref_r_cm, n, pm_fi_min, pm_dfi, pm_min_range_cm, pm_max_range_cm);
    orc_laser_data *ldpcur = orc_from_pmscan(sens_r_cm, n, pm_fi_min, pm_dfi, pm_min_range_cm, pm_max_range_cm);
    p.laser_ref = ldpref;
    p.laser_sens = ldpcur;
    for (int k = 0; k < 3; k++) p.first_guess[k] = first_guess_or_null ? first_guess_or_null[k] : 0.0;
    orc_sm_icp(&p, res);
    int ret = 0;
    if (res->valid) {
        for (int k = 0; k < 3; k++) ldpcur->true_pose[k] = res->x[k];
    } else {
        ret = -1;
    }
    /* lsa->rx = true_pose[0]*100. etc (:242-244): double product narrowed to float */
    pose_out[0] = (float)(ldpcur->true_pose[0] * 100.);
    pose_out[1] = (float)(ldpcur->true_pose[1] * 100.);
    pose_out[2] = (float)(ldpcur->true_pose[2]);
    if (j1_out && j2_out) {
        for (int i = 0; i < n; i++) {
            j1_out[i] = (res->valid && ldpcur->corr_valid[i]) ? ldpcur->corr_j1[i] : -1;
            j2_out[i] = (res->valid && ldpcur->corr_valid[i]) ? ldpcur->corr_j2[i] : -1;
        }
    }
    orc_ld_free(ldpref);
    orc_ld_free(ldpcur);
    return ret;
}

typedef struct {
    const float *scans; int n; const int *ref_idx, *sens_idx; int lo, hi;
    float fi_min, dfi; int minr, maxr; const orc_sm_params *params; const double *fg; orc_sm_result *out;
} batch_job;

static void *batch_worker(void *arg) {
    batch_job *j = (batch_job *)arg;
    float pose[3];
    for (int k = j->lo; k < j->hi; k++) {
        orc_zhcsm_fmatch(j->scans + (size_t)j->ref_idx[k] * j->n, j->scans + (size_t)j->sens_idx[k] * j->n, j->n,
                         j->fi_min, j->dfi, j->minr, j->maxr, j->params, j->fg ? j->fg + 3 * (size_t)k : NULL, pose,
                         &j->out[k], NULL, NULL);
    }
    return NULL;
}

void orc_zhcsm_fmatch_batch(const float *scans, int n, const int *ref_idx, const int *sens_idx, int npairs,
                            float pm_fi_min, float pm_dfi, int pm_min_range_cm, int pm_max_range_cm,
                            const orc_sm_params *params_or_null, const double *first_guess_or_null,
                            orc_sm_result *out, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    batch_job jobs[256];
    for (int t = 0; t < nthreads; t++) {
        batch_job jb = {scans, n, ref_idx, sens_idx, (int)((long long)npairs * t / nthreads),
                        (int)((long long)npairs * (t + 1) / nthreads), pm_fi_min, pm_dfi, pm_min_range_cm,
                        pm_max_range_cm, params_or_null, first_guess_or_null, out};
        jobs[t] = jb;
    }
    if (nthreads == 1) { batch_worker(&jobs[0]); return; }
    for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, batch_worker, &jobs[t]);
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
}

/* median filter over S scans in place (wrapper :233-237 filters every scan once), nthreads pthreads */
typedef struct { float *scans; int n, lo, hi; } mf_job;
static void *mf_worker(void *arg) {
    mf_job *j = (mf_job *)arg;
    for (int k = j->lo; k < j->hi; k++) orc_median_filter(j->scans + (size_t)k * j->n, j->n);
    return NULL;
}
void orc_median_filter_batch(float *scans, int S, int n, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    mf_job jobs[256];
    for (int t = 0; t < nthreads; t++) {
        mf_job jb = {scans, n, (int)((long long)S * t / nthreads), (int)((long long)S * (t + 1) / nthreads)};
        jobs[t] = jb;
    }
    if (nthreads == 1) { mf_worker(&jobs[0]); return; }
    for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, mf_worker, &jobs[t]);
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
}

/* Checker for the CUDA kernel's division-free reading conversion (laser_slam_b200/csrc/plicp2_common.cuh,
 * sens_point): fromPMScan2LDP computes (double)f / 100.0 (src/zhcsm/ZHCanonical_Matcher.cpp:198-208); the kernel
 * forms q0 = f * 0.01, r = fma(-q0, 100, f), q = fma(r, 0.01, q0).  Counts the floats with bit patterns in
 * [lo_bits, hi_bits) for which the two differ (expected: none). */
long long orc_count_div100_mismatches(unsigned lo_bits, unsigned hi_bits) {
    long long bad = 0;
    for (unsigned u = lo_bits; u < hi_bits; u++) {
        float f;
        memcpy(&f, &u, 4);
        const double r = (double)f;
        const double ref = r / 100.0;
        const double q0 = r * 0.01;
        const double q = fma(fma(-q0, 100.0, r), 0.01, q0);
        bad += !(q == ref);
    }
    return bad;
}
