/*
 * oracle/plicp_oracle.h -- TEST INFRASTRUCTURE ONLY (parity oracle, CPU baseline).
 *
 * CPU restatement (plain C, double precision, single thread) of the PL-ICP path
 * that the reference reaches through CCanonicalMatcher::FMatch
 * (src/zhcsm/ZHCanonical_Matcher.cpp:229-285) -> CSM sm_icp().
 *
 * PARITY UNPINNED: CSM (AndreaCensi/csm, pinned by the reference only as the
 * directory name "csm-master", src/CMakeLists.txt:7-9) is NOT vendored under
 * /root/reference and cannot be built here.  The sm_icp arithmetic below restates
 * the published upstream algorithm (sm/csm/icp/{icp.c,icp_loop.c,icp_corr_tricks.c,
 * icp_corr_dumb.c,icp_outliers.c,icp_covariance.c}, sm/csm/laser_data.c,
 * sm/lib/gpc/gpc.c) from its public description; every parameter value, the
 * median filter, the unit conversion and the call protocol are taken from the
 * in-tree wrapper and ARE authoritative.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may link or call this file.  The product library never does.
 */
#ifndef PLICP_ORACLE_H
#define PLICP_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* CSM "struct laser_data" fields the path touches. */
typedef struct orc_laser_data {
    int nrays;
    double min_theta, max_theta;
    double *theta;      /* [nrays] */
    int *valid;         /* [nrays] */
    double *readings;   /* [nrays] metres, -1 when invalid (wrapper :203-204) */
    int *cluster;       /* [nrays] */
    double *px, *py;    /* cartesian, sensor frame */
    double *wx, *wy;    /* cartesian, reference ("world") frame */
    double *wrho, *wphi;
    int *corr_valid, *corr_j1, *corr_j2;
    double *corr_dist2_j1;
    int *up_bigger, *up_smaller, *down_bigger, *down_smaller;
    double odometry[3], estimate[3], true_pose[3];
} orc_laser_data;

/* CSM "struct sm_params": every field the wrapper sets
 * (src/zhcsm/ZHCanonical_Matcher.cpp:83-157,266-272). */
typedef struct orc_sm_params {
    orc_laser_data *laser_ref;
    orc_laser_data *laser_sens;
    double first_guess[3];
    double max_angular_correction_deg;
    double max_linear_correction;
    int max_iterations;
    double epsilon_xy;
    double epsilon_theta;
    double max_correspondence_dist;
    double sigma;
    int use_corr_tricks;        /* 1 tricks (reference), 0 dumb/possible_interval, 2 exhaustive (analysis only) */
    int restart;
    double restart_threshold_mean_error;
    double restart_dt;
    double restart_dtheta;
    double clustering_threshold;
    int orientation_neighbourhood;
    int use_point_to_line_distance;
    int do_alpha_test;
    double do_alpha_test_thresholdDeg;
    double outliers_maxPerc;
    double outliers_adaptive_order;
    double outliers_adaptive_mult;
    int do_visibility_test;
    int outliers_remove_doubles;
    int do_compute_covariance;
    int debug_verify_tricks;
    int use_ml_weights;
    int use_sigma_weights;
    double min_reading;
    double max_reading;
} orc_sm_params;

/* CSM "struct sm_result". cov_x_m is a plain row-major 3x3 here. */
typedef struct orc_sm_result {
    int valid;
    double x[3];
    int iterations;
    int nvalid;
    double error;
    double cov_x_m[9];
    unsigned int corr_hash;   /* ld_corr_hash of the final correspondences (not in CSM's struct) */
    int n_loops;              /* number of icp_loop runs (1 + restarts executed) */
} orc_sm_result;

orc_laser_data *orc_ld_alloc_new(int nrays);
void orc_ld_free(orc_laser_data *ld);

/* wrapper defaults: sm_init (ZHCanonical_Matcher.cpp:83-157); min/max range in cm like Base_PARAM */
void orc_sm_default_params(orc_sm_params *p, int pm_min_range_cm, int pm_max_range_cm);

void orc_sm_icp(orc_sm_params *params, orc_sm_result *res);

/* 5-tap in-place median filter (ZHCanonical_Matcher.cpp:159-189) */
void orc_median_filter(float *r, int n);

/* fromPMScan2LDP (ZHCanonical_Matcher.cpp:193-227); ranges in cm (float), thresholds in cm (int) */
orc_laser_data *orc_from_pmscan(const float *r_cm, int n, float pm_fi_min, float pm_dfi,
                                int pm_min_range_cm, int pm_max_range_cm);

/* CCanonicalMatcher::FMatch(PMScan*,PMScan*,bool) (ZHCanonical_Matcher.cpp:229-285) on already
 * median-filtered ranges.  Returns 0 ok / -1 failed; pose_out = {rx[cm], ry[cm], th[rad]} as floats
 * exactly like lsa->rx,ry,th; res receives the raw sm_result (metres).
 * j1_out/j2_out (may be NULL): final correspondences, -1 when invalid. */
int orc_zhcsm_fmatch(const float *ref_r_cm, const float *sens_r_cm, int n,
                     float pm_fi_min, float pm_dfi, int pm_min_range_cm, int pm_max_range_cm,
                     const orc_sm_params *params_or_null, const double *first_guess_or_null,
                     float pose_out[3], orc_sm_result *res, int *j1_out, int *j2_out);

/* Batch helper for the timed CPU baseline: runs pairs [0,npairs) with `nthreads` pthreads
 * (static partition).  scans: [nscans][n] float cm (already median filtered);
 * ref_idx/sens_idx index scans.  out: npairs results. */
void orc_zhcsm_fmatch_batch(const float *scans, int n, const int *ref_idx, const int *sens_idx,
                            int npairs, float pm_fi_min, float pm_dfi, int pm_min_range_cm,
                            int pm_max_range_cm, const orc_sm_params *params_or_null,
                            const double *first_guess_or_null /* [npairs][3] */,
                            orc_sm_result *out, int nthreads);

void orc_median_filter_batch(float *scans, int S, int n, int nthreads);

/* analysis helpers */
unsigned int orc_ld_corr_hash(const orc_laser_data *ld);
int orc_poly_greatest_real_root(const double a[5], double *root);

#ifdef __cplusplus
}
#endif
#endif
