/*
 * include/csm_compat.h -- the CSM-shaped C entry of libzhsm.so.
 *
 * The reference reaches PL-ICP through one C function of the Canonical Scan Matcher and three helpers
 * (src/zhcsm/ZHCanonical_Matcher.cpp): `sm_icp(struct sm_params*, struct sm_result*)` (:274, :344),
 * `LDP ld_alloc_new(int)` (:196), `void ld_free(LDP)` (:246-247, :367) and `gsl_matrix_get(res->cov_x_m, i, j)`
 * (:292-296), with the headers <csm/csm_all.h>, <csm/algos.h>, <csm/logging.h> (:10-12).  This header declares
 * exactly those names with every field the wrapper sets or reads, and include/csm/ holds three
 * forwarding headers of the same names, so the reference's ZHCanonical_Matcher.cpp compiles UNMODIFIED against
 * -I<repo>/include and links against libzhsm.so instead of libcsm (option C).
 *
 * sm_icp here is a batch of one over a process-wide default engine handle (device $ZHSM_DEVICE, default 0); it is
 * serialised by a mutex, like the reference's single-threaded use.  The batched entry for throughput is
 * zhsm_plicp_batch (zhsm.h).  There is no CPU matcher behind it: without a usable sm_100 GPU sm_icp reports
 * valid = 0 and says why on stderr.
 *
 * What the engine accepts (everything the reference's only producer, fromPMScan2LDP :193-227, makes):
 *   - laser_ref and laser_sens have the same nrays and the same theta[] (one sensor);
 *   - theta[i] == (double)(fi_min + dfi * (float)i) for some floats fi_min, dfi (float arithmetic, as :210);
 *   - valid readings are float-centimetre values: readings[i] == (double)f / 100.0 for a float f (as :207);
 *   - use_corr_tricks = 1, use_point_to_line_distance = 1, do_alpha_test = do_visibility_test = use_ml_weights =
 *     use_sigma_weights = 0 (the wrapper's settings, :83-157).
 * Anything else: valid = 0 and a one-line reason on stderr -- never a silently different algorithm.
 */
#ifndef ZHSM_CSM_COMPAT_H
#define ZHSM_CSM_COMPAT_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- the slice of GSL the wrapper touches: a dense row-major matrix and its element read ------------------- */
#ifndef __GSL_MATRIX_DOUBLE_H__
typedef struct {
    size_t size1, size2; /* rows, columns */
    size_t tda;          /* row stride in elements */
    double *data;
    void *block;         /* unused (GSL's owner block) */
    int owner;
} gsl_matrix;
static inline double gsl_matrix_get(const gsl_matrix *m, size_t i, size_t j) { return m->data[i * m->tda + j]; }
static inline void gsl_matrix_set(gsl_matrix *m, size_t i, size_t j, double v) { m->data[i * m->tda + j] = v; }
#endif
gsl_matrix *zhsm_gsl_matrix_alloc(size_t rows, size_t cols); /* zero-filled */
void zhsm_gsl_matrix_free(gsl_matrix *m);

/* ---- struct laser_data ----------------------------------------------------------------------- */
typedef struct {
    double p[2]; /* Cartesian */
    double rho, phi;
} point2d;

enum correspondence_type { corr_pp = 0, corr_pl = 1 };
struct correspondence {
    int valid;
    int j1, j2; /* closest ref point, second closest */
    enum correspondence_type type;
    double dist2_j1;
};

struct laser_data {
    int nrays;
    double min_theta, max_theta;
    double *theta;    /* [nrays] beam angles, rad */
    int *valid;       /* [nrays] 1 = usable reading; sm_icp clears rays outside [min_reading, max_reading] */
    double *readings; /* [nrays] metres; the wrapper writes -1 on invalid rays (:204) */
    int *cluster;     /* [nrays] -1 = none (:211) */
    double *alpha, *cov_alpha; /* [nrays] orientation estimates: not used with the wrapper's settings */
    int *alpha_valid;
    double *readings_sigma, *true_alpha;
    struct correspondence *corr; /* [nrays] OUT (laser_sens): final correspondences */
    double true_pose[3], odometry[3], estimate[3];
    point2d *points;   /* [nrays] OUT: Cartesian points, sensor frame */
    point2d *points_w; /* [nrays] OUT (laser_sens): points in the ref frame at the result pose */
    int *up_bigger, *up_smaller, *down_bigger, *down_smaller; /* [nrays] jump tables: not exported by the engine */
};
typedef struct laser_data *LDP;

LDP ld_alloc_new(int nrays); /* all arrays allocated; valid = 0, readings/theta = NaN, cluster = -1, corr invalid */
void ld_free(LDP ld);

/* ---- struct sm_params: every field CCanonicalMatcher::sm_init sets (:83-157) ---------------- */
struct sm_params {
    LDP laser_ref, laser_sens;
    double first_guess[3]; /* m, m, rad (:270-272) */
    double max_angular_correction_deg, max_linear_correction;
    int max_iterations;
    double epsilon_xy, epsilon_theta;
    double max_correspondence_dist;
    double sigma;
    int use_corr_tricks;
    int restart;
    double restart_threshold_mean_error, restart_dt, restart_dtheta;
    double clustering_threshold;
    int orientation_neighbourhood;
    int use_point_to_line_distance;
    int do_alpha_test;
    double do_alpha_test_thresholdDeg;
    double outliers_maxPerc, outliers_adaptive_order, outliers_adaptive_mult;
    int do_visibility_test;
    int outliers_remove_doubles;
    int do_compute_covariance;
    int debug_verify_tricks;
    int use_ml_weights, use_sigma_weights;
    double laser[3]; /* pose of the sensor on the robot: never initialised by the wrapper, never read here */
    double min_reading, max_reading; /* m */
};

/* ---- struct sm_result ------------------------------------------------------------------------ */
struct sm_result {
    int valid;
    double x[3]; /* m, m, rad */
    int iterations;
    int nvalid;
    double error;
    gsl_matrix *cov_x_m;  /* 3x3, allocated by sm_icp when do_compute_covariance (the caller owns it, as with CSM) */
    gsl_matrix *dx_dy1_m; /* not produced: always NULL */
    gsl_matrix *dx_dy2_m; /* not produced: always NULL */
    /* extras (beyond CSM's struct, at the end so that CSM-style code never sees them) */
    unsigned int corr_hash; /* ld_corr_hash of the final correspondences */
    int n_loops;            /* icp_loop executions: 1 + restarts */
    int status;             /* ZHSM_JOB_* (zhsm.h); -1: rejected before the engine, see stderr */
};

void sm_icp(struct sm_params *params, struct sm_result *res);

/* the same on an engine handle of the caller's (zhsm.h: one handle per host thread) instead of the process-wide one */
struct zhsm_handle;
void zhsm_sm_icp(struct zhsm_handle *h, struct sm_params *params, struct sm_result *res);

/* CSM's logging switches (csm/logging.h): accepted and ignored */
void sm_debug_write(int enabled);

/* releases the process-wide handle sm_icp created (optional: the process exit reclaims it otherwise) */
void zhsm_csm_shutdown(void);

#ifdef __cplusplus
}
#endif
#endif /* ZHSM_CSM_COMPAT_H */
