/* tests/cpp/csm_plain.c -- a plain-C program that does exactly what the reference's wrapper does around sm_icp
 * (src/zhcsm/ZHCanonical_Matcher.cpp:193-285): allocate two LDPs, fill them from float-centimetre ranges, set the
 * parameters of sm_init (:83-157), call sm_icp, read valid / x / iterations (+ cov_x_m through gsl_matrix_get).
 * It includes <csm/csm_all.h> and links against libzhsm.so ONLY.
 *   csm_plain <n> <fi_min> <dfi> <min_range_cm> <max_range_cm> <scans.f32> <S> <cov 0|1> <out.txt>
 * scans.f32: S median-filtered scans; pair k = (scan k, scan k+1).  Output per pair:
 *   valid x y theta iterations nvalid error corr_hash [c00 c01 c02 c11 c12 c22] */
#include <csm/csm_all.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static LDP from_ranges(const float *r_cm, int n, float fi_min, float dfi, int min_cm, int max_cm) {
    LDP ld = ld_alloc_new(n);
    int i;
    if (!ld) return NULL;
    for (i = 0; i < n; i++) {
        double r = r_cm[i];
        if (r <= min_cm || r >= max_cm) { ld->valid[i] = 0; ld->readings[i] = -1; }
        else { ld->valid[i] = 1; ld->readings[i] = r / 100.0; }
        ld->theta[i] = fi_min + dfi * (unsigned)i; /* float arithmetic, widened on assignment (:210) */
        ld->cluster[i] = -1;
    }
    ld->min_theta = ld->theta[0];
    ld->max_theta = ld->theta[n - 1];
    for (i = 0; i < 3; i++) ld->odometry[i] = ld->estimate[i] = ld->true_pose[i] = 0.0;
    return ld;
}

int main(int argc, char **argv) {
    struct sm_params in;
    struct sm_result out;
    int n, S, k, i, min_cm, max_cm, want_cov;
    float fi_min, dfi, *all;
    FILE *f, *o;
    if (argc < 10) return 2;
    n = atoi(argv[1]); fi_min = (float)atof(argv[2]); dfi = (float)atof(argv[3]);
    min_cm = atoi(argv[4]); max_cm = atoi(argv[5]); S = atoi(argv[7]); want_cov = atoi(argv[8]);
    all = (float *)malloc(sizeof(float) * (size_t)S * n);
    f = fopen(argv[6], "rb");
    if (!f || fread(all, sizeof(float), (size_t)S * n, f) != (size_t)S * n) return 3;
    fclose(f);
    o = fopen(argv[9], "w");
    if (!o) return 3;
    memset(&in, 0, sizeof(in));
    /* CCanonicalMatcher::sm_init (:83-157) */
    in.min_reading = min_cm / 100.; in.max_reading = max_cm / 100.;
    in.max_angular_correction_deg = 45; in.max_linear_correction = 0.5;
    in.max_iterations = 10; in.epsilon_xy = 0.001; in.epsilon_theta = 0.001;
    in.max_correspondence_dist = 2; in.sigma = 0.04;
    in.use_corr_tricks = 1; in.restart = 1; in.restart_threshold_mean_error = 0.01;
    in.restart_dt = 0.01; in.restart_dtheta = 0.0261799;
    in.clustering_threshold = 0.05; in.orientation_neighbourhood = 10; in.use_point_to_line_distance = 1;
    in.do_alpha_test = 0; in.do_alpha_test_thresholdDeg = 20.0;
    in.outliers_maxPerc = 0.95; in.outliers_adaptive_order = 0.7; in.outliers_adaptive_mult = 2.0;
    in.do_visibility_test = 0; in.outliers_remove_doubles = 1; in.do_compute_covariance = want_cov;
    in.debug_verify_tricks = 0; in.use_ml_weights = 0; in.use_sigma_weights = 0;
    for (k = 0; k + 1 < S; k++) {
        LDP ref = from_ranges(all + (size_t)k * n, n, fi_min, dfi, min_cm, max_cm);
        LDP cur = from_ranges(all + (size_t)(k + 1) * n, n, fi_min, dfi, min_cm, max_cm);
        int ncorr = 0;
        in.laser_ref = ref; in.laser_sens = cur;
        in.first_guess[0] = in.first_guess[1] = in.first_guess[2] = 0.0;
        sm_icp(&in, &out);
        for (i = 0; i < n; i++) ncorr += cur->corr[i].valid;
        fprintf(o, "%d %.17g %.17g %.17g %d %d %.17g %u %d", out.valid, out.x[0], out.x[1], out.x[2], out.iterations,
                out.nvalid, out.error, out.corr_hash, ncorr);
        if (want_cov && out.valid && out.cov_x_m)
            fprintf(o, " %.17g %.17g %.17g %.17g %.17g %.17g", gsl_matrix_get(out.cov_x_m, 0, 0),
                    gsl_matrix_get(out.cov_x_m, 0, 1), gsl_matrix_get(out.cov_x_m, 0, 2), gsl_matrix_get(out.cov_x_m, 1, 1),
                    gsl_matrix_get(out.cov_x_m, 1, 2), gsl_matrix_get(out.cov_x_m, 2, 2));
        fprintf(o, "\n");
        if (out.cov_x_m) zhsm_gsl_matrix_free(out.cov_x_m);
        ld_free(ref);
        ld_free(cur);
    }
    fclose(o);
    free(all);
    return 0;
}
