// laser_slam_b200/csrc/csm_compat.cu -- the CSM-shaped C entry (include/csm_compat.h): sm_icp / ld_alloc_new / ld_free
// as the reference calls them (src/zhcsm/ZHCanonical_Matcher.cpp:196, :246-247, :274, :344), implemented as a batch of
// one over a process-wide engine handle.  Host logic only: argument checks, LDP <-> scan-pool conversion, bookkeeping
// of the LDP's output fields.  The matching itself is zhsm_plicp_batch (CUDA); there is no CPU matcher here.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "../../include/csm_compat.h"
#include "zhsm_internal.h"

namespace {

std::mutex g_mu;
zhsm_handle *g_h = nullptr;

void reject(struct sm_result *res, const char *why) {
    fprintf(stderr, "sm_icp (libzhsm): %s\n", why);
    res->valid = 0;
    res->status = -1;
}

// theta[i] == (double)(fi_min + dfi * (float)i) in float arithmetic (fromPMScan2LDP :210)?  Finds such a (fi_min, dfi).
bool float_progression(const double *theta, int n, float *fi_min_out, float *dfi_out) {
    const float fi_min = (float)theta[0];
    if ((double)fi_min != theta[0]) return false;
    const float guess = (float)((theta[n - 1] - theta[0]) / (double)(n - 1));
    auto fits = [&](float dfi) {
        for (int i = 0; i < n; i++) {
            volatile float prod = dfi * (float)(unsigned)i;  // two roundings, like the reference's float expression
            volatile float th = fi_min + prod;
            if ((double)th != theta[i]) return false;
        }
        return true;
    };
    float lo = guess, hi = guess;
    for (int k = 0; k <= 64; k++) {
        if (fits(lo)) { *fi_min_out = fi_min; *dfi_out = lo; return true; }
        if (k && fits(hi)) { *fi_min_out = fi_min; *dfi_out = hi; return true; }
        lo = nextafterf(lo, -INFINITY);
        hi = nextafterf(hi, INFINITY);
    }
    return false;
}

// ld_valid_fields' per-scan and per-ray checks (laser_data.c): a scan that fails them makes sm_icp return valid = 0
// before iterating.  (The "at least 10 % valid rays" check runs on the device, plicp2_kernel.cuh load phase.)
const char *ld_host_checks(const struct laser_data *ld) {
    if (!ld) return "NULL laser_data";
    if (ld->nrays < 10 || ld->nrays > ZHSM_MAX_BEAMS) return "nrays out of [10, 4096]";
    if (!ld->theta || !ld->valid || !ld->readings || !ld->cluster) return "laser_data arrays missing";
    if (isnan(ld->min_theta) || isnan(ld->max_theta)) return "min_theta / max_theta is NaN";
    if (fabs(ld->min_theta - ld->theta[0]) > 1e-8 || fabs(ld->max_theta - ld->theta[ld->nrays - 1]) > 1e-8)
        return "min_theta / max_theta do not match theta[]";
    for (int i = 0; i < ld->nrays; i++) {
        if (isnan(ld->theta[i])) return "theta[] holds NaN";
        if (ld->valid[i] != 0 && ld->valid[i] != 1) return "valid[] must be 0 or 1";
        if (ld->valid[i] == 0 && ld->cluster[i] != -1) return "invalid ray with a cluster";
        if (ld->cluster[i] < -1) return "cluster[] below -1";
        if (ld->valid[i]) {
            const double r = ld->readings[i];
            if (isnan(r) || !(0.0 < r && r < 100.0)) return "a valid ray's reading is NaN or outside (0, 100) m";
        }
    }
    return nullptr;
}

// reading [m] -> the float centimetre value whose (double)f / 100.0 is exactly it (fromPMScan2LDP :207); 0 marks an
// invalid ray for the engine (its laser is set with min_range 0: r <= 0 is "invalid in the LDP")
bool to_float_cm(const struct laser_data *ld, float *out) {
    for (int i = 0; i < ld->nrays; i++) {
        if (!ld->valid[i]) { out[i] = 0.f; continue; }
        const double r = ld->readings[i];  // 0 < r < 100 (ld_host_checks)
        float f = (float)(r * 100.0);
        bool ok = false;
        for (int k = 0; k < 3 && !ok; k++) {
            const float c = k == 0 ? f : k == 1 ? nextafterf(f, INFINITY) : nextafterf(f, -INFINITY);
            if ((double)c / 100.0 == r) { f = c; ok = true; }
        }
        if (!ok) return false;
        out[i] = f;
    }
    return true;
}

}  // namespace

extern "C" gsl_matrix *zhsm_gsl_matrix_alloc(size_t rows, size_t cols) {
    gsl_matrix *m = (gsl_matrix *)calloc(1, sizeof(gsl_matrix));
    if (!m) return nullptr;
    m->size1 = rows; m->size2 = cols; m->tda = cols; m->owner = 1;
    m->data = (double *)calloc(rows * cols ? rows * cols : 1, sizeof(double));
    if (!m->data) { free(m); return nullptr; }
    return m;
}

extern "C" void zhsm_gsl_matrix_free(gsl_matrix *m) {
    if (!m) return;
    free(m->data);
    free(m);
}

extern "C" LDP ld_alloc_new(int nrays) {
    if (nrays < 0) return nullptr;
    LDP ld = (LDP)calloc(1, sizeof(struct laser_data));
    if (!ld) return nullptr;
    const size_t n = (size_t)(nrays > 0 ? nrays : 1);
    ld->nrays = nrays;
    ld->theta = (double *)malloc(sizeof(double) * n);
    ld->valid = (int *)calloc(n, sizeof(int));
    ld->readings = (double *)malloc(sizeof(double) * n);
    ld->cluster = (int *)malloc(sizeof(int) * n);
    ld->alpha = (double *)malloc(sizeof(double) * n);
    ld->cov_alpha = (double *)malloc(sizeof(double) * n);
    ld->alpha_valid = (int *)calloc(n, sizeof(int));
    ld->readings_sigma = (double *)malloc(sizeof(double) * n);
    ld->true_alpha = (double *)malloc(sizeof(double) * n);
    ld->corr = (struct correspondence *)calloc(n, sizeof(struct correspondence));
    ld->points = (point2d *)malloc(sizeof(point2d) * n);
    ld->points_w = (point2d *)malloc(sizeof(point2d) * n);
    ld->up_bigger = (int *)calloc(n, sizeof(int));
    ld->up_smaller = (int *)calloc(n, sizeof(int));
    ld->down_bigger = (int *)calloc(n, sizeof(int));
    ld->down_smaller = (int *)calloc(n, sizeof(int));
    if (!ld->theta || !ld->valid || !ld->readings || !ld->cluster || !ld->alpha || !ld->cov_alpha || !ld->alpha_valid ||
        !ld->readings_sigma || !ld->true_alpha || !ld->corr || !ld->points || !ld->points_w || !ld->up_bigger ||
        !ld->up_smaller || !ld->down_bigger || !ld->down_smaller) {
        ld_free(ld);
        return nullptr;
    }
    const double qnan = nan("");
    ld->min_theta = ld->max_theta = qnan;
    for (int i = 0; i < nrays; i++) {
        ld->theta[i] = ld->readings[i] = ld->alpha[i] = ld->cov_alpha[i] = ld->readings_sigma[i] = ld->true_alpha[i] = qnan;
        ld->cluster[i] = -1;
        ld->corr[i].valid = 0; ld->corr[i].j1 = -1; ld->corr[i].j2 = -1; ld->corr[i].type = corr_pp; ld->corr[i].dist2_j1 = qnan;
        ld->points[i].p[0] = ld->points[i].p[1] = ld->points[i].rho = ld->points[i].phi = qnan;
        ld->points_w[i] = ld->points[i];
    }
    for (int k = 0; k < 3; k++) ld->true_pose[k] = ld->odometry[k] = ld->estimate[k] = qnan;
    return ld;
}

extern "C" void ld_free(LDP ld) {
    if (!ld) return;
    free(ld->theta); free(ld->valid); free(ld->readings); free(ld->cluster); free(ld->alpha); free(ld->cov_alpha);
    free(ld->alpha_valid); free(ld->readings_sigma); free(ld->true_alpha); free(ld->corr); free(ld->points);
    free(ld->points_w); free(ld->up_bigger); free(ld->up_smaller); free(ld->down_bigger); free(ld->down_smaller);
    free(ld);
}

extern "C" void sm_debug_write(int) {}

extern "C" void zhsm_csm_shutdown(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_h) zhsm_destroy(g_h);
    g_h = nullptr;
}

static void sm_icp_on(zhsm_handle *use_h, struct sm_params *params, struct sm_result *res);

extern "C" void sm_icp(struct sm_params *params, struct sm_result *res) { sm_icp_on(nullptr, params, res); }

// the same on a handle of the caller's (one handle per host thread, as everywhere in zhsm.h): no global state touched
extern "C" void zhsm_sm_icp(zhsm_handle *h, struct sm_params *params, struct sm_result *res) {
    if (!h) { if (res) reject(res, "zhsm_sm_icp: NULL handle"); return; }
    sm_icp_on(h, params, res);
}

static void sm_icp_on(zhsm_handle *use_h, struct sm_params *params, struct sm_result *res) {
    if (!res) return;
    res->valid = 0;
    res->x[0] = res->x[1] = res->x[2] = 0;
    res->iterations = 0; res->nvalid = 0; res->error = 0;
    res->cov_x_m = res->dx_dy1_m = res->dx_dy2_m = nullptr;
    res->corr_hash = 0; res->n_loops = 0; res->status = 0;
    if (!params) return reject(res, "NULL params");
    LDP ref = params->laser_ref, sens = params->laser_sens;
    if (const char *why = ld_host_checks(ref)) return reject(res, why);
    if (const char *why = ld_host_checks(sens)) return reject(res, why);
    const int n = ref->nrays;
    if (sens->nrays != n) return reject(res, "laser_ref and laser_sens must have the same nrays (one sensor)");
    if (memcmp(ref->theta, sens->theta, sizeof(double) * n) != 0) return reject(res, "laser_ref and laser_sens must share theta[]");
    if (params->use_corr_tricks != 1 || params->use_point_to_line_distance != 1 || params->do_alpha_test ||
        params->do_visibility_test || params->use_ml_weights || params->use_sigma_weights)
        return reject(res, "only the reference's settings are implemented: use_corr_tricks=1, use_point_to_line_distance=1, "
                           "do_alpha_test=do_visibility_test=use_ml_weights=use_sigma_weights=0");
    float fi_min, dfi;
    if (!float_progression(ref->theta, n, &fi_min, &dfi))
        return reject(res, "theta[] is not fi_min + dfi*i in float arithmetic (what fromPMScan2LDP builds)");
    std::vector<float> pool(2 * (size_t)n);
    if (!to_float_cm(ref, &pool[0]) || !to_float_cm(sens, &pool[n]))
        return reject(res, "a valid reading is not a float-centimetre value (readings[i] == (double)f / 100.0)");

    std::unique_lock<std::mutex> lk(g_mu, std::defer_lock);
    zhsm_handle *h = use_h;
    if (!h) {  // the process-wide default handle, serialised
        lk.lock();
        if (!g_h) {
            int dev = 0;
            if (const char *e = getenv("ZHSM_DEVICE")) dev = atoi(e);
            if (zhsm_create(dev, &g_h) != ZHSM_OK) {
                g_h = nullptr;
                return reject(res, "no usable sm_100 GPU (zhsm_create failed); this library has no CPU matcher");
            }
        }
        h = g_h;
    }
    zhsm_laser lz;
    memset(&lz, 0, sizeof(lz));
    lz.n = n; lz.fi_min = fi_min; lz.dfi = dfi;
    lz.min_range_cm = 0;            // validity comes from the LDP: r <= 0 marks the invalid rays (to_float_cm)
    lz.max_range_cm = 0x7fffffff;
    lz.min_valid_points = n / 5; lz.scan_window = n / 10;  // zhpsm fields, unused by PL-ICP
    if (zhsm_set_laser(h, &lz) != ZHSM_OK) return reject(res, zhsm_last_error(h));  // no-op when unchanged
    zhsm_plicp_params P;
    memset(&P, 0, sizeof(P));
    P.max_angular_correction_deg = params->max_angular_correction_deg;
    P.max_linear_correction = params->max_linear_correction;
    P.epsilon_xy = params->epsilon_xy;
    P.epsilon_theta = params->epsilon_theta;
    P.max_correspondence_dist = params->max_correspondence_dist;
    P.sigma = params->sigma;
    P.restart_threshold_mean_error = params->restart_threshold_mean_error;
    P.restart_dt = params->restart_dt;
    P.restart_dtheta = params->restart_dtheta;
    P.outliers_maxPerc = params->outliers_maxPerc;
    P.outliers_adaptive_order = params->outliers_adaptive_order;
    P.outliers_adaptive_mult = params->outliers_adaptive_mult;
    P.min_reading = params->min_reading;
    P.max_reading = params->max_reading;
    P.max_iterations = params->max_iterations;
    P.use_corr_tricks = 1;
    P.restart = params->restart;
    P.use_point_to_line_distance = 1;
    P.outliers_remove_doubles = params->outliers_remove_doubles;
    P.do_compute_covariance = params->do_compute_covariance;
    const int32_t ri = 0, si = 1;
    zhsm_result r;
    std::vector<int16_t> j1(n), j2(n);
    if (zhsm_plicp_batch(h, &pool[0], 2, &ri, &si, params->first_guess, 1, &P, &r, &j1[0], &j2[0], 0, nullptr) != ZHSM_OK)
        return reject(res, zhsm_last_error(h));
    if (lk.owns_lock()) lk.unlock();

    // ---- what CSM leaves behind in the two LDPs
    // ld_invalid_if_outside (sm_icp step 2): valid = 0 where r <= min_reading or r > max_reading
    for (LDP ld : {ref, sens})
        for (int i = 0; i < n; i++) {
            if (!ld->valid[i]) continue;
            const double rr = ld->readings[i];
            if (rr <= params->min_reading || rr > params->max_reading) ld->valid[i] = 0;
        }
    const double qnan = nan("");
    for (LDP ld : {ref, sens})  // ld_compute_cartesian
        for (int i = 0; i < n; i++) {
            const bool v = ld->valid[i] != 0;
            ld->points[i].p[0] = v ? cos(ld->theta[i]) * ld->readings[i] : qnan;
            ld->points[i].p[1] = v ? sin(ld->theta[i]) * ld->readings[i] : qnan;
            ld->points[i].rho = ld->points[i].phi = qnan;
        }
    res->valid = r.valid;
    res->status = r.status;
    res->iterations = r.iterations;
    res->n_loops = r.n_loops;
    if (!r.valid) {
        for (int i = 0; i < n; i++) { sens->corr[i].valid = 0; sens->corr[i].j1 = -1; sens->corr[i].j2 = -1; }
        return;
    }
    for (int k = 0; k < 3; k++) res->x[k] = r.x[k];
    res->nvalid = r.nvalid;
    res->error = r.error;
    res->corr_hash = r.corr_hash;
    const double c = cos(r.x[2]), s = sin(r.x[2]);
    for (int i = 0; i < n; i++) {  // ld_compute_world_coords(laser_sens, x) + the final correspondences
        point2d &w = sens->points_w[i];
        if (sens->valid[i]) {
            const double px = sens->points[i].p[0], py = sens->points[i].p[1];
            w.p[0] = c * px - s * py + r.x[0];
            w.p[1] = s * px + c * py + r.x[1];
            w.rho = sqrt(w.p[0] * w.p[0] + w.p[1] * w.p[1]);
            w.phi = atan2(w.p[1], w.p[0]);
        } else {
            w.p[0] = w.p[1] = w.rho = w.phi = qnan;
        }
        struct correspondence &co = sens->corr[i];
        if (j1[i] >= 0) {
            co.valid = 1; co.j1 = j1[i]; co.j2 = j2[i]; co.type = corr_pl;
            const double dx = w.p[0] - ref->points[j1[i]].p[0], dy = w.p[1] - ref->points[j1[i]].p[1];
            co.dist2_j1 = dx * dx + dy * dy;
        } else {
            co.valid = 0; co.j1 = -1; co.j2 = -1; co.dist2_j1 = qnan;
        }
    }
    if (params->do_compute_covariance) {
        res->cov_x_m = zhsm_gsl_matrix_alloc(3, 3);
        if (res->cov_x_m) {
            const double *cv = r.cov;  // upper triangle xx xy xt yy yt tt
            const double full[9] = {cv[0], cv[1], cv[2], cv[1], cv[3], cv[4], cv[2], cv[4], cv[5]};
            memcpy(res->cov_x_m->data, full, sizeof(full));
        }
    }
}
