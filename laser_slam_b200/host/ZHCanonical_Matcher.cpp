// laser_slam_b200/host/ZHCanonical_Matcher.cpp -- see ZHCanonical_Matcher.h.  Host logic only; all arithmetic of the
// matcher runs in libzhsm.so's CUDA kernels through the C ABI (include/zhsm.h, include/csm_compat.h).
#include "ZHCanonical_Matcher.h"

#include <math.h>
#include <stdio.h>
#include <string.h>

#include <stdexcept>
#include <string>

static void check(zhsm_handle *h, zhsm_status st, const char *what) {
    if (st != ZHSM_OK) throw std::runtime_error(std::string(what) + ": " + (h ? zhsm_last_error(h) : "no handle"));
}

CCanonicalMatcher::CCanonicalMatcher(Base_PARAM *p, int device)
    : m_pKeyframe(0), threshold_angular_(0), threshold_dist_sq_(0), input(new sm_params), output(new sm_result),
      m_pParam(0), h_(0) {
    memset(input, 0, sizeof(*input));  // the reference leaves the fields sm_init does not set uninitialised (:20)
    memset(output, 0, sizeof(*output));
    if (zhsm_create(device, &h_) != ZHSM_OK) {
        delete input;
        delete output;
        throw std::runtime_error("CCanonicalMatcher: no usable sm_100 GPU (this matcher has no CPU path)");
    }
    if (p) m_pParam = new Base_PARAM(*p);
    sm_init(m_pParam);
    // reference constructor (ZHCanonical_Matcher.cpp:40-42)
    threshold_angular_ = (float)(5. * (3.141592654 / 180.));
    threshold_dist_sq_ = 0.02f * 0.02f;
}

CCanonicalMatcher::~CCanonicalMatcher() {
    if (m_pKeyframe) ld_free(m_pKeyframe);
    drop_cov();
    zhsm_destroy(h_);
    delete input;
    delete output;
    delete m_pParam;
}

// the reference never frees cov_x_m ("leak"); the shim owns the last one
void CCanonicalMatcher::drop_cov() {
    if (output->cov_x_m) zhsm_gsl_matrix_free(output->cov_x_m);
    output->cov_x_m = 0;
}

// sm_init (ZHCanonical_Matcher.cpp:83-157): the values come from zhsm_plicp_default_params, the one place that holds them
void CCanonicalMatcher::sm_init(Base_PARAM *p) {
    zhsm_plicp_params d;
    if (p == 0) zhsm_plicp_default_params(&d, 0, 100000);  // min_reading 0, max_reading 1000 m (:86-88)
    else zhsm_plicp_default_params(&d, p->pm_min_range, p->pm_max_range);
    input->min_reading = d.min_reading;
    input->max_reading = d.max_reading;
    input->max_angular_correction_deg = d.max_angular_correction_deg;
    input->max_linear_correction = d.max_linear_correction;
    input->max_iterations = d.max_iterations;
    input->epsilon_xy = d.epsilon_xy;
    input->epsilon_theta = d.epsilon_theta;
    input->max_correspondence_dist = d.max_correspondence_dist;
    input->sigma = d.sigma;
    input->use_corr_tricks = d.use_corr_tricks;
    input->restart = d.restart;
    input->restart_threshold_mean_error = d.restart_threshold_mean_error;
    input->restart_dt = d.restart_dt;
    input->restart_dtheta = d.restart_dtheta;
    input->clustering_threshold = 0.05;      // :109
    input->orientation_neighbourhood = 10;   // :110
    input->use_point_to_line_distance = d.use_point_to_line_distance;
    input->do_alpha_test = 0;                // :113
    input->do_alpha_test_thresholdDeg = 20.0;
    input->outliers_maxPerc = d.outliers_maxPerc;
    input->outliers_adaptive_order = d.outliers_adaptive_order;
    input->outliers_adaptive_mult = d.outliers_adaptive_mult;
    input->do_visibility_test = 0;           // :135
    input->outliers_remove_doubles = d.outliers_remove_doubles;
    input->do_compute_covariance = d.do_compute_covariance;
    input->debug_verify_tricks = 0;
    input->use_ml_weights = 0;
    input->use_sigma_weights = 0;
}

void CCanonicalMatcher::setCov(bool on) { input->do_compute_covariance = on; }

void CCanonicalMatcher::setThreshold(float dist, float angle) {
    threshold_angular_ = angle;
    threshold_dist_sq_ = dist * dist;
}

// :379-387
bool CCanonicalMatcher::newKeyframeNeeded(double pose[3]) {
    const float x = pose[0], y = pose[1], th = pose[2];  // narrowed like the reference; the angle is NOT taken absolute
    if (th > threshold_angular_) return true;
    if (x * x + y * y > threshold_dist_sq_) return true;
    return false;
}

void CCanonicalMatcher::set_pm_laser() {
    if (!m_pParam) throw std::runtime_error("CCanonicalMatcher: PMScan entry points need the Base_PARAM geometry");
    zhsm_laser lz;
    memset(&lz, 0, sizeof(lz));
    lz.n = m_pParam->pm_l_points;
    lz.fi_min = m_pParam->pm_fi_min;
    lz.dfi = m_pParam->pm_dfi;
    lz.min_range_cm = m_pParam->pm_min_range;
    lz.max_range_cm = m_pParam->pm_max_range;
    lz.min_valid_points = m_pParam->pm_min_valid_points;
    lz.scan_window = m_pParam->pm_scan_window;
    check(h_, zhsm_set_laser(h_, &lz), "zhsm_set_laser");  // a no-op while the geometry is unchanged
}

void CCanonicalMatcher::median_filter(PMScan *ls) {
    set_pm_laser();
    check(h_, zhsm_median_filter_batch(h_, ls->r, 1, 0, 0), "zhsm_median_filter_batch");
}

// :193-227
void CCanonicalMatcher::fromPMScan2LDP(PMScan *pm_scan, LDP *ldp, Base_PARAM *param) {
    const unsigned int n = pm_scan->np;
    *ldp = ld_alloc_new(n);
    if (!*ldp) throw std::bad_alloc();
    LDP ld = *ldp;
    for (unsigned int i = 0; i < n; i++) {
        const double r = pm_scan->r[i];
        if (r <= param->pm_min_range || r >= param->pm_max_range) {
            ld->valid[i] = 0;
            ld->readings[i] = -1;
        } else {
            ld->valid[i] = 1;
            ld->readings[i] = r / 100.0;
        }
        ld->theta[i] = param->pm_fi_min + param->pm_dfi * i;  // float arithmetic, widened on assignment
        ld->cluster[i] = -1;
    }
    ld->min_theta = ld->theta[0];
    ld->max_theta = ld->theta[n - 1];
    for (int k = 0; k < 3; k++) ld->odometry[k] = ld->estimate[k] = ld->true_pose[k] = 0.0;
}

float CCanonicalMatcher::FMatch(PMScan *lsr, PMScan *lsa, bool bFirst) {
    LDP ldpref, ldpcur;
    if (bFirst) median_filter(lsr);
    median_filter(lsa);
    fromPMScan2LDP(lsr, &ldpref, m_pParam);
    fromPMScan2LDP(lsa, &ldpcur, m_pParam);
    const float ret = FMatch(ldpref, ldpcur);
    lsa->rx = ldpcur->true_pose[0] * 100.;
    lsa->ry = ldpcur->true_pose[1] * 100.;
    lsa->th = ldpcur->true_pose[2];
    ld_free(ldpref);
    ld_free(ldpcur);
    return ret;
}

// :251-285
float CCanonicalMatcher::FMatch(LDP lsr, LDP lsa) {
    float ret = 0;
    for (int k = 0; k < 3; k++) lsr->odometry[k] = lsr->estimate[k] = lsr->true_pose[k] = 0.0;
    input->laser_ref = lsr;
    input->laser_sens = lsa;
    input->first_guess[0] = input->first_guess[1] = input->first_guess[2] = 0.0;
    drop_cov();
    zhsm_sm_icp(h_, input, output);  // sm_icp on this matcher's own engine handle
    if (output->valid) {
        for (int i = 0; i < 3; i++) lsa->true_pose[i] = output->x[i];
    } else {
        printf("icp_failed!\n");  // the reference says so on stdout (:281)
        ret = -1;
    }
    return ret;
}

float CCanonicalMatcher::FMatchKeyFrame2(PMScan *lsa, double cov[9]) {
    float ret = FMatchKeyFrame(lsa);
    if (ret < 0 || ret == 2) return ret;
    if (!output->cov_x_m) throw std::runtime_error("FMatchKeyFrame2: setCov(true) first (the reference reads a null matrix here)");
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) cov[3 * i + j] = gsl_matrix_get(output->cov_x_m, i, j);
    return ret;
}

// :300-313
float CCanonicalMatcher::FMatchKeyFrame(PMScan *lsa) {
    LDP ldpcur;
    median_filter(lsa);
    fromPMScan2LDP(lsa, &ldpcur, m_pParam);
    const float ret = FMatchKeyFrame(ldpcur);  // ldpcur becomes the key frame: not freed here
    lsa->rx = ldpcur->true_pose[0] * 100.;
    lsa->ry = ldpcur->true_pose[1] * 100.;
    lsa->th = ldpcur->true_pose[2];
    return ret;
}

// :315-370
float CCanonicalMatcher::FMatchKeyFrame(LDP lsa) {
    float ret = 0;
    if (m_pKeyframe == 0) {
        m_pKeyframe = lsa;
        return 2;  // first laser data
    }
    for (int k = 0; k < 3; k++) m_pKeyframe->odometry[k] = m_pKeyframe->estimate[k] = m_pKeyframe->true_pose[k] = 0.0;
    input->laser_ref = m_pKeyframe;
    input->laser_sens = lsa;
    input->first_guess[0] = input->first_guess[1] = input->first_guess[2] = 0.0;
    drop_cov();
    zhsm_sm_icp(h_, input, output);
    if (output->valid) {
        for (int i = 0; i < 3; i++) lsa->true_pose[i] = output->x[i];
    } else {
        printf("icp_failed!\n");
        ret = -1;
    }
    ld_free(m_pKeyframe);  // the key frame is always replaced by the current scan (:367-368)
    m_pKeyframe = lsa;
    return ret;
}

// :372-377: no filtering here, like the reference
void CCanonicalMatcher::resetKeyFrame(PMScan *lsa) {
    LDP ldpcur;
    fromPMScan2LDP(lsa, &ldpcur, m_pParam);
    if (m_pKeyframe) ld_free(m_pKeyframe);
    m_pKeyframe = ldpcur;
}

zhsm_plicp_params CCanonicalMatcher::engine_params() const {
    zhsm_plicp_params P;
    memset(&P, 0, sizeof(P));
    P.max_angular_correction_deg = input->max_angular_correction_deg;
    P.max_linear_correction = input->max_linear_correction;
    P.epsilon_xy = input->epsilon_xy;
    P.epsilon_theta = input->epsilon_theta;
    P.max_correspondence_dist = input->max_correspondence_dist;
    P.sigma = input->sigma;
    P.restart_threshold_mean_error = input->restart_threshold_mean_error;
    P.restart_dt = input->restart_dt;
    P.restart_dtheta = input->restart_dtheta;
    P.outliers_maxPerc = input->outliers_maxPerc;
    P.outliers_adaptive_order = input->outliers_adaptive_order;
    P.outliers_adaptive_mult = input->outliers_adaptive_mult;
    P.min_reading = input->min_reading;
    P.max_reading = input->max_reading;
    P.max_iterations = input->max_iterations;
    P.use_corr_tricks = input->use_corr_tricks;
    P.restart = input->restart;
    P.use_point_to_line_distance = input->use_point_to_line_distance;
    P.outliers_remove_doubles = input->outliers_remove_doubles;
    P.do_compute_covariance = input->do_compute_covariance;
    return P;
}

void CCanonicalMatcher::FMatchBatch(PMScan **refs, PMScan **acts, int B, bool filter_refs, float *ret) {
    if (B <= 0) return;
    set_pm_laser();
    const int n = m_pParam->pm_l_points;
    std::vector<float> pool(2 * (size_t)B * n);
    std::vector<int32_t> ri(B), si(B);
    for (int k = 0; k < B; k++) {
        memcpy(&pool[(2 * (size_t)k) * n], refs[k]->r, sizeof(float) * n);
        memcpy(&pool[(2 * (size_t)k + 1) * n], acts[k]->r, sizeof(float) * n);
        ri[k] = 2 * k;
        si[k] = 2 * k + 1;
    }
    if (filter_refs) {
        check(h_, zhsm_median_filter_batch(h_, &pool[0], 2 * B, 0, 0), "zhsm_median_filter_batch");
    } else {
        // only the current scans are filtered (bFirst == false)
        std::vector<float> cur((size_t)B * n);
        for (int k = 0; k < B; k++) memcpy(&cur[(size_t)k * n], acts[k]->r, sizeof(float) * n);
        check(h_, zhsm_median_filter_batch(h_, &cur[0], B, 0, 0), "zhsm_median_filter_batch");
        for (int k = 0; k < B; k++) memcpy(&pool[(2 * (size_t)k + 1) * n], &cur[(size_t)k * n], sizeof(float) * n);
    }
    std::vector<zhsm_result> res(B);
    const zhsm_plicp_params P = engine_params();
    check(h_, zhsm_plicp_batch(h_, &pool[0], 2 * B, &ri[0], &si[0], 0, B, &P, &res[0], 0, 0, 0, 0), "zhsm_plicp_batch");
    for (int k = 0; k < B; k++) {
        if (filter_refs) memcpy(refs[k]->r, &pool[(2 * (size_t)k) * n], sizeof(float) * n);
        memcpy(acts[k]->r, &pool[(2 * (size_t)k + 1) * n], sizeof(float) * n);
        const zhsm_result &r = res[k];
        if (r.valid) {
            acts[k]->rx = (float)(r.x[0] * 100.);
            acts[k]->ry = (float)(r.x[1] * 100.);
            acts[k]->th = (float)r.x[2];
            ret[k] = 0;
        } else {
            acts[k]->rx = 0; acts[k]->ry = 0; acts[k]->th = 0;
            ret[k] = -1;
        }
    }
    // the last pair's result, in the shape single-pair callers read
    const zhsm_result &r = res[B - 1];
    drop_cov();
    output->valid = r.valid;
    for (int k = 0; k < 3; k++) output->x[k] = r.x[k];
    output->iterations = r.iterations; output->nvalid = r.nvalid; output->error = r.error;
    output->corr_hash = r.corr_hash; output->n_loops = r.n_loops; output->status = r.status;
}
