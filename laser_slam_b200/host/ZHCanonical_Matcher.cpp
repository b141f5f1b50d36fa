// laser_slam_b200/host/ZHCanonical_Matcher.cpp -- see ZHCanonical_Matcher.h.  Host logic only; all arithmetic of the
// matcher runs in libzhsm.so's CUDA kernels through the C ABI (include/zhsm.h).
#include "ZHCanonical_Matcher.h"

#include <stdio.h>
#include <string.h>

#include <stdexcept>
#include <string>

static void check(zhsm_handle *h, zhsm_status st, const char *what) {
    if (st != ZHSM_OK) throw std::runtime_error(std::string(what) + ": " + (h ? zhsm_last_error(h) : "no handle"));
}

CCanonicalMatcher::CCanonicalMatcher(Base_PARAM *p, int device)
    : m_pParam(0), threshold_angular_(0), threshold_dist_sq_(0), h_(0), have_keyframe_(false) {
    memset(&output, 0, sizeof(output));
    if (zhsm_create(device, &h_) != ZHSM_OK)
        throw std::runtime_error("CCanonicalMatcher: no usable sm_100 GPU (this matcher has no CPU path)");
    if (p) m_pParam = new Base_PARAM(*p);
    sm_init(m_pParam);
    // reference constructor (ZHCanonical_Matcher.cpp:40-42)
    threshold_angular_ = (float)(5. * (3.141592654 / 180.));
    threshold_dist_sq_ = 0.02f * 0.02f;
}

CCanonicalMatcher::~CCanonicalMatcher() {
    zhsm_destroy(h_);
    delete m_pParam;
}

// sm_init (ZHCanonical_Matcher.cpp:83-157): the parameter values live in zhsm_plicp_default_params
void CCanonicalMatcher::sm_init(Base_PARAM *p) {
    if (p == 0) {
        zhsm_plicp_default_params(&input, 0, 100000);  // min_reading 0, max_reading 1000 m (:86-88)
        return;
    }
    zhsm_plicp_default_params(&input, p->pm_min_range, p->pm_max_range);
    zhsm_laser lz;
    memset(&lz, 0, sizeof(lz));
    lz.n = p->pm_l_points;
    lz.fi_min = p->pm_fi_min;
    lz.dfi = p->pm_dfi;
    lz.min_range_cm = p->pm_min_range;
    lz.max_range_cm = p->pm_max_range;
    lz.min_valid_points = p->pm_min_valid_points;
    lz.scan_window = p->pm_scan_window;
    check(h_, zhsm_set_laser(h_, &lz), "zhsm_set_laser");
}

void CCanonicalMatcher::setCov(bool on) { input.do_compute_covariance = on ? 1 : 0; }

void CCanonicalMatcher::setThreshold(float dist, float angle) {
    threshold_angular_ = angle;
    threshold_dist_sq_ = dist * dist;
}

void CCanonicalMatcher::median_filter(PMScan *ls) {
    check(h_, zhsm_median_filter_batch(h_, ls->r, 1, 0, 0), "zhsm_median_filter_batch");
}

float CCanonicalMatcher::match_filtered(const float *ref_r, const float *sens_r, PMScan *lsa) {
    const int n = m_pParam->pm_l_points;
    std::vector<float> pool(2 * (size_t)n);
    memcpy(&pool[0], ref_r, sizeof(float) * n);
    memcpy(&pool[n], sens_r, sizeof(float) * n);
    const int32_t ri = 0, si = 1;
    check(h_, zhsm_plicp_batch(h_, &pool[0], 2, &ri, &si, 0, 1, &input, &output, 0, 0, 0, 0), "zhsm_plicp_batch");
    if (!output.valid) {
        printf("icp_failed!\n");  // the reference says so on stdout (:281)
        // FMatch(LDP,LDP) leaves true_pose at the zeros fromPMScan2LDP wrote (:223-225), and the caller still copies it
        lsa->rx = 0; lsa->ry = 0; lsa->th = 0;
        return -1;
    }
    lsa->rx = (float)(output.x[0] * 100.);
    lsa->ry = (float)(output.x[1] * 100.);
    lsa->th = (float)output.x[2];
    return 0;
}

float CCanonicalMatcher::FMatch(PMScan *lsr, PMScan *lsa, bool bFirst) {
    if (bFirst) median_filter(lsr);
    median_filter(lsa);
    return match_filtered(lsr->r, lsa->r, lsa);
}

float CCanonicalMatcher::FMatchKeyFrame(PMScan *lsa) {
    median_filter(lsa);
    const int n = m_pParam->pm_l_points;
    float ret;
    if (!have_keyframe_) {
        ret = 2;  // first laser data (:318-321); true_pose stays 0
        lsa->rx = 0; lsa->ry = 0; lsa->th = 0;
    } else {
        ret = match_filtered(&keyframe_[0], lsa->r, lsa);
    }
    keyframe_.assign(lsa->r, lsa->r + n);  // the key frame is always replaced by the current scan (:367-368)
    have_keyframe_ = true;
    return ret;
}

float CCanonicalMatcher::FMatchKeyFrame2(PMScan *lsa, double cov[9]) {
    float ret = FMatchKeyFrame(lsa);
    if (ret < 0 || ret == 2) return ret;
    // sm_result.cov_x_m, row-major; zhsm_result keeps the upper triangle xx xy xt yy yt tt
    const double *c = output.cov;
    cov[0] = c[0]; cov[1] = c[1]; cov[2] = c[2];
    cov[3] = c[1]; cov[4] = c[3]; cov[5] = c[4];
    cov[6] = c[2]; cov[7] = c[4]; cov[8] = c[5];
    return ret;
}

void CCanonicalMatcher::resetKeyFrame(PMScan *lsa) {
    keyframe_.assign(lsa->r, lsa->r + m_pParam->pm_l_points);  // no filtering here, like the reference (:372-377)
    have_keyframe_ = true;
}

void CCanonicalMatcher::FMatchBatch(PMScan **refs, PMScan **acts, int B, bool filter_refs, float *ret) {
    if (B <= 0) return;
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
    check(h_, zhsm_plicp_batch(h_, &pool[0], 2 * B, &ri[0], &si[0], 0, B, &input, &res[0], 0, 0, 0, 0), "zhsm_plicp_batch");
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
    output = res[B - 1];
}
