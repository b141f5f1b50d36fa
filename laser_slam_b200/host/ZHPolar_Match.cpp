// laser_slam_b200/host/ZHPolar_Match.cpp -- see ZHPolar_Match.h.  Host logic only; the matchers run in libzhsm.so's
// CUDA kernels (csrc/psm_sm100.cu) through the C ABI.
#include "ZHPolar_Match.h"

#include <math.h>
#include <stdio.h>
#include <string.h>

#include <stdexcept>

static void check(zhsm_handle *h, zhsm_status st, const char *what) {
    if (st != ZHSM_OK) throw std::runtime_error(std::string(what) + ": " + (h ? zhsm_last_error(h) : "no handle"));
}

CPolarMatch::CPolarMatch(std::string laser_name, int device) : m_pParam(0), m_bReady(false), h_(0) {
    if (zhsm_create(device, &h_) != ZHSM_OK)
        throw std::runtime_error("CPolarMatch: no usable sm_100 GPU (this matcher has no CPU path)");
    if (laser_name == "LMS511") m_pParam = new PM_SICK_LMS511();
    else if (laser_name == "LMS211") m_pParam = new PM_SICK_LMS211();
    else if (laser_name == "LMS151") m_pParam = new PM_SICK_LMS151();
    if (!m_pParam) { printf("laser is not right!\n"); return; }  // the reference's message (ZHPolar_Match.cpp:35)
    m_bReady = true;
    pm_init();
}

// pm_init (ZHPolar_Match.cpp:68-78) rebuilds the bearing tables from m_pParam; here that is the device-side geometry.
// Call it again after changing m_pParam (readCarmon overwrites pm_fi_min / pm_max_range from the log header, :227-235).
void CPolarMatch::pm_init() {
    if (!m_bReady) return;
    zhsm_laser lz;
    memset(&lz, 0, sizeof(lz));
    lz.n = m_pParam->pm_l_points;
    lz.fi_min = m_pParam->pm_fi_min;
    lz.dfi = m_pParam->pm_dfi;
    lz.min_range_cm = m_pParam->pm_min_range;
    lz.max_range_cm = m_pParam->pm_max_range;
    lz.min_valid_points = m_pParam->pm_min_valid_points;
    lz.scan_window = m_pParam->pm_scan_window;
    check(h_, zhsm_set_laser(h_, &lz), "zhsm_set_laser");
}

void CPolarMatch::clearRecord() {
    for (size_t i = 0; i < m_SickScans.size(); i++) delete m_SickScans[i];
    m_SickScans.clear();
}

bool CPolarMatch::readCarmon(std::string filename, std::string laser) {
    if (!m_bReady) { printf("Laser is not right initialized!\n"); return false; }
    clearRecord();
    ZhCarmenLog log;
    if (!zh_read_carmen(filename, laser, m_pParam, &log)) return false;
    pm_init();  // the header may have changed pm_fi_min / pm_max_range (:227-235)
    for (int k = 0; k < log.S; k++) {
        PMScan *ls = new PMScan(log.n);
        memcpy(ls->r, &log.r[(size_t)k * log.n], sizeof(float) * log.n);
        ls->rx = (float)log.pose[3 * k]; ls->ry = (float)log.pose[3 * k + 1]; ls->th = (float)log.pose[3 * k + 2];
        m_SickScans.push_back(ls);
    }
    return !m_SickScans.empty();
}

CPolarMatch::~CPolarMatch() {
    clearRecord();
    zhsm_destroy(h_);
    delete m_pParam;
}

namespace {
// PMScan keeps bad / seg as int; the C ABI packs them as uint8 / int16
struct Packed {
    std::vector<float> r;
    std::vector<uint8_t> bad;
    std::vector<int16_t> seg;
    int n;
    Packed(int scans, int n_) : r((size_t)scans * n_), bad((size_t)scans * n_), seg((size_t)scans * n_), n(n_) {}
    void put(int k, const PMScan *s) {
        memcpy(&r[(size_t)k * n], s->r, sizeof(float) * n);
        for (int i = 0; i < n; i++) { bad[(size_t)k * n + i] = (uint8_t)s->bad[i]; seg[(size_t)k * n + i] = (int16_t)s->seg[i]; }
    }
    void get(int k, PMScan *s) const {
        memcpy(s->r, &r[(size_t)k * n], sizeof(float) * n);
        for (int i = 0; i < n; i++) { s->bad[i] = bad[(size_t)k * n + i]; s->seg[i] = seg[(size_t)k * n + i]; }
    }
};
}  // namespace

void CPolarMatch::pm_median_filter(PMScan *ls) {
    check(h_, zhsm_median_filter_batch(h_, ls->r, 1, 0, 0), "zhsm_median_filter_batch");
}

void CPolarMatch::pm_preprocessScan(PMScan *ls) {
    Packed p(1, m_pParam->pm_l_points);
    p.put(0, ls);
    check(h_, zhsm_preprocess_batch(h_, &p.r[0], &p.bad[0], &p.seg[0], 1, 0, 0), "zhsm_preprocess_batch");
    p.get(0, ls);
}

void CPolarMatch::pm_scan_project(const PMScan *act, float *new_r, int *new_bad) {
    const int n = m_pParam->pm_l_points;
    Packed p(1, n);
    p.put(0, act);
    const int32_t idx = 0;
    const float pose[3] = {act->rx, act->ry, act->th};
    std::vector<uint8_t> nb(n);
    check(h_, zhsm_scan_project_batch(h_, &p.r[0], &p.bad[0], &p.seg[0], 1, &idx, pose, 1, new_r, &nb[0]), "zhsm_scan_project_batch");
    for (int i = 0; i < n; i++) new_bad[i] = nb[i];
}

bool CPolarMatch::pm_is_corridor(PMScan *act) {
    Packed p(1, m_pParam->pm_l_points);
    p.put(0, act);
    int8_t out = 0;
    check(h_, zhsm_is_corridor_batch(h_, &p.r[0], &p.bad[0], &p.seg[0], 1, &out, 0, 0), "zhsm_is_corridor_batch");
    if (out < 0) throw (int)(-out);
    return out == 1;
}

// pm_psm / pm_icp move the current scan into the reference scan's frame first (:918-925, PM_LASER_Y = 0), in float
// with the host's sinf/cosf exactly as the reference does; the kernel then works with the ref scan at the origin.
float CPolarMatch::match(int algo, const PMScan *lsr, PMScan *lsa) {
    const float rx = lsr->rx, ry = lsr->ry, rth = lsr->th, ax = lsa->rx, ay = lsa->ry, ath = lsa->th, LASER_Y = 0;
    const float t13 = sinf(rth - ath) * LASER_Y + cosf(rth) * ax + sinf(rth) * ay - sinf(rth) * ry - rx * cosf(rth);
    const float t23 = cosf(rth - ath) * LASER_Y - sinf(rth) * ax + cosf(rth) * ay - cosf(rth) * ry + rx * sinf(rth) - LASER_Y;
    const float init[3] = {t13, t23, ath - rth};
    Packed p(2, m_pParam->pm_l_points);
    p.put(0, lsr);
    p.put(1, lsa);
    const int32_t ri = 0, ai = 1;
    zhsm_result res;
    check(h_, zhsm_psm_batch(h_, &p.r[0], &p.bad[0], &p.seg[0], 2, &ri, &ai, init, 1, algo, &res, 0, 0), "zhsm_psm_batch");
    if (!res.valid) throw 1;  // the reference's error signal; lsa is left as it was
    lsa->rx = (float)res.x[0];
    lsa->ry = (float)res.x[1];
    lsa->th = (float)res.x[2];
    return (float)res.error;
}

float CPolarMatch::pm_psm(const PMScan *lsr, PMScan *lsa) { return match(ZHSM_ALGO_PSM, lsr, lsa); }
float CPolarMatch::pm_icp(const PMScan *lsr, PMScan *lsa) { return match(ZHSM_ALGO_PM_ICP, lsr, lsa); }

std::pair<float, float> CPolarMatch::FMatch(PMScan *lsr, PMScan *lsa, bool bFirst) {
    std::pair<float, float> out(-1.f, -1.f);
    if (lsr == 0 || lsa == 0) { printf("Error in FMatch!\n"); return out; }
    PMScan *refs[1] = {lsr}, *acts[1] = {lsa};
    FMatchBatch(refs, acts, 1, bFirst, &out);
    return out;
}

void CPolarMatch::FMatchBatch(PMScan **refs, PMScan **acts, int B, bool preprocess_refs, std::pair<float, float> *out) {
    if (B <= 0) return;
    const int n = m_pParam->pm_l_points;
    Packed p(2 * B, n);
    std::vector<int32_t> ri(B), ai(B);
    for (int k = 0; k < B; k++) { p.put(2 * k, refs[k]); p.put(2 * k + 1, acts[k]); ri[k] = 2 * k; ai[k] = 2 * k + 1; }
    if (preprocess_refs) {
        check(h_, zhsm_preprocess_batch(h_, &p.r[0], &p.bad[0], &p.seg[0], 2 * B, 0, 0), "zhsm_preprocess_batch");
    } else {
        Packed c(B, n);
        for (int k = 0; k < B; k++) c.put(k, acts[k]);
        check(h_, zhsm_preprocess_batch(h_, &c.r[0], &c.bad[0], &c.seg[0], B, 0, 0), "zhsm_preprocess_batch");
        for (int k = 0; k < B; k++) {
            memcpy(&p.r[(size_t)(2 * k + 1) * n], &c.r[(size_t)k * n], sizeof(float) * n);
            memcpy(&p.bad[(size_t)(2 * k + 1) * n], &c.bad[(size_t)k * n], n);
            memcpy(&p.seg[(size_t)(2 * k + 1) * n], &c.seg[(size_t)k * n], sizeof(int16_t) * n);
        }
    }
    std::vector<zhsm_result> res(B);
    check(h_, zhsm_psm_batch(h_, &p.r[0], &p.bad[0], &p.seg[0], 2 * B, &ri[0], &ai[0], 0, B, ZHSM_ALGO_FMATCH, &res[0], 0, 0),
          "zhsm_psm_batch");
    for (int k = 0; k < B; k++) {
        if (preprocess_refs) p.get(2 * k, refs[k]);
        p.get(2 * k + 1, acts[k]);
        refs[k]->rx = refs[k]->ry = refs[k]->th = 0;  // FMatch zeroes both poses before matching (:97-99)
        acts[k]->rx = acts[k]->ry = acts[k]->th = 0;
        if (res[k].valid) {
            acts[k]->rx = (float)res[k].x[0]; acts[k]->ry = (float)res[k].x[1]; acts[k]->th = (float)res[k].x[2];
            out[k] = std::make_pair((float)res[k].cov[0], (float)res[k].cov[3]);
        } else {
            if (B == 1) printf("frontend match failed!\n");  // deterministic matcher: the retry (:108-117) fails the same way
            out[k] = std::make_pair(-1.f, -1.f);
        }
    }
}
