// oracle/zhpsm_ref_driver.cpp -- TEST INFRASTRUCTURE ONLY.
//
// extern "C" driver around the reference's OWN zhpsm sources, which are compiled unmodified from
// /root/reference/src/zhpsm/{ZHPolar_Match,point}.cpp by oracle/Makefile into oracle/_ref/libzhpsm_ref.so
// (git-ignored; travels to the GPU box as a prebuilt file).  Nothing here re-implements the matcher:
// every call below lands in the reference's CPolarMatch (src/zhpsm/ZHPolar_Match.h:16-88).
//
// Status codes mirror the reference's error behaviour: 0 ok, 1 = the callee threw `int 1`
// (too few points / singular / search failed: ZHPolar_Match.cpp:1095,1106,1239,1831).
#include <cstdio>
#include <cstring>
#include <iostream>
#include <pthread.h>
#include <streambuf>
#include <string>
#include <unistd.h>
#include <fcntl.h>

#include "ZHPolar_Match.h"
#include "point.h"

namespace {
struct NullBuf : std::streambuf {
    int overflow(int c) override { return c; }
};
NullBuf g_nullbuf;
bool g_silenced = false;
void silence() {
    // the reference chats on cerr/cout from inside the hot loop; keep timing and test logs clean
    if (!g_silenced) {
        std::cerr.rdbuf(&g_nullbuf);
        std::cout.rdbuf(&g_nullbuf);
        g_silenced = true;
    }
}

struct Ref {
    CPolarMatch *m;
    int n;
};

void fill_scan(Ref *h, PMScan &ls, const float *r, const int *bad, const int *seg, const float pose[3]) {
    int n = h->n;
    memcpy(ls.r, r, sizeof(float) * n);
    for (int i = 0; i < n; i++) {
        // readCarmon (:253-255): x,y from the unfiltered range; bad starts at 0
        ls.x[i] = ls.r[i] * h->m->pm_co[i];
        ls.y[i] = ls.r[i] * h->m->pm_si[i];
        ls.bad[i] = bad ? bad[i] : 0;
        ls.seg[i] = seg ? seg[i] : 0;
    }
    ls.rx = pose ? pose[0] : 0;
    ls.ry = pose ? pose[1] : 0;
    ls.th = pose ? pose[2] : 0;
    ls.t = 0;
}
}  // namespace

extern "C" {

// laser: "LMS211" | "LMS511" | "LMS151" (src/zhpsm/PolarParameter.h:70-83).  fi_min/max_range_cm override the
// table exactly like readCarmon does from the first log record (ZHPolar_Match.cpp:227-235); pass NAN / <=0 to keep.
void *zhpsm_ref_create(const char *laser, float fi_min_override, int max_range_cm_override) {
    silence();
    CPolarMatch *m = new CPolarMatch(std::string(laser));
    if (!m->m_bReady) { delete m; return 0; }
    if (fi_min_override == fi_min_override) {  // not NaN
        m->m_pParam->pm_fi_min = fi_min_override;
        m->pm_init();
    }
    if (max_range_cm_override > 0) m->m_pParam->pm_max_range = max_range_cm_override;
    Ref *h = new Ref;
    h->m = m;
    h->n = m->m_pParam->pm_l_points;
    return h;
}
// pm_orientation_search keys one branch on the sensor's NAME (:1182-1191, "HOKUYO_UTM_30LX") although no parameter
// table exists for it: renaming a ready matcher is the only way to execute that branch of the reference
void zhpsm_ref_set_laser_name(void *hv, const char *name) {
    Ref *h = (Ref *)hv;
    h->m->m_pParam->pm_laser_name = std::string(name);
}
void zhpsm_ref_destroy(void *hv) {
    Ref *h = (Ref *)hv;
    if (!h) return;
    delete h->m;
    delete h;
}
void zhpsm_ref_params(void *hv, int *n, float *fi_min, float *dfi, int *min_range, int *max_range, int *min_valid,
                      int *window) {
    Ref *h = (Ref *)hv;
    Base_PARAM *p = h->m->m_pParam;
    *n = p->pm_l_points; *fi_min = p->pm_fi_min; *dfi = p->pm_dfi; *min_range = p->pm_min_range;
    *max_range = p->pm_max_range; *min_valid = p->pm_min_valid_points; *window = p->pm_scan_window;
}
void zhpsm_ref_tables(void *hv, float *fi, float *si, float *co) {
    Ref *h = (Ref *)hv;
    memcpy(fi, h->m->pm_fi, sizeof(float) * h->n);
    memcpy(si, h->m->pm_si, sizeof(float) * h->n);
    memcpy(co, h->m->pm_co, sizeof(float) * h->n);
}

// pm_preprocessScan (:861-866): r (in/out, median filtered in place), bad (in/out), seg (out)
void zhpsm_ref_preprocess(void *hv, float *r, int *bad, int *seg) {
    Ref *h = (Ref *)hv;
    PMScan ls(h->n);
    fill_scan(h, ls, r, bad, 0, 0);
    h->m->pm_preprocessScan(&ls);
    memcpy(r, ls.r, sizeof(float) * h->n);
    memcpy(bad, ls.bad, sizeof(int) * h->n);
    memcpy(seg, ls.seg, sizeof(int) * h->n);
}
void zhpsm_ref_median(void *hv, float *r) {
    Ref *h = (Ref *)hv;
    PMScan ls(h->n);
    fill_scan(h, ls, r, 0, 0, 0);
    h->m->pm_median_filter(&ls);
    memcpy(r, ls.r, sizeof(float) * h->n);
}

// pm_scan_project (:1356-1479) of a preprocessed scan at pose (cm,cm,rad)
void zhpsm_ref_scan_project(void *hv, const float *r, const int *bad, const int *seg, const float pose[3],
                            float *new_r, int *new_bad) {
    Ref *h = (Ref *)hv;
    PMScan ls(h->n);
    fill_scan(h, ls, r, bad, seg, pose);
    h->m->pm_scan_project(&ls, new_r, new_bad);
}

// algo 0: pm_psm (:890-1003); 1: pm_icp (:1653-2021); 2: pm_psm then pm_icp seeded with its result (runlog :561-563)
// Scans must be preprocessed.  pose_io: act pose in/out (cm,cm,rad); ref pose = ref_pose.  Returns status.
int zhpsm_ref_match(void *hv, int algo, const float *ref_r, const int *ref_bad, const int *ref_seg,
                    const float ref_pose[3], const float *act_r, const int *act_bad, const int *act_seg,
                    float pose_io[3], float *ret_err) {
    Ref *h = (Ref *)hv;
    PMScan lsr(h->n), lsa(h->n);
    fill_scan(h, lsr, ref_r, ref_bad, ref_seg, ref_pose);
    fill_scan(h, lsa, act_r, act_bad, act_seg, pose_io);
    int status = 0;
    float e = 0;
    try {
        if (algo == 0) e = h->m->pm_psm(&lsr, &lsa);
        else if (algo == 1) e = h->m->pm_icp(&lsr, &lsa);
        else { h->m->pm_psm(&lsr, &lsa); e = h->m->pm_icp(&lsr, &lsa); }
    } catch (int err) {
        status = 1;
    }
    pose_io[0] = lsa.rx; pose_io[1] = lsa.ry; pose_io[2] = lsa.th;
    if (ret_err) *ret_err = e;
    return status;
}

// pm_error_index2 (:1279-1339); scans preprocessed, poses given
void zhpsm_ref_error_index2(void *hv, const float *ref_r, const int *ref_bad, const int *ref_seg,
                            const float ref_pose[3], const float *act_r, const int *act_bad, const int *act_seg,
                            const float act_pose[3], float out[2], int *assoc) {
    Ref *h = (Ref *)hv;
    PMScan lsr(h->n), lsa(h->n);
    fill_scan(h, lsr, ref_r, ref_bad, ref_seg, ref_pose);
    fill_scan(h, lsa, act_r, act_bad, act_seg, act_pose);
    std::pair<float, float> p = h->m->pm_error_index2(&lsr, &lsa, assoc);
    out[0] = p.first; out[1] = p.second;
}

// CPolarMatch::FMatch (:81-121) on RAW scans (it preprocesses lsa, and lsr when bFirst).
// r arrays are updated in place like the reference mutates its inputs.  out = returned pair.
void zhpsm_ref_fmatch(void *hv, float *ref_r, int *ref_bad, int *ref_seg, float *act_r, int *act_bad, int *act_seg,
                      int bFirst, float pose_out[3], float out[2]) {
    Ref *h = (Ref *)hv;
    PMScan lsr(h->n), lsa(h->n);
    fill_scan(h, lsr, ref_r, ref_bad, ref_seg, 0);
    fill_scan(h, lsa, act_r, act_bad, act_seg, 0);
    std::pair<float, float> p = h->m->FMatch(&lsr, &lsa, bFirst != 0);
    memcpy(ref_r, lsr.r, sizeof(float) * h->n); memcpy(ref_bad, lsr.bad, sizeof(int) * h->n); memcpy(ref_seg, lsr.seg, sizeof(int) * h->n);
    memcpy(act_r, lsa.r, sizeof(float) * h->n); memcpy(act_bad, lsa.bad, sizeof(int) * h->n); memcpy(act_seg, lsa.seg, sizeof(int) * h->n);
    pose_out[0] = lsa.rx; pose_out[1] = lsa.ry; pose_out[2] = lsa.th;
    out[0] = p.first; out[1] = p.second;
}

int zhpsm_ref_is_corridor(void *hv, const float *r, const int *bad, const int *seg) {
    Ref *h = (Ref *)hv;
    PMScan ls(h->n);
    fill_scan(h, ls, r, bad, seg, 0);
    try { return h->m->pm_is_corridor(&ls) ? 1 : 0; } catch (int e) { return -e; }
}

// Batch for the timed CPU baseline: preprocessed scans [S][n] (r,bad,seg), pairs (ref_idx,act_idx), zero poses.
// One CPolarMatch per thread (pm_* are re-entrant on a const instance, but keep the reference's
// "one matcher per thread" discipline).  poses_out [npairs][3], status_out [npairs].
struct BatchJob { const char *laser; float fi_min; int max_range; int algo; const float *r; const int *bad; const int *seg;
                  int n; const int *ri; const int *ai; int lo, hi; float *poses; int *status; };
static void *batch_worker(void *a) {
    BatchJob *j = (BatchJob *)a;
    void *h = zhpsm_ref_create(j->laser, j->fi_min, j->max_range);
    float zero[3] = {0, 0, 0};
    for (int k = j->lo; k < j->hi; k++) {
        float pose[3] = {0, 0, 0};
        size_t ro = (size_t)j->ri[k] * j->n, ao = (size_t)j->ai[k] * j->n;
        j->status[k] = zhpsm_ref_match(h, j->algo, j->r + ro, j->bad + ro, j->seg + ro, zero, j->r + ao, j->bad + ao,
                                       j->seg + ao, pose, 0);
        memcpy(j->poses + 3 * (size_t)k, pose, sizeof(pose));
    }
    zhpsm_ref_destroy(h);
    return 0;
}
void zhpsm_ref_match_batch(const char *laser, float fi_min, int max_range, int algo, const float *r, const int *bad,
                           const int *seg, int n, const int *ri, const int *ai, int npairs, float *poses, int *status,
                           int nthreads) {
    silence();
    // the reference's laser-name table is a lazily filled function-static map (ZHPolar_Match.cpp:15-27):
    // touch it once on this thread before workers race on it
    zhpsm_ref_destroy(zhpsm_ref_create(laser, fi_min, max_range));
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    BatchJob jobs[256];
    for (int t = 0; t < nthreads; t++) {
        BatchJob jb = {laser, fi_min, max_range, algo, r, bad, seg, n, ri, ai, (int)((long long)npairs * t / nthreads),
                       (int)((long long)npairs * (t + 1) / nthreads), poses, status};
        jobs[t] = jb;
    }
    if (nthreads == 1) { batch_worker(&jobs[0]); return; }
    for (int t = 0; t < nthreads; t++) pthread_create(&th[t], 0, batch_worker, &jobs[t]);
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], 0);
}

// OrientedPoint2D::oplus(Point2D) / ominus(OrientedPoint2D) of the reference (src/zhpsm/point.cpp:46-69), for pinning the
// restatement in oracle/graph_oracle.c
void zhpsm_ref_oplus_point(const double pose[3], const double pt[2], double out[2]) {
    zhpsm::OrientedPoint2D p(pose[0], pose[1], pose[2]);
    zhpsm::Point2D r = p.oplus(zhpsm::Point2D(pt[0], pt[1]));
    out[0] = r.x; out[1] = r.y;
}
void zhpsm_ref_ominus_pose(const double a[3], const double b[3], double out[3]) {
    zhpsm::OrientedPoint2D p(a[0], a[1], a[2]);
    zhpsm::OrientedPoint2D r = p.ominus(zhpsm::OrientedPoint2D(b[0], b[1], b[2]));
    out[0] = r.x; out[1] = r.y; out[2] = r.theta;
}

// CPolarMatch::readCarmon (ZHPolar_Match.cpp:172-284) executed by the reference: scans [cap][n] float cm, poses [cap][3],
// geometry after the header override (fi_min, max_range).  Returns the number of scans read (-1 on failure).
int zhpsm_ref_read_carmen(const char *laser_name, const char *file, const char *tag, float *scans, float *poses, int cap,
                          float *fi_min, int *max_range) {
    silence();
    CPolarMatch m(laser_name);
    if (!m.m_bReady) return -1;
    if (!m.readCarmon(file, tag)) return -1;
    const int n = m.m_pParam->pm_l_points, S = (int)m.m_SickScans.size();
    for (int k = 0; k < S && k < cap; k++) {
        memcpy(scans + (size_t)k * n, m.m_SickScans[k]->r, sizeof(float) * n);
        poses[3 * k] = m.m_SickScans[k]->rx; poses[3 * k + 1] = m.m_SickScans[k]->ry; poses[3 * k + 2] = m.m_SickScans[k]->th;
    }
    *fi_min = m.m_pParam->pm_fi_min;
    *max_range = m.m_pParam->pm_max_range;
    return S;
}
}
