// laser_slam_b200/host/ZHCanonical_Matcher.h -- C++ host shim with the reference's PL-ICP matcher interface
// (class CCanonicalMatcher, src/zhcsm/ZHCanonical_Matcher.h:20-47) on top of the C ABI (include/zhsm.h and the
// CSM-shaped entry include/csm_compat.h).
//
// Same member names, argument meaning, side effects and return codes as the reference:
//   input / output             public `sm_params*` / `sm_result*` exactly like the reference's: callers tune
//                              `matcher.input->max_iterations` etc. and read `matcher.output->x`, `->cov_x_m`
//   fromPMScan2LDP             cm -> m, validity, theta[i] = fi_min + dfi*i in float, zero poses (:193-227)
//   FMatch(lsr, lsa, bFirst)   median-filters lsa->r (and lsr->r when bFirst) IN PLACE, matches, writes the result
//                              into lsa->rx, ry [cm], th [rad]; returns 0, or -1 when sm_icp reports valid == 0 (:229-249)
//   FMatch(LDP, LDP)           zeroes the ref poses and first_guess, sm_icp, copies x into lsa->true_pose (:251-285)
//   FMatchKeyFrame(PMScan*)    returns 2 for the very first scan, afterwards matches against the previous scan (:300-313)
//   FMatchKeyFrame(LDP)        ... and always replaces (and frees) the key frame with the current scan (:315-370)
//   FMatchKeyFrame2(lsa, cov)  + the 3x3 covariance of output->cov_x_m (:287-298); setCov(true) must be on
//   setCov / resetKeyFrame / setThreshold / newKeyframeNeeded as in the reference
// and one addition the reference cannot offer: FMatchBatch() matches many independent pairs in one launch.
// There is no CPU path: construction throws std::runtime_error when no sm_100 GPU / libzhsm.so is usable.
#ifndef ZHSM_HOST_ZHCANONICAL_MATCHER_H
#define ZHSM_HOST_ZHCANONICAL_MATCHER_H
#include <vector>

#include "../../include/csm_compat.h"
#include "../../include/zhsm.h"
#include "zh_scan.h"

class CCanonicalMatcher {
public:
    explicit CCanonicalMatcher(Base_PARAM *p = 0, int device = 0);
    ~CCanonicalMatcher();
    void sm_init(Base_PARAM *p);
    void setCov(bool on);
    void fromPMScan2LDP(PMScan *pm_scan, LDP *ldp, Base_PARAM *param);
    void median_filter(PMScan *ls);
    float FMatch(PMScan *lsr, PMScan *lsa, bool bFirst = false);
    float FMatch(LDP lsr, LDP lsa);
    void resetKeyFrame(PMScan *lsa);
    float FMatchKeyFrame(PMScan *lsa);
    float FMatchKeyFrame(LDP lsa);
    float FMatchKeyFrame2(PMScan *lsa, double cov[9]);
    template <class Matrix3>
    float FMatchKeyFrame2(PMScan *lsa, Matrix3 &cor) {  // Eigen::Matrix3d in the reference
        double c[9];
        float ret = FMatchKeyFrame2(lsa, c);
        if (ret < 0 || ret == 2) return ret;
        for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) cor(i, j) = c[3 * i + j];
        return ret;
    }
    bool newKeyframeNeeded(double pose[3]);
    void setThreshold(float dist, float angle);
    // many independent pairs at once: refs[k] vs acts[k]; same side effects per pair as FMatch(.., bFirst=filter_refs)
    void FMatchBatch(PMScan **refs, PMScan **acts, int B, bool filter_refs, float *ret);

    LDP m_pKeyframe;
    float threshold_angular_, threshold_dist_sq_;
    sm_params *input;
    sm_result *output;
    Base_PARAM *m_pParam;

private:
    zhsm_handle *h_;
    void drop_cov();
    void set_pm_laser();
    zhsm_plicp_params engine_params() const;
    CCanonicalMatcher &operator=(const CCanonicalMatcher &);
    CCanonicalMatcher(const CCanonicalMatcher &);
};
#endif
