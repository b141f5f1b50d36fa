// laser_slam_b200/host/ZHCanonical_Matcher.h -- C++ host shim with the reference's PL-ICP matcher interface
// (class CCanonicalMatcher, src/zhcsm/ZHCanonical_Matcher.h:20-47) on top of the C ABI in include/zhsm.h.
//
// Same method names, argument meaning, side effects and return codes as the reference:
//   FMatch(lsr, lsa, bFirst)   median-filters lsa->r (and lsr->r when bFirst) IN PLACE, matches, writes the result
//                              into lsa->rx, ry [cm], th [rad]; returns 0, or -1 when sm_icp reports valid == 0
//                              (ZHCanonical_Matcher.cpp:229-285)
//   FMatchKeyFrame(lsa)        returns 2 for the very first scan, afterwards matches against the previous scan and
//                              always replaces the key frame with the current scan (:300-370)
//   FMatchKeyFrame2(lsa, cov)  + the 3x3 covariance of sm_result.cov_x_m (:287-298); setCov(true) must be on
//   setCov / resetKeyFrame / setThreshold as in the reference
// and one addition the reference cannot offer: FMatchBatch() matches many independent pairs in one launch.
// There is no CPU path: construction throws std::runtime_error when no sm_100 GPU / libzhsm.so is usable.
#ifndef ZHSM_HOST_ZHCANONICAL_MATCHER_H
#define ZHSM_HOST_ZHCANONICAL_MATCHER_H
#include <vector>

#include "../../include/zhsm.h"
#include "zh_scan.h"

class CCanonicalMatcher {
public:
    explicit CCanonicalMatcher(Base_PARAM *p = 0, int device = 0);
    ~CCanonicalMatcher();
    void sm_init(Base_PARAM *p);
    void setCov(bool on);
    void median_filter(PMScan *ls);
    float FMatch(PMScan *lsr, PMScan *lsa, bool bFirst = false);
    float FMatchKeyFrame(PMScan *lsa);
    float FMatchKeyFrame2(PMScan *lsa, double cov[9]);
    template <class Matrix3>
    float FMatchKeyFrame2(PMScan *lsa, Matrix3 &cor) {  // Eigen::Matrix3d in the reference
        double c[9];
        float ret = FMatchKeyFrame2(lsa, c);
        if (ret < 0 || ret == 2) return ret;
        for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) cor(i, j) = c[3 * i + j];
        return ret;
    }
    void resetKeyFrame(PMScan *lsa);
    void setThreshold(float dist, float angle);
    // many independent pairs at once: refs[k] vs acts[k]; same side effects per pair as FMatch(.., bFirst=filter_refs)
    void FMatchBatch(PMScan **refs, PMScan **acts, int B, bool filter_refs, float *ret);

    zhsm_plicp_params input;   // the reference's public `sm_params* input`
    zhsm_result output;        // the reference's public `sm_result* output` (last single-pair result)
    Base_PARAM *m_pParam;
    float threshold_angular_, threshold_dist_sq_;

private:
    zhsm_handle *h_;
    std::vector<float> keyframe_;  // filtered ranges of the key frame (the reference keeps an LDP)
    bool have_keyframe_;
    float match_filtered(const float *ref_r, const float *sens_r, PMScan *lsa);
    CCanonicalMatcher &operator=(const CCanonicalMatcher &);
    CCanonicalMatcher(const CCanonicalMatcher &);
};
#endif
