// laser_slam_b200/host/ZHIcp_Warpper.h -- C++ host shim with the reference's ICP wrapper interface (class CICPWarpper,
// src/zhicp/ZHIcp_Warpper.h:20-45) on top of the C ABI in include/zhsm.h.  No MRPT types appear: the reference's
// public MRPT members (m_pICP, m_pTrans, m_pPose) become `options`, and the result record read by getResult /
// getMatchQuality.  Same method names, argument order and meaning:
//   ICPMatch(rx, ry, N1, ax, ay, N2, pini, pout)   aligns the current cloud (ax, ay) to the map cloud (rx, ry) from
//                                                  pini; writes the pose to pout; returns goodness (:125-142)
//   ICPMatch(rx, ry, ax, ay, N, pini, pout)        same with equal sizes (:144-161)
//   ICPMatch(vector rx, ry, ax, ay, px, py, pth)   same, result only through getResult / getMatchQuality (:163-176)
//   getResult(p, cov)                              pose + upper triangle of the covariance (:238-244)
//   getMatchQuality(iterations, goodness, quality) (:245-250)
//   setDebugWindowOn/Off                           kept; there is no plot window on a GPU box
// plus ICPMatchMultiStart(): G initial guesses for one pair in one launch, best goodness wins (BASELINE config 3).
// There is no CPU path: construction throws std::runtime_error when no sm_100 GPU / libzhsm.so is usable.
#ifndef ZHSM_HOST_ZHICP_WARPPER_H
#define ZHSM_HOST_ZHICP_WARPPER_H
#include <vector>

#include "../../include/zhsm.h"

class CICPWarpper {
public:
    explicit CICPWarpper(int device = 0);
    ~CICPWarpper();
    void initICP();
    float ICPMatch(float *rx, float *ry, float *ax, float *ay, int N, double pini[3], double pout[3]);
    float ICPMatch(float *rx, float *ry, int N1, float *ax, float *ay, int N2, double pini[3], double pout[3]);
    void ICPMatch(std::vector<float> &rx, std::vector<float> &ry, std::vector<float> &ax, std::vector<float> &ay,
                  double px, double py, double pth);
    // G guesses pini[G][3]; returns the best goodness, pout = its pose; results (optional) receives all G records
    float ICPMatchMultiStart(const float *rx, const float *ry, int N1, const float *ax, const float *ay, int N2,
                             const double *pini, int G, double pout[3], zhsm_result *results = 0);
    void getResult(double p[3], double cov[6]);
    void getMatchQuality(int &iterations, float &good, float &quality);
    void setDebugWindowOn() { skip_window = false; }
    void setDebugWindowOff() { skip_window = true; }

    bool skip_window;
    int ICP_method;             // 0 = icpClassic, the only algorithm the reference selects (ZHIcp_Warpper.cpp:18)
    zhsm_icp_options options;   // the reference's m_pICP->options
    zhsm_result last;           // the reference's m_pPose / m_pTrans / TReturnInfo

private:
    zhsm_handle *h_;
    CICPWarpper(const CICPWarpper &);
    CICPWarpper &operator=(const CICPWarpper &);
};
#endif
