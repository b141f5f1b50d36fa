// laser_slam_b200/host/ZHPolar_Match.h -- C++ host shim with the reference's polar matcher interface (class
// CPolarMatch, src/zhpsm/ZHPolar_Match.h:16-88) on top of the C ABI in include/zhsm.h: the matching entry points only
// (the log readers / run loops are file IO and stay with the reference).
//
// Same names, argument meaning, side effects and error behaviour:
//   CPolarMatch(laser_name)      "LMS511" / "LMS211" / "LMS151" (getParam, ZHPolar_Match.cpp:14-27); m_bReady false
//                                for an unknown name
//   pm_init()                    re-reads m_pParam (call after changing it, as readCarmon does, :227-235)
//   pm_preprocessScan(ls)        median filter ls->r in place, ls->bad |= PM_RANGE, ls->seg (:861-866)
//   pm_psm(lsr, lsa) / pm_icp    match preprocessed scans; the result replaces lsa->rx, ry [cm], th [rad]; the return
//                                value is the reference's average error; failures `throw 1` like the reference
//                                (:1095,1106,1239,1831) and leave lsa untouched
//   FMatch(lsr, lsa, bFirst)     preprocess (lsr too when bFirst), zero both poses, pm_icp, pm_error_index2;
//                                (-1,-1) on failure (:81-121)
//   pm_scan_project(act, new_r, new_bad)   (:1356-1479)
//   pm_is_corridor(act)          (:2058-2186)
// and FMatchBatch(): many independent pairs in one launch.
// There is no CPU path: construction throws std::runtime_error when no sm_100 GPU / libzhsm.so is usable.
#ifndef ZHSM_HOST_ZHPOLAR_MATCH_H
#define ZHSM_HOST_ZHPOLAR_MATCH_H
#include <string>
#include <utility>
#include <vector>

#include "../../include/zhsm.h"
#include "zh_scan.h"
#include "zh_carmen.h"

class CPolarMatch {
public:
    explicit CPolarMatch(std::string laser_name, int device = 0);
    ~CPolarMatch();
    std::pair<float, float> FMatch(PMScan *lsr, PMScan *lsa, bool bFirst = false);
    void pm_init();
    void pm_preprocessScan(PMScan *ls);
    void pm_median_filter(PMScan *ls);
    void pm_scan_project(const PMScan *act, float *new_r, int *new_bad);
    bool pm_is_corridor(PMScan *act);  // throws 3 / 2 like the reference when fewer than two line directions exist
    float pm_psm(const PMScan *lsr, PMScan *lsa);
    float pm_icp(const PMScan *lsr, PMScan *lsa);
    // pairs (refs[k], acts[k]); preprocesses acts (and refs when preprocess_refs); out[k] = FMatch's return value
    void FMatchBatch(PMScan **refs, PMScan **acts, int B, bool preprocess_refs, std::pair<float, float> *out);

    // log input (:172-289): fills m_SickScans and overrides the geometry from the log header, then pm_init()
    bool readCarmon(std::string filename, std::string laser = "");
    bool readFlirt(std::string filename) { return readCarmon(filename, "ROBOTLASER1"); }
    void clearRecord();

    Base_PARAM *m_pParam;
    bool m_bReady;
    std::vector<PMScan *> m_SickScans;

private:
    zhsm_handle *h_;
    float match(int algo, const PMScan *lsr, PMScan *lsa);
    CPolarMatch(const CPolarMatch &);
    CPolarMatch &operator=(const CPolarMatch &);
};
#endif
