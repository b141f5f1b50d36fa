// laser_slam_b200/host/zh_carmen.h -- CARMEN log reader of the reference (CPolarMatch::readCarmon,
// src/zhpsm/ZHPolar_Match.cpp:172-284; readFlirt :286-289) as a host-side loader that fills a scan pool directly:
// SURVEY 8f-4, the data format on the input side of the matchers.  Text parsing stays on the host; the pool it
// produces is what zhsm_preprocess_batch / zhsm_plicp_batch / zhsm_psm_batch take.
#ifndef ZHSM_HOST_ZH_CARMEN_H
#define ZHSM_HOST_ZH_CARMEN_H
#include <string>
#include <vector>

#include "../../include/zhsm.h"
#include "zh_scan.h"

struct ZhCarmenLog {
    std::vector<float> r;       // scan pool [S][n], cm, exactly the PMScan::r the reference builds (:250-279)
    std::vector<double> pose;   // [S][3] laser pose fields of the record (0 for "LMS*" logs, which carry none)
    int S, n;
    ZhCarmenLog() : S(0), n(0) {}
};

// Reads every record whose tag is `laser` ("ROBOTLASER1", "FLASER" ... ; a name starting with "LMS" selects the
// reference's own log flavour whose tag is the laser's name and which has no pose fields).  Like readCarmon the FIRST
// record overwrites param->pm_fi_min / pm_fi_max / pm_max_range from the log header (:227-235).  Readings below
// PM_MIN_RANGE and the padded last beam of (n-1)-beam records become pm_max_range + 1.  Returns false when the file
// cannot be opened, a record has an unexpected beam count ("ERROR READING DATA!") or no record matched.
bool zh_read_carmen(const std::string &filename, std::string laser, Base_PARAM *param, ZhCarmenLog *out);

extern "C" {
// C entry for bindings: laser_name = "LMS511" / "LMS211" / "LMS151" (the parameter table), tag as above.
// scans [cap][n] / poses [cap][3] may be NULL to only count.  geom receives the geometry after the header override.
// Returns the number of scans in the file (-1 on failure).
int zhsm_read_carmen(const char *filename, const char *laser_name, const char *tag, float *scans, double *poses, int cap,
                     zhsm_laser *geom);
}
#endif
