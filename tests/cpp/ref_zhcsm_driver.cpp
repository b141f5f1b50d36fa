// tests/cpp/ref_zhcsm_driver.cpp -- drives the REFERENCE'S OWN CCanonicalMatcher (compiled unmodified from
// /root/reference/src/zhcsm/ZHCanonical_Matcher.cpp against -I<repo>/include, where <csm/csm_all.h> forwards to
// include/csm_compat.h) linked against libzhsm.so instead of libcsm: "link, change nothing" (INTEGRATION.md option C).
//   ref_zhcsm_on_zhsm <laser> <fi_min> <max_range_cm> <scans.f32> <S> <mode> <out.txt>
// mode 0: FMatch(prev, cur, bFirst = first pair)      -> "ret rx ry th"
// mode 1: FMatchKeyFrame(cur)                         -> "ret rx ry th"
// mode 2: setCov(true); FMatchKeyFrame2(cur, Matrix3d) -> "ret rx ry th c00 c01 c02 c11 c12 c22"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "PolarParameter.h"        // the reference's (src/zhpsm)
#include "ZHCanonical_Matcher.h"   // the reference's (src/zhcsm)

int main(int argc, char **argv) {
    if (argc < 8) return 2;
    Base_PARAM *p = 0;
    if (!strcmp(argv[1], "LMS211")) p = new PM_SICK_LMS211();
    else if (!strcmp(argv[1], "LMS511")) p = new PM_SICK_LMS511();
    else if (!strcmp(argv[1], "LMS151")) p = new PM_SICK_LMS151();
    else return 2;
    p->pm_fi_min = (float)atof(argv[2]);  // readCarmon overwrites these from the log header (ZHPolar_Match.cpp:227-235)
    p->pm_max_range = atoi(argv[3]);
    const int S = atoi(argv[5]), mode = atoi(argv[6]), n = p->pm_l_points;
    std::vector<float> all((size_t)S * n);
    FILE *f = fopen(argv[4], "rb");
    if (!f || fread(&all[0], sizeof(float), all.size(), f) != all.size()) return 3;
    fclose(f);
    FILE *out = fopen(argv[7], "w");
    if (!out) return 3;
    std::vector<PMScan *> scans(S);
    for (int k = 0; k < S; k++) {
        scans[k] = new PMScan(n);
        memcpy(scans[k]->r, &all[(size_t)k * n], sizeof(float) * n);
    }
    CCanonicalMatcher m(p);
    if (mode == 0) {
        for (int k = 1; k < S; k++) {
            float ret = m.FMatch(scans[k - 1], scans[k], k == 1);
            fprintf(out, "%g %.9g %.9g %.9g\n", ret, scans[k]->rx, scans[k]->ry, scans[k]->th);
        }
    } else if (mode == 1) {
        for (int k = 0; k < S; k++) {
            float ret = m.FMatchKeyFrame(scans[k]);
            fprintf(out, "%g %.9g %.9g %.9g\n", ret, scans[k]->rx, scans[k]->ry, scans[k]->th);
        }
    } else {
        m.setCov(true);
        for (int k = 0; k < S; k++) {
            Eigen::Matrix3d c;
            float ret = m.FMatchKeyFrame2(scans[k], c);
            fprintf(out, "%g %.9g %.9g %.9g %.17g %.17g %.17g %.17g %.17g %.17g\n", ret, scans[k]->rx, scans[k]->ry,
                    scans[k]->th, c(0, 0), c(0, 1), c(0, 2), c(1, 1), c(1, 2), c(2, 2));
        }
    }
    fclose(out);
    return 0;
}
