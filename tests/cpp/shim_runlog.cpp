// tests/cpp/shim_runlog.cpp -- drives the C++ shim the way the reference's own drivers drive CCanonicalMatcher
// (runlog-style consecutive-scan odometry, src/zhcsm/ZHCanonical_Matcher.cpp runlog / testKeyFrameMatch):
//   shim_runlog <laser> <fi_min> <max_range_cm> <scans.f32> <S> <mode> <out.txt>
// mode 0: FMatch(prev, cur, bFirst = first pair)   mode 1: FMatchKeyFrame(cur)   mode 2: FMatchBatch
// Output: one line per scan: ret rx ry th
// CPolarMatch (src/zhpsm/ZHPolar_Match.cpp testAccuracy / runlog style):
// mode 3: FMatch(prev, cur, bFirst = first pair): "ex ey rx ry th"      mode 4: FMatchBatch over copies
// mode 5: preprocess all, then per pair pm_psm from zero ("status err rx ry th", status 1 = threw) and pm_icp
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "ZHCanonical_Matcher.h"
#include "ZHPolar_Match.h"

int main(int argc, char **argv) {
    if (argc < 8) return 2;
    Base_PARAM *p = 0;
    if (!strcmp(argv[1], "LMS211")) p = new PM_SICK_LMS211();
    else if (!strcmp(argv[1], "LMS511")) p = new PM_SICK_LMS511();
    else if (!strcmp(argv[1], "LMS151")) p = new PM_SICK_LMS151();
    else return 2;
    p->pm_fi_min = (float)atof(argv[2]);     // readCarmon overwrites these from the log header (ZHPolar_Match.cpp:227-235)
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
    if (mode >= 3) {
        CPolarMatch pm(argv[1]);
        if (!pm.m_bReady) return 2;
        pm.m_pParam->pm_fi_min = p->pm_fi_min;
        pm.m_pParam->pm_max_range = p->pm_max_range;
        pm.pm_init();
        if (mode == 3) {
            for (int k = 1; k < S; k++) {
                std::pair<float, float> e = pm.FMatch(scans[k - 1], scans[k], k == 1);
                fprintf(out, "%.9g %.9g %.9g %.9g %.9g\n", e.first, e.second, scans[k]->rx, scans[k]->ry, scans[k]->th);
            }
        } else if (mode == 4) {
            std::vector<PMScan *> refs, acts;
            for (int k = 0; k + 1 < S; k++) { refs.push_back(new PMScan(*scans[k])); acts.push_back(new PMScan(*scans[k + 1])); }
            std::vector<std::pair<float, float> > e(refs.size());
            pm.FMatchBatch(&refs[0], &acts[0], (int)refs.size(), true, &e[0]);
            for (size_t k = 0; k < refs.size(); k++)
                fprintf(out, "%.9g %.9g %.9g %.9g %.9g\n", e[k].first, e[k].second, acts[k]->rx, acts[k]->ry, acts[k]->th);
        } else {
            for (int k = 0; k < S; k++) pm.pm_preprocessScan(scans[k]);
            for (int k = 1; k < S; k++) {
                for (int alg = 0; alg < 2; alg++) {
                    PMScan ref(*scans[k - 1]), act(*scans[k]);
                    ref.rx = ref.ry = ref.th = 0; act.rx = act.ry = act.th = 0;
                    int threw = 0;
                    float err = 0;
                    try { err = alg == 0 ? pm.pm_psm(&ref, &act) : pm.pm_icp(&ref, &act); } catch (int) { threw = 1; }
                    fprintf(out, "%d %.9g %.9g %.9g %.9g%s", threw, err, act.rx, act.ry, act.th, alg == 0 ? " " : "\n");
                }
            }
        }
        fclose(out);
        delete p;
        return 0;
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
        // batch over independent copies: pair k = (scan k, scan k+1), both filtered like bFirst = true
        std::vector<PMScan *> refs, acts;
        for (int k = 0; k + 1 < S; k++) { refs.push_back(new PMScan(*scans[k])); acts.push_back(new PMScan(*scans[k + 1])); }
        std::vector<float> ret(refs.size());
        m.FMatchBatch(&refs[0], &acts[0], (int)refs.size(), true, &ret[0]);
        for (size_t k = 0; k < refs.size(); k++) fprintf(out, "%g %.9g %.9g %.9g\n", ret[k], acts[k]->rx, acts[k]->ry, acts[k]->th);
    }
    fclose(out);
    delete p;
    return 0;
}
