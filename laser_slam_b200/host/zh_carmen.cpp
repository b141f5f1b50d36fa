// laser_slam_b200/host/zh_carmen.cpp -- see zh_carmen.h.
#include "zh_carmen.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <fstream>

namespace {
// next blank-separated field of a record, or NULL at the end of the line
struct Fields {
    char *p;
    explicit Fields(char *line) : p(line) {}
    const char *next() {
        while (*p == ' ') p++;
        if (!*p) return 0;
        char *s = p;
        while (*p && *p != ' ') p++;
        if (*p) *p++ = 0;
        return s;
    }
    double num() { const char *s = next(); return s ? atof(s) : 0.0; }
};
}  // namespace

bool zh_read_carmen(const std::string &filename, std::string laser, Base_PARAM *param, ZhCarmenLog *out) {
    std::ifstream inf(filename.c_str());
    if (!inf.is_open()) { printf("failed to open file: %s\n", filename.c_str()); return false; }
    const int n = param->pm_l_points;
    out->r.clear(); out->pose.clear(); out->S = 0; out->n = n;
    bool ours = false, first = true;
    if (laser.compare(0, 3, "LMS") == 0) { laser = param->pm_laser_name; ours = true; }
    std::string line;
    std::vector<char> buf;
    std::vector<float> r(n);
    while (std::getline(inf, line)) {
        buf.assign(line.begin(), line.end());
        buf.push_back(0);
        Fields f(&buf[0]);
        const char *tag = f.next();
        if (!tag || laser != tag) continue;
        f.num();                                   // laser type
        const double start = f.num(), fov = f.num();
        f.num();                                   // resolution (the reference ignores it, :230)
        const double maxRange = f.num();
        f.num();                                   // accuracy
        f.num();                                   // remission mode
        const int num_points = (int)f.num();
        if (num_points == 0) continue;
        if (first) {
            first = false;
            param->pm_fi_min = (float)start;
            param->pm_fi_max = (float)(start + fov * ZH_PM_D2R);
            if (maxRange * 100 != param->pm_max_range) param->pm_max_range = (int)(maxRange * 100);
        }
        if (num_points != n - 1 && num_points != n) { printf("ERROR READING DATA!\n"); return false; }
        for (int i = 0; i < num_points; i++) {
            const char *s = f.next();
            if (!s) { printf("not enough points in file!\n"); return false; }
            r[i] = (float)(atof(s) * 100.0);      // [m] -> [cm]
            if (r[i] < 10.0f) r[i] = (float)(param->pm_max_range + 1);   // PM_MIN_RANGE
        }
        double pose[3] = {0, 0, 0};
        if (!ours) {
            f.next();                              // remission count
            pose[0] = f.num(); pose[1] = f.num(); pose[2] = f.num();
        }
        if (num_points == n - 1) r[num_points] = (float)(param->pm_max_range + 1);
        out->r.insert(out->r.end(), r.begin(), r.end());
        out->pose.insert(out->pose.end(), pose, pose + 3);
        out->S++;
    }
    printf("Scan size: %d\n", out->S);
    return out->S > 0;
}

extern "C" int zhsm_read_carmen(const char *filename, const char *laser_name, const char *tag, float *scans, double *poses,
                                int cap, zhsm_laser *geom) {
    if (!filename || !laser_name || !tag) return -1;
    Base_PARAM *p = 0;
    if (!strcmp(laser_name, "LMS511")) p = new PM_SICK_LMS511();
    else if (!strcmp(laser_name, "LMS211")) p = new PM_SICK_LMS211();
    else if (!strcmp(laser_name, "LMS151")) p = new PM_SICK_LMS151();
    else return -1;
    ZhCarmenLog log;
    const bool ok = zh_read_carmen(filename, tag, p, &log);
    if (geom) {
        memset(geom, 0, sizeof(*geom));
        geom->n = p->pm_l_points; geom->fi_min = p->pm_fi_min; geom->dfi = p->pm_dfi;
        geom->min_range_cm = p->pm_min_range; geom->max_range_cm = p->pm_max_range;
        geom->min_valid_points = p->pm_min_valid_points; geom->scan_window = p->pm_scan_window;
    }
    delete p;
    if (!ok) return -1;
    const int ncopy = log.S < cap ? log.S : cap;
    if (scans && ncopy > 0) memcpy(scans, &log.r[0], sizeof(float) * (size_t)ncopy * log.n);
    if (poses && ncopy > 0) memcpy(poses, &log.pose[0], sizeof(double) * 3 * (size_t)ncopy);
    return log.S;
}
