// laser_slam_b200/host/zh_scan.h -- the scan and sensor-geometry types at the matcher boundary.
//
// The reference passes scans as `PMScan` and geometry as `Base_PARAM` (src/zhpsm/PolarParameter.h:42-184).  Inside the
// reference tree include ITS header first (it defines POLAR_PARAMETER) and these fall away; stand-alone users of the
// shim get layout- and name-compatible minimal versions: only the members the matchers touch.
#ifndef ZH_SCAN_H
#define ZH_SCAN_H
#ifndef POLAR_PARAMETER
#include <string.h>

#include <string>

#define ZH_PM_PI 3.141592654  // the reference's truncated pi (PolarParameter.h:38); all its geometry derives from it
#define ZH_PM_D2R (ZH_PM_PI / 180.0)

struct _PARAM {
    std::string pm_laser_name;
    int pm_l_points;          // beams per scan
    int pm_fov;               // [deg]
    int pm_max_range;         // [cm]
    int pm_min_range;         // [cm]
    int pm_min_valid_points;
    int pm_scan_window;
    float pm_corridor_threshold;
    float pm_fi_min, pm_fi_max;  // [rad]
    float pm_dfi;                // [rad]
    _PARAM(std::string name, int points, int fov, int max_range, int valid_points, int scan_window, int corridor,
           int fi_min_deg, int fi_max_deg)
        : pm_laser_name(name), pm_l_points(points), pm_fov(fov), pm_max_range(max_range), pm_min_range(10),
          pm_min_valid_points(valid_points), pm_scan_window(scan_window), pm_corridor_threshold((float)corridor),
          pm_fi_min((float)(fi_min_deg * ZH_PM_D2R)), pm_fi_max((float)(fi_max_deg * ZH_PM_D2R)) {
        pm_dfi = (float)((pm_fov * ZH_PM_D2R) / (pm_l_points - 1.0));
    }
    virtual ~_PARAM() {}
};
typedef _PARAM Base_PARAM;
struct PM_SICK_LMS511 : Base_PARAM { PM_SICK_LMS511() : Base_PARAM("LMS511", 361, 180, 5000, 80, 40, 25, 0, 180) {} };
struct PM_SICK_LMS211 : Base_PARAM { PM_SICK_LMS211() : Base_PARAM("LMS211", 181, 180, 5000, 40, 20, 25, -90, 90) {} };
struct PM_SICK_LMS151 : Base_PARAM { PM_SICK_LMS151() : Base_PARAM("LMS151", 541, 270, 5000, 100, 50, 25, -45, 225) {} };

struct _PMScan {
    double t;
    float rx, ry, th;  // [cm, cm, rad]: initial pose in, matched pose out
    int np;
    float *r, *x, *y;  // [cm]
    int *bad, *seg;
    explicit _PMScan(int n) : t(0), rx(0), ry(0), th(0), np(n) {
        r = new float[n]; x = new float[n]; y = new float[n]; bad = new int[n]; seg = new int[n];
        memset(bad, 0, sizeof(int) * n); memset(seg, 0, sizeof(int) * n);
    }
    _PMScan(const _PMScan &o) : t(o.t), rx(o.rx), ry(o.ry), th(o.th), np(o.np) {
        r = new float[np]; x = new float[np]; y = new float[np]; bad = new int[np]; seg = new int[np];
        memcpy(r, o.r, sizeof(float) * np); memcpy(x, o.x, sizeof(float) * np); memcpy(y, o.y, sizeof(float) * np);
        memcpy(bad, o.bad, sizeof(int) * np); memcpy(seg, o.seg, sizeof(int) * np);
    }
    ~_PMScan() { delete[] r; delete[] x; delete[] y; delete[] bad; delete[] seg; }
private:
    _PMScan &operator=(const _PMScan &);
};
typedef _PMScan PMScan;
#endif  // POLAR_PARAMETER
#endif  // ZH_SCAN_H
