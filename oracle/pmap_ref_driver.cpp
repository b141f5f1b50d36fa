// oracle/pmap_ref_driver.cpp -- TEST INFRASTRUCTURE ONLY.
// extern "C" driver around the reference's OWN occupancy-map sources src/mapGraph/PMap.cpp and line.cpp (compiled
// unmodified where they lie, see Makefile) = CPMap::initializeMap / updateMap / computeProbsOfMap / getPointCloud and
// the Bresenham grid_line.  The few lines of CMapNode that feed it (MapNode.cpp, which needs g2o/FLIRT and cannot be
// compiled) are restated here: translate2GlobalFrame :387-460, computeBoundary :714-725, computePMAP :726-757.
#include <cmath>
#include <vector>

#include "PMap.h"

extern "C" {

// scans [K][n] float cm; traj [K][3] double (x, y, theta of each scan's pose in the node frame); geometry as floats
// (Angle[i] = minAngle + dfi*i, MapNode.cpp:421-427); readings above max_range_cm are skipped (:446).
// Returns the number of cloud points; writes at most cap of them.  dims (optional): size_x, size_y, hits.
int pmap_ref_submap_cloud(const float *scans, const double *traj, int K, int n, float min_angle, float dfi,
                          int max_range_cm, float *out_x, float *out_y, int cap, int *dims) {
    static bool warmed = false;
    if (!warmed) {
        // CPMap::updateMap keeps a function-static scratch line sized by the FIRST map it ever sees (PMap.cpp:63) and
        // writes past it for bigger maps; give it the largest size once
        warmed = true;
        CPMap big;
        big.initializeMap(8, 8, 4, 4, 0, 0, 0.05);
        big.m_max_line_lens = 1 << 16;
        float gx = 0.1f, gy = 0.1f;
        big.updateMap(&gx, &gy, 1, 0, 0, 0);
    }
    std::vector<float> cosA(n), sinA(n);
    for (int i = 0; i < n; i++) {
        const float a = min_angle + dfi * i;
        cosA[i] = cosf(a);
        sinA[i] = sinf(a);
    }
    std::vector<std::vector<float> > gxs(K), gys(K);
    float min_x = 10000, min_y = 10000, max_x = -10000, max_y = -10000;
    for (int k = 0; k < K; k++) {
        const double rx = traj[3 * k], ry = traj[3 * k + 1], th = traj[3 * k + 2];
        const float fcos = cosf(th), fsin = sinf(th);
        for (int i = 0; i < n; i++) {
            const float b = scans[(size_t)k * n + i];
            if (b > max_range_cm) continue;
            const float x = (b / 100.0) * cosA[i], y = (b / 100.0) * sinA[i];
            const float tx = fcos * x - fsin * y + rx, ty = fsin * x + fcos * y + ry;
            gxs[k].push_back(tx);
            gys[k].push_back(ty);
            if (tx < min_x) min_x = tx;
            if (tx > max_x) max_x = tx;
            if (ty < min_y) min_y = ty;
            if (ty > max_y) max_y = ty;
        }
    }
    const int size_x = ((max_x - min_x + 0.05) / 0.05) * 2, size_y = ((max_y - min_y + 0.05) / 0.05) * 2;
    if (dims) { dims[0] = size_x; dims[1] = size_y; }
    if (size_x <= 0 || size_y <= 0) return 0;
    CPMap map;
    map.initializeMap(size_x, size_y, size_x / 2, size_y / 2, 0, 0, 0.05);
    for (int k = 0; k < K; k++)
        map.updateMap(gxs[k].empty() ? 0 : &gxs[k][0], gys[k].empty() ? 0 : &gys[k][0], (int)gxs[k].size(), traj[3 * k],
                      traj[3 * k + 1], traj[3 * k + 2]);
    map.computeProbsOfMap();
    std::vector<float> cx, cy;
    map.getPointCloud(cx, cy);
    for (size_t i = 0; i < cx.size() && (int)i < cap; i++) { out_x[i] = cx[i]; out_y[i] = cy[i]; }
    return (int)cx.size();
}
}
