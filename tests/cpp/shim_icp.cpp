// tests/cpp/shim_icp.cpp -- drives the C++ shim CICPWarpper the way the reference's src/zhicp/testWarp.cpp does:
//   shim_icp <pool.bin> <out.txt>
// pool.bin: int32 C, P, G; int32 off[C], cnt[C]; int32 pairs[P][2]; double init[P][3]; double grid[G][3];
//           float pts[2 * sum cnt] (x block then y block per cloud)
// out: per pair "goodness px py pth iterations good quality cov0..cov5", once through each ICPMatch overload that
// applies (array + sizes, vector), then "best_goodness px py pth" of the G-start run (init + grid).
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "ZHIcp_Warpper.h"

template <class T>
static bool rd(FILE *f, std::vector<T> &v, size_t n) {
    v.resize(n);
    return n == 0 || fread(&v[0], sizeof(T), n, f) == n;
}

int main(int argc, char **argv) {
    if (argc < 3) return 2;
    FILE *f = fopen(argv[1], "rb");
    if (!f) return 3;
    int hdr[3];
    if (fread(hdr, sizeof(int), 3, f) != 3) return 3;
    const int C = hdr[0], P = hdr[1], G = hdr[2];
    std::vector<int> off, cnt, pairs;
    std::vector<double> init, grid;
    std::vector<float> pts;
    if (!rd(f, off, C) || !rd(f, cnt, C) || !rd(f, pairs, 2 * (size_t)P) || !rd(f, init, 3 * (size_t)P) || !rd(f, grid, 3 * (size_t)G)) return 3;
    size_t np = 0;
    for (int c = 0; c < C; c++) np += cnt[c];
    if (!rd(f, pts, 2 * np)) return 3;
    fclose(f);
    FILE *out = fopen(argv[2], "w");
    if (!out) return 3;
    CICPWarpper icp;
    icp.setDebugWindowOff();
    for (int k = 0; k < P; k++) {
        const int r = pairs[2 * k], c = pairs[2 * k + 1];
        float *rx = &pts[2 * (size_t)off[r]], *ry = rx + cnt[r], *ax = &pts[2 * (size_t)off[c]], *ay = ax + cnt[c];
        double pout[3], p[3], cov[6];
        int it;
        float good, quality;
        float g = icp.ICPMatch(rx, ry, cnt[r], ax, ay, cnt[c], &init[3 * k], pout);
        icp.getResult(p, cov);
        icp.getMatchQuality(it, good, quality);
        fprintf(out, "%.9g %.17g %.17g %.17g %d %.9g %.9g", g, pout[0], pout[1], pout[2], it, good, quality);
        for (int j = 0; j < 6; j++) fprintf(out, " %.17g", cov[j]);
        fprintf(out, "\n");
        std::vector<float> vrx(rx, rx + cnt[r]), vry(ry, ry + cnt[r]), vax(ax, ax + cnt[c]), vay(ay, ay + cnt[c]);
        icp.ICPMatch(vrx, vry, vax, vay, init[3 * k], init[3 * k + 1], init[3 * k + 2]);
        icp.getResult(p, cov);
        if (p[0] != pout[0] || p[1] != pout[1] || p[2] != pout[2]) { fprintf(stderr, "overloads disagree on pair %d\n", k); return 4; }
        std::vector<double> starts(3 * (size_t)G);
        for (int s = 0; s < G; s++) for (int j = 0; j < 3; j++) starts[3 * s + j] = init[3 * k + j] + grid[3 * s + j];
        g = icp.ICPMatchMultiStart(rx, ry, cnt[r], ax, ay, cnt[c], &starts[0], G, pout);
        fprintf(out, "%.9g %.17g %.17g %.17g\n", g, pout[0], pout[1], pout[2]);
    }
    fclose(out);
    return 0;
}
