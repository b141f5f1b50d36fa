// laser_slam_b200/host/ZHIcp_Warpper.cpp -- see ZHIcp_Warpper.h.  Host logic only; the ICP runs in libzhsm.so's CUDA
// kernel (csrc/icp_sm100.cu) through zhsm_icp_batch.
#include "ZHIcp_Warpper.h"

#include <string.h>

#include <stdexcept>
#include <string>

static void check(zhsm_handle *h, zhsm_status st, const char *what) {
    if (st != ZHSM_OK) throw std::runtime_error(std::string(what) + ": " + (h ? zhsm_last_error(h) : "no handle"));
}

CICPWarpper::CICPWarpper(int device) : skip_window(true), ICP_method(0), h_(0) {
    memset(&last, 0, sizeof(last));
    if (zhsm_create(device, &h_) != ZHSM_OK)
        throw std::runtime_error("CICPWarpper: no usable sm_100 GPU (this matcher has no CPU path)");
    initICP();
}

CICPWarpper::~CICPWarpper() { zhsm_destroy(h_); }

void CICPWarpper::initICP() { zhsm_icp_default_options(&options); }  // the values of ZHIcp_Warpper.cpp:43-53

float CICPWarpper::ICPMatchMultiStart(const float *rx, const float *ry, int N1, const float *ax, const float *ay, int N2,
                                      const double *pini, int G, double pout[3], zhsm_result *results) {
    if (ICP_method != 0) throw std::runtime_error("CICPWarpper: only icpClassic (ICP_method = 0) is implemented");
    if (N1 < 0 || N2 < 0 || G < 1) throw std::invalid_argument("CICPWarpper::ICPMatch: negative size");
    std::vector<float> pool(2 * ((size_t)N1 + N2) + 1);
    if (N1) { memcpy(&pool[0], rx, sizeof(float) * N1); memcpy(&pool[N1], ry, sizeof(float) * N1); }
    if (N2) { memcpy(&pool[2 * (size_t)N1], ax, sizeof(float) * N2); memcpy(&pool[2 * (size_t)N1 + N2], ay, sizeof(float) * N2); }
    const int32_t off[2] = {0, N1}, cnt[2] = {N1, N2};
    std::vector<int32_t> ri(G, 0), ci(G, 1);
    std::vector<zhsm_result> res(G);
    check(h_, zhsm_icp_batch(h_, &pool[0], (long long)N1 + N2, off, cnt, 2, &ri[0], &ci[0], pini, G, &options, &res[0], 0, 0),
          "zhsm_icp_batch");
    int best = 0;
    if (G > 1) {
        int32_t b = -1;
        check(h_, zhsm_best_of_starts(h_, &res[0], 1, G, ZHSM_BEST_GOODNESS, &b, 0, 0, 0), "zhsm_best_of_starts");
        best = b < 0 ? 0 : b;
    }
    last = res[best];
    if (results) memcpy(results, &res[0], sizeof(zhsm_result) * G);
    pout[0] = last.x[0]; pout[1] = last.x[1]; pout[2] = last.x[2];
    return last.goodness;
}

float CICPWarpper::ICPMatch(float *rx, float *ry, int N1, float *ax, float *ay, int N2, double pini[3], double pout[3]) {
    return ICPMatchMultiStart(rx, ry, N1, ax, ay, N2, pini, 1, pout, 0);
}

float CICPWarpper::ICPMatch(float *rx, float *ry, float *ax, float *ay, int N, double pini[3], double pout[3]) {
    return ICPMatchMultiStart(rx, ry, N, ax, ay, N, pini, 1, pout, 0);
}

void CICPWarpper::ICPMatch(std::vector<float> &rx, std::vector<float> &ry, std::vector<float> &ax, std::vector<float> &ay,
                           double px, double py, double pth) {
    double pini[3] = {px, py, pth}, pout[3];
    static float none = 0;
    ICPMatchMultiStart(rx.empty() ? &none : &rx[0], ry.empty() ? &none : &ry[0], (int)rx.size(), ax.empty() ? &none : &ax[0],
                       ay.empty() ? &none : &ay[0], (int)ax.size(), pini, 1, pout, 0);
}

void CICPWarpper::getResult(double p[3], double cov[6]) {
    for (int k = 0; k < 3; k++) p[k] = last.x[k];
    for (int k = 0; k < 6; k++) cov[k] = last.cov[k];
}

void CICPWarpper::getMatchQuality(int &iterations, float &good, float &quality) {
    iterations = last.iterations;
    good = last.goodness;
    quality = last.quality;
}
