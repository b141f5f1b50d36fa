// laser_slam_b200/csrc/graph_sm100.cu -- SURVEY 8f-1: the arithmetic either side of the ICP batches in the reference's
// loop-closure search (src/mapGraph/MapGraph.cpp:1342-1442): candidate gating produces the (node, candidate) job list,
// consensus consumes the matched poses.  Small, data-parallel, no g2o involved.
//   gate_kernel       isOverlappedArea :957-994 + isLoopyArea :996-1032, one thread per (current node, other node)
//   available_kernel  isAvaliableArea :918-944, one thread per matched transform
//   consensus_kernel  deleteMisMatch :169-209 (statis :48-61), calculateWeightedPose :146-167 and the choice made by
//                     calculateFinalPose :110-123; matched sets are a handful of poses, the reference's float sums are
//                     order-sensitive, so one thread walks one set in the reference's order
// float / double are used exactly where the reference declares them (corners pass through float in calBounding).
#include <math.h>
#include <string.h>

#include "zhsm_internal.h"

namespace graphk {

__device__ __forceinline__ void oplus_point(const double *p, double x, double y, double *out) {  // point.cpp:63-69
    const double c = cos(p[2]), s = sin(p[2]);
    out[0] = p[0] + x * c - y * s;
    out[1] = p[1] + x * s + y * c;
}

__device__ __forceinline__ void cal_bounding(double *ld, double *ru) {
    const float x_min = ld[0] < ru[0] ? ld[0] : ru[0], x_max = ld[0] > ru[0] ? ld[0] : ru[0];
    const float y_min = ld[1] < ru[1] ? ld[1] : ru[1], y_max = ld[1] > ru[1] ? ld[1] : ru[1];
    ld[0] = x_min; ld[1] = y_min; ru[0] = x_max; ru[1] = y_max;
}

__device__ __forceinline__ float cov_radius2(const double *cov_xy, float constant, float beta) {
    if (!cov_xy) return constant;
    const float dis_x = (float)cov_xy[0], dis_y = (float)cov_xy[1];
    return (fabsf(dis_x) + fabsf(dis_y)) * beta;
}

__global__ void gate_kernel(const double *pose, const double *ld, const double *ru, const double *cov_xy, int N,
                            const int32_t *cur_idx, int M, float beta, uint8_t *flags) {
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= (long long)M * N) return;
    const int m = (int)(g / N), j = (int)(g % N), c = cur_idx[m];
    const double *pc = pose + 3 * (size_t)c, *pr = pose + 3 * (size_t)j;
    int f = 0;
    {
        double ld_cur[2], ru_cur[2], ld_ref[2], ru_ref[2];
        oplus_point(pc, ld[2 * c], ld[2 * c + 1], ld_cur); oplus_point(pc, ru[2 * c], ru[2 * c + 1], ru_cur);
        oplus_point(pr, ld[2 * j], ld[2 * j + 1], ld_ref); oplus_point(pr, ru[2 * j], ru[2 * j + 1], ru_ref);
        cal_bounding(ld_cur, ru_cur);
        cal_bounding(ld_ref, ru_ref);
        bool ok = !(ru_cur[0] <= ld_cur[0] || ru_cur[1] <= ld_cur[1] || ru_ref[0] <= ld_ref[0] || ru_ref[1] <= ld_ref[1]);
        ok = ok && !(ru_cur[0] <= ld_ref[0] || ru_cur[1] <= ld_ref[1] || ru_ref[0] <= ld_cur[0] || ru_ref[1] <= ld_cur[1]);
        if (ok) {
            const float x_min = ld_cur[0] > ld_ref[0] ? ld_cur[0] : ld_ref[0], y_min = ld_cur[1] > ld_ref[1] ? ld_cur[1] : ld_ref[1];
            const float x_max = ru_cur[0] < ru_ref[0] ? ru_cur[0] : ru_ref[0], y_max = ru_cur[1] < ru_ref[1] ? ru_cur[1] : ru_ref[1];
            const float overlapped = (x_max - x_min) * (y_max - y_min);
            const float curArea = (float)((ru_cur[0] - ld_cur[0]) * (ru_cur[1] - ld_cur[1]));
            const float ratio = overlapped / curArea;
            if (overlapped > 0 && (double)ratio >= 0.4) f |= 1;
        }
    }
    {
        const float dis = cov_radius2(cov_xy ? cov_xy + 2 * (size_t)c : nullptr, 2 * 2, beta);
        const float cur_dis = (float)((pc[0] - pr[0]) * (pc[0] - pr[0]) + (pc[1] - pr[1]) * (pc[1] - pr[1]));
        if (dis >= cur_dis) f |= 2;
    }
    flags[g] = (uint8_t)f;
}

__global__ void available_kernel(const double *trans, const double *cov_xy, int M, uint8_t *ok) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    const float beta = 25, angle_thre = (float)(M_PI / 3.0);
    const double *t = trans + 3 * (size_t)m;
    if (fabs(t[2]) >= (double)angle_thre) { ok[m] = 0; return; }
    const float dis = cov_radius2(cov_xy ? cov_xy + 2 * (size_t)m : nullptr, (float)(2.5 * 2.5), beta);
    const float cur_dis = (float)(t[0] * t[0] + t[1] * t[1]);
    ok[m] = dis >= cur_dis;
}

__device__ __forceinline__ float dis_poses(const double *a, const double *b) {
    return (float)sqrt((a[0] - b[0]) * (a[0] - b[0]) + (a[1] - b[1]) * (a[1] - b[1]));
}

// dis: scratch [total] floats
__global__ void consensus_kernel(const double *poses, const float *good, const int32_t *set_off, int S, uint8_t *keep,
                                 double *final_pose, int32_t *final_idx, float *dis) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const int lo = set_off[s], n = set_off[s + 1] - lo;
    for (int i = 0; i < n; i++) keep[lo + i] = 1;
    final_idx[s] = -1;
    final_pose[3 * s] = final_pose[3 * s + 1] = final_pose[3 * s + 2] = 0;
    if (n <= 0) return;
    if (n >= 3) {
        for (int i = 0; i < n; i++) {
            float mean_dis = 0;
            for (int j = 0; j < n; j++) {
                if (j == i) continue;
                mean_dis += dis_poses(poses + 3 * (size_t)(lo + i), poses + 3 * (size_t)(lo + j));
            }
            dis[lo + i] = mean_dis / (float)(n - 1);
        }
        float sum = 0;
        for (int i = 0; i < n; i++) sum += dis[lo + i];
        const float mean = sum / (float)n;
        float sq = 0;
        for (int i = 0; i < n; i++) sq += (dis[lo + i] - mean) * (dis[lo + i] - mean);
        const float sigma = sq / (float)(n - 1);
        const float scale = 1.731f;
        for (int i = 0; i < n; i++)
            if (dis[lo + i] > mean + scale * sigma) keep[lo + i] = 0;
    }
    float total = 0;
    for (int i = 0; i < n; i++) if (keep[lo + i]) total += good[lo + i];
    float px = 0, py = 0, pth = 0;
    for (int i = 0; i < n; i++) {
        if (!keep[lo + i]) continue;
        const float w = good[lo + i] / total;
        px = (float)((double)px + (double)w * poses[3 * (size_t)(lo + i)]);
        py = (float)((double)py + (double)w * poses[3 * (size_t)(lo + i) + 1]);
        pth = (float)((double)pth + (double)w * poses[3 * (size_t)(lo + i) + 2]);
    }
    const double fp[3] = {px, py, pth};
    final_pose[3 * s] = px; final_pose[3 * s + 1] = py; final_pose[3 * s + 2] = pth;
    float min_dis = 1000000;
    int best = -1;
    for (int i = 0; i < n; i++) {
        if (!keep[lo + i]) continue;
        const float d = dis_poses(fp, poses + 3 * (size_t)(lo + i));
        if (d < min_dis) { min_dis = d; best = lo + i; }
    }
    final_idx[s] = best;
}

}  // namespace graphk

// ---- C ABI (include/zhsm.h).  All arrays are HOST arrays: these calls sit on the graph side of the pipeline, where
// the reference keeps its node table; sizes are tiny next to the scan pools.
namespace {
struct Dev {
    zhsm_handle *h;
    int slot;
    zhsm_status st;
    cudaStream_t s;
    Dev(zhsm_handle *h_, cudaStream_t s_) : h(h_), slot(5), st(ZHSM_OK), s(s_) {}
    template <class T>
    T *up(const T *src, size_t n) {
        void *p = nullptr;
        if (st != ZHSM_OK) return nullptr;
        if ((st = zhsm_stage(h, slot++, sizeof(T) * (n ? n : 1), &p)) != ZHSM_OK) return nullptr;
        if (src && n && cudaMemcpyAsync(p, src, sizeof(T) * n, cudaMemcpyHostToDevice, s) != cudaSuccess) { h->err = "H2D copy failed"; st = ZHSM_ERR_CUDA; }
        return (T *)p;
    }
};
}  // namespace

extern "C" zhsm_status zhsm_gate_candidates(zhsm_handle *h, const double *pose, const double *ld, const double *ru,
                                            const double *cov_xy, int N, const int32_t *cur_idx, int M, float beta,
                                            uint8_t *flags) {
    if (!h) return ZHSM_ERR_ARG;
    if (!pose || !ld || !ru || !cur_idx || !flags || N <= 0 || M < 0) { h->err = "zhsm_gate_candidates: bad argument"; return ZHSM_ERR_ARG; }
    for (int m = 0; m < M; m++) if (cur_idx[m] < 0 || cur_idx[m] >= N) { h->err = "zhsm_gate_candidates: node index out of range"; return ZHSM_ERR_ARG; }
    if (M == 0) return ZHSM_OK;
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    StreamScope scope(h, h->stream);
    Dev d(h, h->stream);
    const double *dp = d.up(pose, 3 * (size_t)N), *dl = d.up(ld, 2 * (size_t)N), *dr = d.up(ru, 2 * (size_t)N);
    const double *dc = cov_xy ? d.up(cov_xy, 2 * (size_t)N) : nullptr;
    const int32_t *di = d.up(cur_idx, (size_t)M);
    uint8_t *df = d.up((const uint8_t *)nullptr, (size_t)M * N);
    if (d.st != ZHSM_OK) return d.st;
    const long long tot = (long long)M * N;
    graphk::gate_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, h->stream>>>(dp, dl, dr, dc, N, di, M, beta, df);
    ZHSM_CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(flags, df, (size_t)tot, cudaMemcpyDeviceToHost, h->stream));
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return ZHSM_OK;
}

extern "C" zhsm_status zhsm_available_area(zhsm_handle *h, const double *trans, const double *cov_xy, int M, uint8_t *ok) {
    if (!h) return ZHSM_ERR_ARG;
    if (!trans || !ok || M < 0) { h->err = "zhsm_available_area: bad argument"; return ZHSM_ERR_ARG; }
    if (M == 0) return ZHSM_OK;
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    StreamScope scope(h, h->stream);
    Dev d(h, h->stream);
    const double *dt = d.up(trans, 3 * (size_t)M);
    const double *dc = cov_xy ? d.up(cov_xy, 2 * (size_t)M) : nullptr;
    uint8_t *dk = d.up((const uint8_t *)nullptr, (size_t)M);
    if (d.st != ZHSM_OK) return d.st;
    graphk::available_kernel<<<(M + 127) / 128, 128, 0, h->stream>>>(dt, dc, M, dk);
    ZHSM_CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(ok, dk, (size_t)M, cudaMemcpyDeviceToHost, h->stream));
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return ZHSM_OK;
}

extern "C" zhsm_status zhsm_consensus(zhsm_handle *h, const double *poses, const float *good, const int32_t *set_off,
                                      int S, uint8_t *keep, double *final_pose, int32_t *final_idx) {
    if (!h) return ZHSM_ERR_ARG;
    if (!poses || !good || !set_off || !keep || !final_pose || !final_idx || S < 0) { h->err = "zhsm_consensus: bad argument"; return ZHSM_ERR_ARG; }
    if (S == 0) return ZHSM_OK;
    for (int s = 0; s < S; s++) if (set_off[s + 1] < set_off[s] || set_off[0] != 0) { h->err = "zhsm_consensus: set offsets must start at 0 and not decrease"; return ZHSM_ERR_ARG; }
    const size_t total = (size_t)set_off[S];
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    StreamScope scope(h, h->stream);
    Dev d(h, h->stream);
    const double *dp = d.up(poses, 3 * total);
    const float *dg = d.up(good, total);
    const int32_t *dof = d.up(set_off, (size_t)S + 1);
    uint8_t *dk = d.up((const uint8_t *)nullptr, total);
    double *dfp = d.up((const double *)nullptr, 3 * (size_t)S);
    int32_t *dfi = d.up((const int32_t *)nullptr, (size_t)S);
    float *dd = d.up((const float *)nullptr, total);
    if (d.st != ZHSM_OK) return d.st;
    graphk::consensus_kernel<<<(S + 63) / 64, 64, 0, h->stream>>>(dp, dg, dof, S, dk, dfp, dfi, dd);
    ZHSM_CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    if (total) ZHSM_CUDA_TRY(h, cudaMemcpyAsync(keep, dk, total, cudaMemcpyDeviceToHost, h->stream));
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(final_pose, dfp, sizeof(double) * 3 * S, cudaMemcpyDeviceToHost, h->stream));
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(final_idx, dfi, sizeof(int32_t) * S, cudaMemcpyDeviceToHost, h->stream));
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return ZHSM_OK;
}
