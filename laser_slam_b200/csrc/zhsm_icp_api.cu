// laser_slam_b200/csrc/zhsm_icp_api.cu -- C-ABI entry points of zhicp (include/zhsm.h): host logic only.
#include <math.h>
#include <string.h>

#include "zhsm_internal.h"

extern "C" void zhsm_icp_default_options(zhsm_icp_options *o) {
    if (!o) return;
    memset(o, 0, sizeof(*o));
    // CICPWarpper::initICP (src/zhicp/ZHIcp_Warpper.cpp:43-53)
    o->maxIterations = 100;
    o->thresholdAng = (float)(10.0f * M_PI / 180.0);
    o->thresholdDist = 0.75f;
    o->ALFA = 0.5f;
    o->smallestThresholdDist = 0.05f;
    // mrpt::slam::CICP::TConfigParams defaults
    o->onlyClosestCorrespondences = 1;
    o->minAbsStep_trans = 1e-6f;
    o->minAbsStep_rot = 1e-6f;
    o->corresponding_points_decimation = 5;
    o->rotate_decimation_offset = 2;
    o->kernel_rho = 0.07f;
    o->use_kernel = 1;
    o->Axy_aprox_derivatives = 0.05f;
}

extern "C" zhsm_status zhsm_icp_batch(zhsm_handle *h, const float *pts, long long npts, const int32_t *off,
                                      const int32_t *cnt, int C, const int32_t *ref_idx, const int32_t *cur_idx,
                                      const double *init, int B, const zhsm_icp_options *opt, zhsm_result *results,
                                      unsigned flags, void *cuda_stream) {
    if (!h) return ZHSM_ERR_ARG;
    if (!pts || !off || !cnt || !ref_idx || !cur_idx || !results || C <= 0 || B < 0 || npts < 0) {
        h->err = "zhsm_icp_batch: bad argument";
        return ZHSM_ERR_ARG;
    }
    if (B == 0) return ZHSM_OK;
    IcpArgs a;
    memset(&a, 0, sizeof(a));
    if (opt) a.opt = *opt; else zhsm_icp_default_options(&a.opt);
    if (!(a.opt.ALFA >= 0 && a.opt.ALFA < 1) || a.opt.maxIterations < 1 || a.opt.smallestThresholdDist <= 0) {
        h->err = "zhsm_icp_batch: options out of range (0 <= ALFA < 1, maxIterations >= 1, smallestThresholdDist > 0)";
        return ZHSM_ERR_ARG;
    }
    int cap = 1;
    for (int c = 0; c < C; c++) {
        if (cnt[c] < 0 || off[c] < 0 || (long long)off[c] + cnt[c] > npts) { h->err = "zhsm_icp_batch: cloud table out of range"; return ZHSM_ERR_ARG; }
        if (cnt[c] > cap) cap = cnt[c];
    }
    a.cap = cap;
    a.B = B;
    a.C = C;
    if (icp_smem_bytes(cap) > 227 * 1024) { h->err = "zhsm_icp_batch: cloud too large for shared memory"; return ZHSM_ERR_UNSUPPORTED; }
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
    StreamScope scope(h, st);
    zhsm_status s;
    void *p;
    if (flags & ZHSM_SCANS_ON_DEVICE) a.pts = pts;
    else {
        if ((s = zhsm_stage(h, 5, sizeof(float) * 2 * (size_t)npts + 16, &p)) != ZHSM_OK) return s;
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(p, pts, sizeof(float) * 2 * (size_t)npts, cudaMemcpyHostToDevice, st));
        a.pts = (const float *)p;
    }
    if ((s = zhsm_stage(h, 6, sizeof(int32_t) * 2 * (size_t)C, &p)) != ZHSM_OK) return s;
    {
        int32_t *d = (int32_t *)p;
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(d, off, sizeof(int32_t) * C, cudaMemcpyHostToDevice, st));
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(d + C, cnt, sizeof(int32_t) * C, cudaMemcpyHostToDevice, st));
        a.off = d; a.cnt = d + C;
    }
    if (flags & ZHSM_JOBS_ON_DEVICE) { a.ref_idx = ref_idx; a.cur_idx = cur_idx; a.init = init; }
    else {
        for (int k = 0; k < B; k++)
            if (ref_idx[k] < 0 || ref_idx[k] >= C || cur_idx[k] < 0 || cur_idx[k] >= C) { h->err = "zhsm_icp_batch: cloud index out of range"; return ZHSM_ERR_ARG; }
        if ((s = zhsm_stage(h, 8, sizeof(int32_t) * 2 * (size_t)B, &p)) != ZHSM_OK) return s;
        int32_t *di = (int32_t *)p;
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(di, ref_idx, sizeof(int32_t) * B, cudaMemcpyHostToDevice, st));
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(di + B, cur_idx, sizeof(int32_t) * B, cudaMemcpyHostToDevice, st));
        a.ref_idx = di; a.cur_idx = di + B;
        if (init) {
            if ((s = zhsm_stage(h, 9, sizeof(double) * 3 * (size_t)B, &p)) != ZHSM_OK) return s;
            ZHSM_CUDA_TRY(h, cudaMemcpyAsync(p, init, sizeof(double) * 3 * B, cudaMemcpyHostToDevice, st));
            a.init = (const double *)p;
        }
    }
    if (flags & ZHSM_RESULTS_ON_DEVICE) a.results = results;
    else {
        if ((s = zhsm_stage(h, 10, sizeof(zhsm_result) * (size_t)B, &p)) != ZHSM_OK) return s;
        a.results = (zhsm_result *)p;
    }
    ZHSM_CUDA_TRY(h, cudaEventRecord(h->ev0, st));
    ZHSM_CUDA_TRY(h, icp_launch(a, st));
    h->launches++;
    ZHSM_CUDA_TRY(h, cudaEventRecord(h->ev1, st));
    h->ev_valid = true;
    if (!(flags & ZHSM_RESULTS_ON_DEVICE)) {
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(results, a.results, sizeof(zhsm_result) * (size_t)B, cudaMemcpyDeviceToHost, st));
        ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    } else {
        // off/cnt were staged from host memory that the caller may free: the copies above are async on pageable
        // memory, which the runtime completes before returning; nothing else to wait for
    }
    return ZHSM_OK;
}

extern "C" zhsm_status zhsm_best_of_starts(zhsm_handle *h, const zhsm_result *results, int P, int G, int criterion,
                                           int32_t *best, zhsm_result *best_results, unsigned flags, void *cuda_stream) {
    if (!h) return ZHSM_ERR_ARG;
    if (!results || P < 0 || G < 1 || criterion < 0 || criterion > 1 || (!best && !best_results)) { h->err = "zhsm_best_of_starts: bad argument"; return ZHSM_ERR_ARG; }
    if (P == 0) return ZHSM_OK;
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
    StreamScope scope(h, st);
    if (flags & ZHSM_RESULTS_ON_DEVICE) {
        ZHSM_CUDA_TRY(h, best_of_starts_launch(results, P, G, criterion, best, best_results, st));
        h->launches++;
        return ZHSM_OK;
    }
    zhsm_status s;
    void *p;
    const size_t nin = sizeof(zhsm_result) * (size_t)P * G;
    if ((s = zhsm_stage(h, 11, nin, &p)) != ZHSM_OK) return s;
    zhsm_result *din = (zhsm_result *)p;
    if ((s = zhsm_stage(h, 12, sizeof(zhsm_result) * (size_t)P, &p)) != ZHSM_OK) return s;
    zhsm_result *dout = (zhsm_result *)p;
    if ((s = zhsm_stage(h, 13, sizeof(int32_t) * (size_t)P, &p)) != ZHSM_OK) return s;
    int32_t *dbest = (int32_t *)p;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(din, results, nin, cudaMemcpyHostToDevice, st));
    ZHSM_CUDA_TRY(h, best_of_starts_launch(din, P, G, criterion, dbest, dout, st));
    h->launches++;
    if (best) ZHSM_CUDA_TRY(h, cudaMemcpyAsync(best, dbest, sizeof(int32_t) * (size_t)P, cudaMemcpyDeviceToHost, st));
    if (best_results) ZHSM_CUDA_TRY(h, cudaMemcpyAsync(best_results, dout, sizeof(zhsm_result) * (size_t)P, cudaMemcpyDeviceToHost, st));
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    return ZHSM_OK;
}
