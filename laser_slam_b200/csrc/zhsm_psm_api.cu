// laser_slam_b200/csrc/zhsm_psm_api.cu -- C-ABI entry points of the zhpsm family (include/zhsm.h): host logic only.
#include <string.h>

#include "zh_libm.cuh"
#include "zhsm_internal.h"

static LaserDev laser_dev(const zhsm_handle *h) {
    LaserDev L;
    L.n = h->laser.n; L.fi_min = h->laser.fi_min; L.dfi = h->laser.dfi;
    L.min_range_cm = h->laser.min_range_cm; L.max_range_cm = h->laser.max_range_cm;
    L.min_valid_points = h->laser.min_valid_points; L.scan_window = h->laser.scan_window;
    L.flags = h->laser.flags;
    return L;
}

// stages host arrays scans/bad/seg into slots 5,6,7 (or passes device pointers through)
static zhsm_status stage_scans(zhsm_handle *h, const float *scans, const uint8_t *bad, const int16_t *seg, int S,
                               unsigned flags, cudaStream_t st, const float **ds, const uint8_t **db, const int16_t **dg) {
    const size_t n = h->laser.n, cnt = (size_t)S * n;
    if (flags & ZHSM_SCANS_ON_DEVICE) { *ds = scans; *db = bad; *dg = seg; return ZHSM_OK; }
    void *p;
    zhsm_status s;
    if ((s = zhsm_stage(h, 5, cnt * 4, &p)) != ZHSM_OK) return s;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(p, scans, cnt * 4, cudaMemcpyHostToDevice, st));
    *ds = (const float *)p;
    if ((s = zhsm_stage(h, 6, cnt, &p)) != ZHSM_OK) return s;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(p, bad, cnt, cudaMemcpyHostToDevice, st));
    *db = (const uint8_t *)p;
    if ((s = zhsm_stage(h, 7, cnt * 2, &p)) != ZHSM_OK) return s;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(p, seg, cnt * 2, cudaMemcpyHostToDevice, st));
    *dg = (const int16_t *)p;
    return ZHSM_OK;
}

extern "C" zhsm_status zhsm_preprocess_batch(zhsm_handle *h, float *scans, uint8_t *bad, int16_t *seg, int S,
                                             unsigned flags, void *cuda_stream) {
    if (!h || !scans || !bad || !seg || S < 0) return ZHSM_ERR_ARG;
    if (!h->have_laser) { h->err = "zhsm_set_laser not called"; return ZHSM_ERR_ARG; }
    if (S == 0) return ZHSM_OK;
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
    StreamScope scope(h, st);
    const size_t n = h->laser.n, cnt = (size_t)S * n;
    float *ds = scans;
    uint8_t *db = bad;
    int16_t *dg = seg;
    const bool host = !(flags & ZHSM_SCANS_ON_DEVICE);
    if (host) {
        void *p;
        zhsm_status s;
        if ((s = zhsm_stage(h, 5, cnt * 4, &p)) != ZHSM_OK) return s;
        ds = (float *)p;
        if ((s = zhsm_stage(h, 6, cnt, &p)) != ZHSM_OK) return s;
        db = (uint8_t *)p;
        if ((s = zhsm_stage(h, 7, cnt * 2, &p)) != ZHSM_OK) return s;
        dg = (int16_t *)p;
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(ds, scans, cnt * 4, cudaMemcpyHostToDevice, st));
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(db, bad, cnt, cudaMemcpyHostToDevice, st));
    }
    ZHSM_CUDA_TRY(h, median_filter_launch(ds, S, (int)n, st));
    ZHSM_CUDA_TRY(h, far_segment_launch(ds, db, (short *)dg, S, (int)n, h->laser.max_range_cm, st));
    h->launches += 2;
    if (host) {
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(scans, ds, cnt * 4, cudaMemcpyDeviceToHost, st));
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(bad, db, cnt, cudaMemcpyDeviceToHost, st));
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(seg, dg, cnt * 2, cudaMemcpyDeviceToHost, st));
        ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    }
    return ZHSM_OK;
}

extern "C" zhsm_status zhsm_psm_batch(zhsm_handle *h, const float *scans, const uint8_t *bad, const int16_t *seg, int S,
                                      const int32_t *ref_idx, const int32_t *act_idx, const float *init_pose, int B,
                                      int algo, zhsm_result *results, unsigned flags, void *cuda_stream) {
    if (!h) return ZHSM_ERR_ARG;
    if (!scans || !bad || !seg || !ref_idx || !act_idx || !results || S <= 0 || B < 0 || algo < 0 || algo > 3) {
        h->err = "zhsm_psm_batch: bad argument";
        return ZHSM_ERR_ARG;
    }
    if (!h->have_laser) { h->err = "zhsm_set_laser not called"; return ZHSM_ERR_ARG; }
    if (2 * h->laser.scan_window + 1 > h->laser.n || h->laser.n > 32767) { h->err = "scan_window / beam count unsupported"; return ZHSM_ERR_ARG; }
    if (B == 0) return ZHSM_OK;
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
    StreamScope scope(h, st);
    PsmArgs a;
    memset(&a, 0, sizeof(a));
    a.B = B; a.algo = algo; a.S = S; a.L = laser_dev(h);
    a.fi = h->d_fi; a.si = h->d_si; a.co = h->d_co;
    zhsm_status s;
    if ((s = stage_scans(h, scans, bad, seg, S, flags, st, &a.scans, &a.bad, (const int16_t **)&a.seg)) != ZHSM_OK) return s;
    void *p;
    if (flags & ZHSM_JOBS_ON_DEVICE) { a.ref_idx = ref_idx; a.act_idx = act_idx; a.init_pose = init_pose; }
    else {
        for (int k = 0; k < B; k++)
            if (ref_idx[k] < 0 || ref_idx[k] >= S || act_idx[k] < 0 || act_idx[k] >= S) { h->err = "zhsm_psm_batch: scan index out of range"; return ZHSM_ERR_ARG; }
        if ((s = zhsm_stage(h, 8, sizeof(int32_t) * 2 * (size_t)B, &p)) != ZHSM_OK) return s;
        int32_t *di = (int32_t *)p;
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(di, ref_idx, sizeof(int32_t) * B, cudaMemcpyHostToDevice, st));
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(di + B, act_idx, sizeof(int32_t) * B, cudaMemcpyHostToDevice, st));
        a.ref_idx = di; a.act_idx = di + B;
        if (init_pose) {
            if ((s = zhsm_stage(h, 9, sizeof(float) * 3 * (size_t)B, &p)) != ZHSM_OK) return s;
            ZHSM_CUDA_TRY(h, cudaMemcpyAsync(p, init_pose, sizeof(float) * 3 * B, cudaMemcpyHostToDevice, st));
            a.init_pose = (const float *)p;
        }
    }
    if (flags & ZHSM_RESULTS_ON_DEVICE) a.results = results;
    else {
        if ((s = zhsm_stage(h, 10, sizeof(zhsm_result) * (size_t)B, &p)) != ZHSM_OK) return s;
        a.results = (zhsm_result *)p;
    }
    ZHSM_CUDA_TRY(h, cudaEventRecord(h->ev0, st));
    ZHSM_CUDA_TRY(h, psm_launch(a, 0, st));
    h->launches++;
    ZHSM_CUDA_TRY(h, cudaEventRecord(h->ev1, st));
    h->ev_valid = true;
    if (!(flags & ZHSM_RESULTS_ON_DEVICE)) {
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(results, a.results, sizeof(zhsm_result) * (size_t)B, cudaMemcpyDeviceToHost, st));
        ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    }
    return ZHSM_OK;
}

extern "C" zhsm_status zhsm_scan_project_batch(zhsm_handle *h, const float *scans, const uint8_t *bad, const int16_t *seg,
                                               int S, const int32_t *idx, const float *pose, int B, float *new_r,
                                               uint8_t *new_bad) {
    if (!h || !scans || !bad || !seg || !idx || !pose || !new_r || !new_bad || S <= 0 || B <= 0) return ZHSM_ERR_ARG;
    if (!h->have_laser) { h->err = "zhsm_set_laser not called"; return ZHSM_ERR_ARG; }
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    StreamScope scope(h, st);
    const size_t n = h->laser.n;
    for (int k = 0; k < B; k++)
        if (idx[k] < 0 || idx[k] >= S) { h->err = "zhsm_scan_project_batch: scan index out of range"; return ZHSM_ERR_ARG; }
    PsmArgs a;
    memset(&a, 0, sizeof(a));
    a.B = B; a.S = S; a.L = laser_dev(h);
    a.fi = h->d_fi; a.si = h->d_si; a.co = h->d_co;
    zhsm_status s;
    if ((s = stage_scans(h, scans, bad, seg, S, 0, st, &a.scans, &a.bad, (const int16_t **)&a.seg)) != ZHSM_OK) return s;
    void *p;
    if ((s = zhsm_stage(h, 8, sizeof(int32_t) * (size_t)B, &p)) != ZHSM_OK) return s;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(p, idx, sizeof(int32_t) * B, cudaMemcpyHostToDevice, st));
    a.act_idx = (const int32_t *)p; a.ref_idx = a.act_idx;
    if ((s = zhsm_stage(h, 9, sizeof(float) * 3 * (size_t)B, &p)) != ZHSM_OK) return s;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(p, pose, sizeof(float) * 3 * B, cudaMemcpyHostToDevice, st));
    a.init_pose = (const float *)p;
    if ((s = zhsm_stage(h, 11, 4 * n * B, &p)) != ZHSM_OK) return s;
    a.proj_r = (float *)p;
    if ((s = zhsm_stage(h, 12, n * B, &p)) != ZHSM_OK) return s;
    a.proj_bad = (unsigned char *)p;
    ZHSM_CUDA_TRY(h, psm_launch(a, 1, st));
    h->launches++;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(new_r, a.proj_r, 4 * n * B, cudaMemcpyDeviceToHost, st));
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(new_bad, a.proj_bad, n * B, cudaMemcpyDeviceToHost, st));
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    return ZHSM_OK;
}

extern "C" zhsm_status zhsm_is_corridor_batch(zhsm_handle *h, const float *scans, const uint8_t *bad, const int16_t *seg,
                                              int S, int8_t *out, unsigned flags, void *cuda_stream) {
    if (!h || !scans || !bad || !seg || !out || S < 0) return ZHSM_ERR_ARG;
    if (!h->have_laser) { h->err = "zhsm_set_laser not called"; return ZHSM_ERR_ARG; }
    if (S == 0) return ZHSM_OK;
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
    StreamScope scope(h, st);
    const float *ds; const uint8_t *db; const int16_t *dg;
    zhsm_status s;
    if ((s = stage_scans(h, scans, bad, seg, S, flags, st, &ds, &db, &dg)) != ZHSM_OK) return s;
    int8_t *dout = out;
    if (!(flags & ZHSM_RESULTS_ON_DEVICE)) {
        void *p;
        if ((s = zhsm_stage(h, 10, (size_t)S, &p)) != ZHSM_OK) return s;
        dout = (int8_t *)p;
    }
    ZHSM_CUDA_TRY(h, corridor_launch(ds, db, (const short *)dg, h->d_fi, h->d_si, h->d_co, S, h->laser.n, (signed char *)dout, st));
    h->launches++;
    if (!(flags & ZHSM_RESULTS_ON_DEVICE)) {
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(out, dout, (size_t)S, cudaMemcpyDeviceToHost, st));
        ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    }
    return ZHSM_OK;
}

__global__ void libm_kernel(int which, const float *x, const float *y, float *out, int N) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    out[i] = which == 0 ? zhm::sinf_glibc(x[i]) : which == 1 ? zhm::cosf_glibc(x[i]) : zhm::atan2f_glibc(x[i], y[i]);
}

extern "C" zhsm_status zhsm_debug_libm(zhsm_handle *h, int which, const float *x, const float *y, float *out, int N) {
    if (!h || !x || !y || !out || N <= 0 || which < 0 || which > 2) return ZHSM_ERR_ARG;
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    float *d = nullptr;
    ZHSM_CUDA_TRY(h, cudaMalloc(&d, sizeof(float) * 3 * (size_t)N));
    ZHSM_CUDA_TRY(h, cudaMemcpy(d, x, sizeof(float) * N, cudaMemcpyHostToDevice));
    ZHSM_CUDA_TRY(h, cudaMemcpy(d + N, y, sizeof(float) * N, cudaMemcpyHostToDevice));
    libm_kernel<<<(N + 255) / 256, 256, 0, h->stream>>>(which, d, d + N, d + 2 * (size_t)N, N);
    h->launches++;
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    ZHSM_CUDA_TRY(h, cudaMemcpy(out, d + 2 * (size_t)N, sizeof(float) * N, cudaMemcpyDeviceToHost));
    cudaFree(d);
    return ZHSM_OK;
}
