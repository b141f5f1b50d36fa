// laser_slam_b200/csrc/psm_sm100.cu -- zhpsm kernels for sm_100a: pm_preprocessScan, pm_scan_project, and the matchers
// pm_psm / pm_icp / CPolarMatch::FMatch, one CTA per job (src/zhpsm/ZHPolar_Match.cpp).
#include "psm_icp.cuh"

namespace psm {

using Args = PsmArgs;

__device__ __forceinline__ void load_job(const Sh &s, const Args &a, int job, bool with_ref) {
    const int n = a.L.n, t = threadIdx.x;
    const size_t ao = (size_t)a.act_idx[job] * n;
    const size_t ro = with_ref ? (size_t)a.ref_idx[job] * n : 0;
    for (int i = t; i < n; i += T) {
        s.act_r[i] = a.scans[ao + i];
        s.act_bad[i] = a.bad[ao + i];
        s.act_seg[i] = a.seg[ao + i];
        if (with_ref) { s.ref_r[i] = a.scans[ro + i]; s.ref_bad[i] = a.bad[ro + i]; }
        s.fi[i] = a.fi[i];
        s.si[i] = a.si[i];
        s.co[i] = a.co[i];
    }
    __syncthreads();
}

__global__ void __launch_bounds__(T) psm_kernel(Args a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const Sh s = carve(smem_raw, a.L.n);
    const int job = blockIdx.x;
    if ((unsigned)a.act_idx[job] >= (unsigned)a.S || (unsigned)a.ref_idx[job] >= (unsigned)a.S) {
        // device-resident job lists are not checked by the host: a bad index fails its own job, never the batch
        if (threadIdx.x == 0) {
            zhsm_result r = {};
            r.status = ZHSM_JOB_INVALID_SCAN;
            a.results[job] = r;
        }
        return;
    }
    load_job(s, a, job, true);
    float pose[3] = {0, 0, 0};
    if (a.init_pose) for (int k = 0; k < 3; k++) pose[k] = a.init_pose[3 * (size_t)job + k];
    int iters = 0, iters2 = 0, status = ST_OK;
    float err = 0, ei[3] = {0, 0, 0};
    if (a.algo == ZHSM_ALGO_PSM) {
        err = pm_psm(s, a.L, pose, &iters, &status);
    } else if (a.algo == ZHSM_ALGO_PM_ICP) {
        err = pm_icp(s, a.L, pose, &iters, &status);
    } else if (a.algo == ZHSM_ALGO_PSM_ICP) {
        pm_psm(s, a.L, pose, &iters, &status);
        if (status == ST_OK) err = pm_icp(s, a.L, pose, &iters2, &status);
    } else {  // ZHSM_ALGO_FMATCH: CPolarMatch::FMatch (:81-121) after preprocessing: zero poses, pm_icp, pm_error_index2
        pose[0] = pose[1] = pose[2] = 0;
        err = pm_icp(s, a.L, pose, &iters, &status);
        if (status == ST_OK) error_index2(s, a.L, pose, ei);
    }
    if (threadIdx.x == 0) {
        zhsm_result r = {};
        r.valid = status == ST_OK;
        r.status = status;
        r.iterations = iters + iters2;
        r.x[0] = pose[0]; r.x[1] = pose[1]; r.x[2] = pose[2];
        r.error = err;
        if (a.algo == ZHSM_ALGO_FMATCH) {
            // the pair FMatch returns: (avg_err_x, avg_err_y), or (-1,-1) when the matcher threw (twice)
            r.cov[0] = status == ST_OK ? ei[0] : -1.0;
            r.cov[3] = status == ST_OK ? ei[1] : -1.0;
            r.nvalid = (int)ei[2];
        }
        a.results[job] = r;
    }
}

__global__ void __launch_bounds__(T) scan_project_kernel(Args a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const Sh s = carve(smem_raw, a.L.n);
    const int job = blockIdx.x, n = a.L.n;
    load_job(s, a, job, false);
    const float *p = a.init_pose + 3 * (size_t)job;
    scan_project(s, a.L, p[0], p[1], p[2]);
    for (int i = threadIdx.x; i < n; i += T) {
        a.proj_r[(size_t)job * n + i] = s.new_r[i];
        a.proj_bad[(size_t)job * n + i] = s.new_bad[i];
    }
}

// pm_find_far_points (:1583-1590) + pm_segment_scan (:1495-1576) on median-filtered ranges: a left-to-right state
// machine with look-back writes (seg[i-1], seg[i-2]); one thread walks one scan.
__global__ void far_segment_kernel(const float *scans, unsigned char *bad, short *seg, int S, int n, int max_range_cm) {
    const int sidx = blockIdx.x * blockDim.x + threadIdx.x;
    if (sidx >= S) return;
    const float *r = scans + (size_t)sidx * n;
    unsigned char *b = bad + (size_t)sidx * n;
    short *sg = seg + (size_t)sidx * n;
    for (int i = 0; i < n; i++)
        if (r[i] > (float)max_range_cm) b[i] |= PM_RANGE;
    const float MAX_DIST = 20.0;
    int seg_cnt = 1, cnt;
    if (fabsf(r[0] - r[1]) < MAX_DIST) { sg[0] = seg_cnt; sg[1] = seg_cnt; cnt = 2; }
    else { sg[0] = 0; sg[1] = seg_cnt; cnt = 1; }
    // carried state: r and seg of i-1, i-2 (seg[i-1], seg[i-2] can be rewritten by the singleton / rescue rules)
    for (int i = 2; i < n; i++) {
        bool break_seg = false;
        if (b[i]) { break_seg = true; sg[i] = 0; }
        else {
            const float dr = (float)((double)r[i] - (2.0 * (double)r[i - 1] - (double)r[i - 2]));
            if (fabsf(r[i] - r[i - 1]) < MAX_DIST || ((sg[i - 1] == sg[i - 2]) && fabsf(dr) < MAX_DIST)) { cnt++; sg[i] = seg_cnt; }
            else break_seg = true;
        }
        if (break_seg) {
            if (cnt == 1) {
                const float dr = (float)((double)r[i] - (2.0 * (double)r[i - 1] - (double)r[i - 2]));
                if (sg[i - 2] == 0 && b[i] == 0 && b[i - 1] == 0 && b[i - 2] == 0 && fabsf(dr) < MAX_DIST) {
                    sg[i] = seg_cnt; sg[i - 1] = seg_cnt; sg[i - 2] = seg_cnt; cnt = 3;
                } else { sg[i - 1] = 0; sg[i] = seg_cnt; cnt = 1; }
            } else { seg_cnt++; sg[i] = seg_cnt; cnt = 1; }
        }
    }
}

// pm_is_corridor (ZHPolar_Match.cpp:2058-2186): standard deviation of the (3-tap median filtered) directions of the
// lines through neighbouring points of a segment, computed twice (the second time rotated by pi/5, without the median:
// the reference's second window assignment order makes all three taps equal).  Sequential float sums: one thread
// walks one scan.  out: 1 corridor, 0 room, -3 / -2 = the reference's `throw 3` / `throw 2` (fewer than 2 lines).
__global__ void corridor_kernel(const float *scans, const unsigned char *bad, const short *seg, const float *fi_t,
                                const float *si_t, const float *co_t, int S, int n, signed char *out) {
    const int sidx = blockIdx.x * blockDim.x + threadIdx.x;
    if (sidx >= S) return;
    const float *r = scans + (size_t)sidx * n;
    const unsigned char *b = bad + (size_t)sidx * n;
    const short *sg = seg + (size_t)sidx * n;
    const double R2D = 180.0 / PI_BODY;
    float fi1 = 0, fi2 = 0, fi3 = 0, sxx = 0, sx = 0, cnt = 0, std1, std2;
    for (int i = 0; i < n - 1; i++) {
        if (sg[i] == sg[i + 1] && sg[i] != 0 && !b[i]) {
            const float x = r[i] * co_t[i], y = r[i] * si_t[i];
            const float x1 = r[i + 1] * co_t[i + 1], y1 = r[i + 1] * si_t[i + 1];
            float fi = (float)((double)zhm::atan2f_glibc(y1 - y, x1 - x) * R2D);
            if (fi < 0) fi = (float)((double)fi + 180.0);
            if (i == 0) { fi1 = fi; fi2 = fi; fi3 = fi; }
            fi3 = fi2; fi2 = fi1; fi1 = fi;
            if (fi1 <= fi2 && fi2 <= fi3) fi = fi2;
            if (fi2 <= fi3 && fi3 <= fi1) fi = fi3;
            if (fi3 <= fi1 && fi1 <= fi2) fi = fi1;
            sx += fi;
            sxx += fi * fi;
            cnt = (float)((double)cnt + 1.0);
        }
    }
    if (cnt > 1) {
        const float var = (float)((double)(sxx - sx * sx / cnt) / ((double)cnt - 1.0));
        std1 = var > 0 ? sqrtf(var) : 0.f;
    } else { out[sidx] = -3; return; }
    sxx = 0; cnt = 0; sx = 0;
    for (int i = 0; i < n - 1; i++) {
        if (sg[i] == sg[i + 1] && sg[i] != 0 && !b[i]) {
            const float a0 = (float)((double)fi_t[i] + PI_BODY / 5.0), a1 = (float)((double)fi_t[i + 1] + PI_BODY / 5.0);
            const float x = r[i] * zhm::cosf_glibc(a0), y = r[i] * zhm::sinf_glibc(a0);
            const float x1 = r[i + 1] * zhm::cosf_glibc(a1), y1 = r[i + 1] * zhm::sinf_glibc(a1);
            float fi = (float)((double)zhm::atan2f_glibc(y1 - y, x1 - x) * R2D);
            if (fi < 0) fi = (float)((double)fi + 180.0);
            sx += fi;
            sxx += fi * fi;
            cnt = (float)((double)cnt + 1.0);
        }
    }
    if (cnt > 1) {
        const float var = (float)((double)(sxx - sx * sx / cnt) / ((double)cnt - 1.0));
        std2 = var > 0 ? sqrtf(var) : 0.f;
    } else { out[sidx] = -2; return; }
    const float st = std1 < std2 ? std1 : std2;
    out[sidx] = st < 25.0f ? 1 : 0;  // PM_CORRIDOR_THRESHOLD (PolarParameter.h:24)
}

}  // namespace psm

cudaError_t corridor_launch(const float *scans, const unsigned char *bad, const short *seg, const float *fi,
                            const float *si, const float *co, int S, int n, signed char *out, cudaStream_t st) {
    psm::corridor_kernel<<<(S + 63) / 64, 64, 0, st>>>(scans, bad, seg, fi, si, co, S, n, out);
    return cudaGetLastError();
}

cudaError_t psm_launch(const psm::Args &a, int mode, cudaStream_t st) {
    const size_t smem = psm::sh_bytes(a.L.n);
    cudaError_t e;
    if (mode == 0) {
        e = cudaFuncSetAttribute(psm::psm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        psm::psm_kernel<<<a.B, psm::T, smem, st>>>(a);
    } else {
        e = cudaFuncSetAttribute(psm::scan_project_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        psm::scan_project_kernel<<<a.B, psm::T, smem, st>>>(a);
    }
    return cudaGetLastError();
}

cudaError_t far_segment_launch(const float *scans, unsigned char *bad, short *seg, int S, int n, int max_range_cm,
                               cudaStream_t st) {
    psm::far_segment_kernel<<<(S + 127) / 128, 128, 0, st>>>(scans, bad, seg, S, n, max_range_cm);
    return cudaGetLastError();
}
