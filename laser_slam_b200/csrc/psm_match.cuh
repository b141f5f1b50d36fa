// laser_slam_b200/csrc/psm_match.cuh -- pm_psm: orientation search + translation estimation.
//
// The reference accumulates every sum sequentially in float, so the order of additions is part of the result.  It is
// kept: the orientation search gives each shift of the window to one thread, which walks the beams in order; the
// translation step gives each of its seven running sums to one lane of warp 0.  No tree reductions on this path.
#pragma once
#include "psm_common.cuh"

namespace psm {

enum { ST_OK = 0, ST_TOO_FEW = ZHSM_JOB_TOO_FEW_POINTS, ST_SINGULAR = ZHSM_JOB_SINGULAR, ST_SEARCH = ZHSM_JOB_SEARCH_FAILED };

// pm_orientation_search (ZHPolar_Match.cpp:1152-1261).  Whole-CTA call; returns dth and a status on every thread.
__device__ __forceinline__ float orientation_search(const Sh &s, const LaserDev &L, int *status) {
    const int n = L.n, t = threadIdx.x, window = L.scan_window;
    const int K = 2 * window + 1;
    const float LARGE_NUMBER = 10000;
    for (int k = t; k < K; k += T) {
        const int di = k - window;
        int min_i, max_i;
        if (di <= 0) { min_i = -di; max_i = n; } else { min_i = 0; max_i = n - di; }
        float e = 0, cnt = 0;
        if (L.flags & ZHSM_LASER_UTM_GATE) {
            // the "HOKUYO_UTM_30LX" branch (:1182-1191): delta < PM_MAX_ERROR (100 cm) picks the right local minimum
            for (int i = min_i; i < max_i; i++) {
                const float delta = fabsf(s.new_r[i] - s.ref_r[i + di]);
                if (!s.new_bad[i] && !s.ref_bad[i + di] && delta < 100.f) { e += delta; cnt++; }
            }
        } else {
            for (int i = min_i; i < max_i; i++) {
                const float delta = fabsf(s.new_r[i] - s.ref_r[i + di]);
                if (!s.new_bad[i] && !s.ref_bad[i + di]) { e += delta; cnt++; }
            }
        }
        s.err[k] = cnt > 0 ? e / cnt : LARGE_NUMBER;
    }
    __syncthreads();
    if (t == 0) {
        float emin = (float)(LARGE_NUMBER * 10.0);
        int imin = -1;
        for (int i = 0; i < K; i++)
            if (s.err[i] < emin) { emin = s.err[i]; imin = i; }
        int st = ST_OK;
        float dth = 0;
        if (imin < 0 || s.err[imin] >= LARGE_NUMBER) st = ST_SEARCH;
        else {
            dth = (float)(imin - window) * L.dfi;
            if (imin >= 1 && imin < (K - 1)) {
                const float D = (float)((double)(s.err[imin - 1] + s.err[imin + 1]) - 2.0 * (double)s.err[imin]);
                float d = LARGE_NUMBER;
                if ((double)fabsf(D) > 0.01 && s.err[imin - 1] > s.err[imin] && s.err[imin + 1] > s.err[imin])
                    d = (float)((double)((s.err[imin - 1] - s.err[imin + 1]) / D) / 2.0);
                if ((double)fabsf(d) < 1.0) dth += d * L.dfi;
            }
        }
        s.ctl[0] = dth;
        s.ctl[1] = __int_as_float(st);
    }
    __syncthreads();
    *status = __float_as_int(s.ctl[1]);
    const float r = s.ctl[0];
    __syncthreads();
    return r;
}

// One lane adds arr[0..n) to a running float sum strictly in index order (the reference's order); the loads do not
// depend on the sum, so they are issued four at a time and only the FADD chain is serial.
__device__ __forceinline__ float seq_sum(const float *arr, int n) {
    float acc = 0;
    int i = 0;
    for (; i + 4 <= n; i += 4) {
        const float a = arr[i], b = arr[i + 1], c = arr[i + 2], d = arr[i + 3];
        acc += a; acc += b; acc += c; acc += d;
    }
    for (; i < n; i++) acc += arr[i];
    return acc;
}

// pm_translation_estimation (ZHPolar_Match.cpp:1015-1131).  Whole-CTA call.  The reference keeps seven running float
// sums over the beams (abs_err, n, hw1, hw2, hwh11, hwh12, hwh22).  All threads first evaluate the per-beam terms
// (products are rounded to float before they are added, exactly as `hw1 += hwi1*dr` does; a beam that fails the
// association test contributes +0.0f, which leaves a running sum that started at +0 unchanged), then lanes 0..5 of
// warp 0 each add one term array in beam order; the count is an integer and order-free.
__device__ __forceinline__ float translation_estimation(const Sh &s, const LaserDev &L, float C, float *dx, float *dy,
                                                        int *status) {
    const int n = L.n, t = threadIdx.x;
    float *t0 = s.nx, *t1 = s.ny, *t2 = s.refx, *t3 = s.refy, *t4 = s.actx, *t5 = s.acty;  // pm_icp's arrays are free here
    if (t == 0) s.scan[0] = 0;
    __syncthreads();
    int mine = 0;
    for (int i = t; i < n; i += T) {
        const float dr = s.ref_r[i] - s.new_r[i];
        const bool ok = s.ref_bad[i] == 0 && s.new_bad[i] == 0 && s.new_r[i] < (float)L.max_range_cm && s.new_r[i] > 10.0f &&
                        fabsf(dr) < 100.0f;
        const float w = C / (dr * dr + C);
        const float hi1 = s.co[i], hi2 = s.si[i];
        const float hwi1 = hi1 * w, hwi2 = hi2 * w;
        t0[i] = fabsf(dr);
        t1[i] = ok ? hwi1 * dr : 0.0f;
        t2[i] = ok ? hwi2 * dr : 0.0f;
        t3[i] = ok ? hwi1 * hi1 : 0.0f;
        t4[i] = ok ? hwi1 * hi2 : 0.0f;
        t5[i] = ok ? hwi2 * hi2 : 0.0f;
        mine += ok;
    }
    mine = __reduce_add_sync(0xffffffffu, mine);
    if ((t & 31) == 0 && mine) atomicAdd(&s.scan[0], mine);
    __syncthreads();
    if (t < 32) {
        float acc = 0;
        if (t < 6) acc = seq_sum(t == 0 ? t0 : t == 1 ? t1 : t == 2 ? t2 : t == 3 ? t3 : t == 4 ? t4 : t5, n);
        const float abs_err = __shfl_sync(0xffffffffu, acc, 0);
        const float hw1 = __shfl_sync(0xffffffffu, acc, 1), hw2 = __shfl_sync(0xffffffffu, acc, 2);
        const float hwh11 = __shfl_sync(0xffffffffu, acc, 3), hwh12 = __shfl_sync(0xffffffffu, acc, 4);
        const float hwh22 = __shfl_sync(0xffffffffu, acc, 5);
        const int nn = s.scan[0];
        if (t == 0) {
            int st = ST_OK;
            float ddx = 0, ddy = 0, ret = 0;
            if (nn < L.min_valid_points) st = ST_TOO_FEW;
            else {
                const float hwh21 = 0;
                const float D = hwh11 * hwh22 - hwh12 * hwh21;
                if ((double)D < 0.001) st = ST_SINGULAR;
                else {
                    const float inv11 = hwh22 / D, inv12 = -hwh12 / D, inv21 = -hwh12 / D, inv22 = hwh11 / D;
                    ddx = inv11 * hw1 + inv12 * hw2;
                    ddy = inv21 * hw1 + inv22 * hw2;
                    ret = abs_err / (float)nn;
                }
            }
            s.ctl[0] = ddx; s.ctl[1] = ddy; s.ctl[2] = ret; s.ctl[3] = __int_as_float(st);
        }
    }
    __syncthreads();
    *dx = s.ctl[0];
    *dy = s.ctl[1];
    const float ret = s.ctl[2];
    *status = __float_as_int(s.ctl[3]);
    __syncthreads();
    return ret;
}

// pm_psm (ZHPolar_Match.cpp:890-1003) with the ref scan at the origin (its callers zero both poses, :97-99).
// pose in/out (cm, cm, rad); returns avg_err; *iters = final value of the reference's `iter`.
__device__ __forceinline__ float pm_psm(const Sh &s, const LaserDev &L, float pose[3], int *iters, int *status) {
    float ax = pose[0], ay = pose[1], ath = pose[2];
    float C = 70 * 70;
    int iter = -1, small_corr_cnt = 0;
    float dx = 0, dy = 0, dth = 0, avg_err = 100000000.0f;
    *status = ST_OK;
    while (++iter < 30 && small_corr_cnt < 3) {
        if ((double)(fabsf(dx) + fabsf(dy) + fabsf(dth)) < 0.4) small_corr_cnt++;
        else small_corr_cnt = 0;
        scan_project(s, L, ax, ay, ath);
        if (iter % 2 == 0) {
            dth = orientation_search(s, L, status);
            if (*status) break;
            ath += dth;
            continue;
        }
        if (iter == 10) C = (float)((double)C / 50.0);
        avg_err = translation_estimation(s, L, C, &dx, &dy, status);
        if (*status) break;
        ax += dx;
        ay += dy;
    }
    *iters = iter;
    if (*status == ST_OK) { pose[0] = ax; pose[1] = ay; pose[2] = ath; }  // a throw leaves lsa's pose untouched
    return avg_err;
}

// pm_error_index2 (ZHPolar_Match.cpp:1279-1339): mean squared x / y components of the range residual (in metres)
// after projecting the current scan at `pose`; lanes 0..3 own e, sum_err_x, sum_err_y, n.  out = (avg_err_x,
// avg_err_y, associated points); NaN when no beam qualifies (the reference returns uninitialised floats there).
__device__ __forceinline__ void error_index2(const Sh &s, const LaserDev &L, const float pose[3], float out[3]) {
    scan_project(s, L, pose[0], pose[1], pose[2]);
    const int n = L.n, t = threadIdx.x;
    if (t < 32) {
        float acc = 0;
        int cnt = 0;
        if (t < 4) {
            for (int i = 0; i < n; i++) {
                const float delta = (float)((double)fabsf(s.new_r[i] - s.ref_r[i]) * 0.01);
                if (!s.new_bad[i] && !s.ref_bad[i] && (double)delta < 100 / 2.0 && delta < 1) {
                    if (t == 0) acc += delta;
                    else if (t == 1) { const float v = delta * s.co[i]; acc += v * v; }
                    else if (t == 2) { const float v = delta * s.si[i]; acc += v * v; }
                    else cnt++;
                }
            }
        }
        const float sx = __shfl_sync(0xffffffffu, acc, 1), sy = __shfl_sync(0xffffffffu, acc, 2);
        const int nn = __shfl_sync(0xffffffffu, cnt, 3);
        if (t == 0) {
            s.ctl[0] = nn > 0 ? sx / (float)nn : nanf("");
            s.ctl[1] = nn > 0 ? sy / (float)nn : nanf("");
            s.ctl[2] = (float)nn;
        }
    }
    __syncthreads();
    out[0] = s.ctl[0]; out[1] = s.ctl[1]; out[2] = s.ctl[2];
    __syncthreads();
}

}  // namespace psm
