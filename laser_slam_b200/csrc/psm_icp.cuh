// laser_slam_b200/csrc/psm_icp.cuh -- pm_icp (ZHPolar_Match.cpp:1653-2021): point-to-point ICP over an
// index window with point-to-line interpolated targets and a 20 % trim.
//
// Order-dependent pieces of the reference that are reproduced, not "fixed":
//   * the trim runs imax = (int)(n*0.2) passes of an adjacent-swap bubble sort; besides moving the worst matches to
//     the end, the passes permute the kept matches.  One warp runs up to 32 passes at once as a systolic pipeline
//     (pass p trails pass p-1 by two positions), which is exactly the sequential result;
//   * foot points are stored per REF index, so when several kept matches share a ref point they all end up using
//     the foot point of the LAST of them in (permuted) list order: an atomicMax over list positions finds it;
//   * means and cross sums are accumulated sequentially in float: one lane per running sum, in list order.
#pragma once
#include "psm_match.cuh"

namespace psm {

// point_line_distance (ZHPolar_Match.cpp:2036-2056); returns -1 when the projection falls outside the segment
__device__ __forceinline__ float point_line_distance(float x1, float y1, float x2, float y2, float x3, float y3,
                                                     float *x, float *y) {
    const float ax = x2 - x1, ay = y2 - y1;
    const float D = sqrtf(ax * ax + ay * ay);
    if ((double)D < 0.0001) return -1;
    const float t1 = -(-ax * x3 + ay * y1 + ax * x1 - ay * y3) / (ax * ax + ay * ay);
    if (t1 < 0 || t1 > 1) return -1;
    *x = x1 + t1 * ax;
    *y = y1 + t1 * ay;
    return sqrtf((x3 - *x) * (x3 - *x) + (y3 - *y) * (y3 - *y));
}

// imax bubble passes over the match list (dist[], pairs[]) of length n, by warp 0.  Pass p keeps a carry (the
// running maximum) and writes min(carry, a[j]) to a[j-1]; lane l of a group starts two steps after lane l-1.
__device__ __forceinline__ void bubble_passes(const Sh &s, int n, int imax) {
    const int lane = threadIdx.x & 31;
    for (int g = 0; g < imax; g += 32) {
        const int p = g + lane;
        const bool active = p < imax;
        const int m = n - p;  // this pass works on positions 0..m-1
        float cd = 0;
        int cp = 0;
        const int steps = 2 * 31 + (n - g) + 1;
        // one instruction stream for all lanes (no divergent paths): j == 0 loads the carry, 1 <= j < m compares the
        // next element with the carry and stores the smaller one behind, j == m stores the carry (the element beyond the
        // end is read as +inf, so the same min/max logic writes the carry)
        for (int tau = 0; tau < steps; tau++) {
            const int j = tau - 2 * lane;
            const bool in_range = active && j >= 0 && j <= m;
            const bool has_in = in_range && j < m;
            const int jj = has_in ? j : 0;
            float vd = s.dist[jj];
            int vp = s.pairs[jj];
            if (!has_in) vd = __int_as_float(0x7f800000);
            const bool lt = vd < cd;
            const float wd = lt ? vd : cd;
            const int wp = lt ? vp : cp;
            if (in_range && j >= 1) { s.dist[j - 1] = wd; s.pairs[j - 1] = wp; }
            if (in_range && (j == 0 || !lt)) { cd = vd; cp = vp; }
            __syncwarp();
        }
    }
}

// pm_icp with the ref scan at the origin.  pose in/out (cm, cm, rad).  Whole-CTA call.
__device__ __forceinline__ float pm_icp(const Sh &s, const LaserDev &L, float pose[3], int *iters, int *status) {
    const int n = L.n, t = threadIdx.x, window = L.scan_window;
    float ax = pose[0], ay = pose[1], ath = pose[2];
    for (int i = t; i < n; i += T) {
        s.refx[i] = s.ref_r[i] * s.co[i];
        s.refy[i] = s.ref_r[i] * s.si[i];
        s.actx[i] = s.act_r[i] * s.co[i];
        s.acty[i] = s.act_r[i] * s.si[i];
    }
    int iter = -1, small_corr_cnt = 0, n_valid_points = 0, nm = 0;
    float dx = 0, dy = 0, dth = 0;
    *status = ST_OK;
    __syncthreads();
    while (++iter < 60 && small_corr_cnt < 3) {
        if ((double)(fabsf(dx) + fabsf(dy)) + (double)fabsf(dth) * (180.0 / PI_BODY) < 0.1) small_corr_cnt++;
        else small_corr_cnt = 0;
        scan_project(s, L, ax, ay, ath);
        const float co = zhm::cosf_glibc(ath), si = zhm::sinf_glibc(ath);
        for (int i = t; i < n; i += T) {
            s.nx[i] = s.actx[i] * co - s.acty[i] * si + ax;
            s.ny[i] = s.actx[i] * si + s.acty[i] * co + ay;
        }
        __syncthreads();
        // nearest ref point inside the index window [i-window, i+window)
        const int per = (n + T - 1) / T;
        const int lo = t * per, hi = min(n, lo + per);
        int mine = 0;
        for (int i = lo; i < hi; i++) {
            float min_d = 1000000;
            int min_idx = -1;
            if (!s.new_bad[i]) {
                const int imin = max(0, i - window), imx = min(n, i + window);
                const float px = s.nx[i], py = s.ny[i];
                for (int j = imin; j < imx; j++) {
                    if (!s.ref_bad[j]) {
                        const float ex = px - s.refx[j], ey = py - s.refy[j];
                        const float d = ex * ex + ey * ey;
                        if (d < min_d) { min_d = d; min_idx = j; }
                    }
                }
            }
            const bool ok = min_idx >= 0 && sqrtf(min_d) < 100.0f;
            s.tmpj[i] = ok ? (short)min_idx : (short)-1;
            s.tmpd[i] = sqrtf(min_d);
            mine += ok;
        }
        // compaction in beam order: exclusive scan of the per-thread counts
        s.scan[t] = mine;
        __syncthreads();
        if (t < 32) {
            int run = 0;
            for (int b = 0; b < T; b += 32) {
                const int v = s.scan[b + t];
                int incl = v;
                for (int o = 1; o < 32; o <<= 1) {
                    const int u = __shfl_up_sync(0xffffffffu, incl, o);
                    if (t >= o) incl += u;
                }
                s.scan[b + t] = run + incl - v;
                run += __shfl_sync(0xffffffffu, incl, 31);
            }
            if (t == 0) s.scan[T] = run;
        }
        __syncthreads();
        nm = s.scan[T];
        {
            int pos = s.scan[t];
            for (int i = lo; i < hi; i++)
                if (s.tmpj[i] >= 0) { s.pairs[pos] = i | ((int)s.tmpj[i] << 16); s.dist[pos] = s.tmpd[i]; pos++; }
        }
        for (int j = t; j < n; j += T) s.lastpos[j] = -1;
        __syncthreads();
        if (nm < L.min_valid_points) { *status = ST_TOO_FEW; break; }
        int imax = (int)((double)nm * 0.2);
        n_valid_points = imax;
        if (t < 32) bubble_passes(s, nm, imax);
        __syncthreads();
        // point-to-line interpolation of the targets for the kept matches (list positions 0 .. nm-imax-1)
        const int kept = nm - imax;
        for (int i = t; i < kept; i += T) {
            const int a = s.pairs[i] & 0xffff, j = s.pairs[i] >> 16;
            const float px = s.nx[a], py = s.ny[a];
            float d1 = -1, d2 = -1, minx1 = 0, miny1 = 0, minx2 = 0, miny2 = 0;
            if (j > 0) d1 = point_line_distance(s.refx[j - 1], s.refy[j - 1], s.refx[j], s.refy[j], px, py, &minx1, &miny1);
            if (j < n - 1) d2 = point_line_distance(s.refx[j], s.refy[j], s.refx[j + 1], s.refy[j + 1], px, py, &minx2, &miny2);
            float fx = s.refx[j], fy = s.refy[j];
            const float ex = s.refx[j] - px, ey = s.refy[j] - py;
            float d0 = sqrtf(ex * ex + ey * ey);
            if (d1 > 0 && d1 < d0) { fx = minx1; fy = miny1; d0 = d1; }
            if (d2 > 0 && d2 < d0) { fx = minx2; fy = miny2; }
            s.fx[i] = fx;
            s.fy[i] = fy;
            atomicMax(&s.lastpos[j], i);
        }
        __syncthreads();
        for (int i = t; i < kept; i += T) {
            const int j = s.pairs[i] >> 16;
            if (s.lastpos[j] == i) { s.ix[j] = s.fx[i]; s.iy[j] = s.fy[i]; }
        }
        __syncthreads();
        // means and cross sums are running float sums in list order: all threads gather / evaluate the terms (a product
        // is rounded to float before it is added, as in the reference), then lanes 0..3 of warp 0 each add one array
        {
            float *g0 = s.pr, *g1 = s.pfi, *g2 = s.err, *g3 = s.tmpd;  // scratch that is free at this point
            for (int i = t; i < kept; i += T) {
                const int a = s.pairs[i] & 0xffff, j = s.pairs[i] >> 16;
                g0[i] = s.nx[a]; g1[i] = s.ny[a]; g2[i] = s.ix[j]; g3[i] = s.iy[j];
            }
            __syncthreads();
            if (t < 32) {
                float acc = 0;
                if (t < 4) acc = seq_sum(t == 0 ? g0 : t == 1 ? g1 : t == 2 ? g2 : g3, kept);
                acc /= (float)kept;
                if (t < 4) s.ctl[8 + t] = acc;
            }
            __syncthreads();
            const float meanpx = s.ctl[8], meanpy = s.ctl[9], meanppx = s.ctl[10], meanppy = s.ctl[11];
            for (int i = t; i < kept; i += T) {
                const float px = g0[i] - meanpx, py = g1[i] - meanpy, qx = g2[i] - meanppx, qy = g3[i] - meanppy;
                g0[i] = px * qx;  // Sxx
                g1[i] = px * qy;  // Sxy
                g2[i] = py * qx;  // Syx
                g3[i] = py * qy;  // Syy
            }
            __syncthreads();
            if (t < 32) {
                float sacc = 0;
                if (t < 4) sacc = seq_sum(t == 0 ? g0 : t == 1 ? g1 : t == 2 ? g2 : g3, kept);
                const float sxx = __shfl_sync(0xffffffffu, sacc, 0), sxy = __shfl_sync(0xffffffffu, sacc, 1);
                const float syx = __shfl_sync(0xffffffffu, sacc, 2), syy = __shfl_sync(0xffffffffu, sacc, 3);
                if (t == 0) {
                    const float ddth = zhm::atan2f_glibc(sxy - syx, sxx + syy);
                    const float c = zhm::cosf_glibc(ddth), sn = zhm::sinf_glibc(ddth);
                    const float ddx = meanppx - ax - (c * (meanpx - ax) - sn * (meanpy - ay));
                    const float ddy = meanppy - ay - (sn * (meanpx - ax) + c * (meanpy - ay));
                    s.ctl[0] = ddx; s.ctl[1] = ddy; s.ctl[2] = ddth;
                }
            }
        }
        __syncthreads();
        dx = s.ctl[0]; dy = s.ctl[1]; dth = s.ctl[2];
        ax += dx;
        ay += dy;
        ath += dth;
        ath = norm_a(ath);
        __syncthreads();
    }
    *iters = iter;
    float ret = 0;
    if (*status == ST_OK) {
        pose[0] = ax; pose[1] = ay; pose[2] = ath;
        // return value: mean residual over the first n_valid_points matches of the last iteration's list
        // (this_x/that_x hold each match's own foot point at the time it was processed)
        if (t == 0) {
            float abs_err = 0;
            for (int i = 0; i < n_valid_points; i++) {
                const int a = s.pairs[i] & 0xffff;
                const float ex = s.nx[a] - s.fx[i], ey = s.ny[a] - s.fy[i];
                abs_err += sqrtf(ex * ex + ey * ey);
            }
            s.ctl[4] = abs_err / (float)n_valid_points;
        }
        __syncthreads();
        ret = s.ctl[4];
        __syncthreads();
    }
    return ret;
}

}  // namespace psm
