// laser_slam_b200/csrc/plicp_kernel.cuh -- the PL-ICP kernel body (included by plicp_sm100.cu).
#pragma once

namespace plicp {

// One sens ray's find_correspondences_tricks search (icp_corr_tricks.c) given the last_best that the
// sequential algorithm would carry into it.  Returns j1 (>=0) or -1; fills d2 and j2.
__device__ __noinline__ int tricks_one(const Smem &s, const LaserDev &L, int n, double C1, double min_theta, double maxd2,
                                 double pwx, double pwy, double rho, double phi, int last_best, double *d2out,
                                 int *j2out) {
    const double HALF_PI = 3.14159265358979323846 * 0.5;
    const int from = 0, to = n - 1;
    int start_cell = (int)((phi - min_theta) * C1);
    int j1 = -1;
    double best_dist = 42;
    int we_start_at = (last_best == -1) ? start_cell : last_best + 1;
    if (we_start_at > to) we_start_at = to;
    if (we_start_at < from) we_start_at = from;
    int up = we_start_at + 1;
    int down = we_start_at;
    double last_dist_up = 0;
    double last_dist_down = -1;
    bool up_stopped = false, down_stopped = false;
    while (!up_stopped || !down_stopped) {
        bool now_up = up_stopped ? false : down_stopped ? true : (last_dist_up < last_dist_down);
        if (now_up) {
            if (up > to) { up_stopped = true; continue; }
            double rr = s.refr[up];
            if (rr < 0) { ++up; continue; }
            double2 q = s.refp[up];
            last_dist_up = dist2(pwx, pwy, q.x, q.y);
            if (last_dist_up < best_dist || j1 == -1) { j1 = up; best_dist = last_dist_up; }
            if (up > start_cell) {
                double delta_theta = theta_of(L, up) - phi;
                double min_dist_up = rho * ((delta_theta > HALF_PI) ? 1 : sin(delta_theta));
                if (min_dist_up * min_dist_up > best_dist) { up_stopped = true; continue; }
                up += (rr < rho) ? s.jt[0 * n + up] : s.jt[1 * n + up];
            } else {
                ++up;
            }
        } else {
            if (down < from) { down_stopped = true; continue; }
            double rr = s.refr[down];
            if (rr < 0) { --down; continue; }
            double2 q = s.refp[down];
            last_dist_down = dist2(pwx, pwy, q.x, q.y);
            if (last_dist_down < best_dist || j1 == -1) { j1 = down; best_dist = last_dist_down; }
            if (down < start_cell) {
                double delta_theta = phi - theta_of(L, down);
                double min_dist_down = rho * ((delta_theta > HALF_PI) ? 1 : sin(delta_theta));
                if (min_dist_down * min_dist_down > best_dist) { down_stopped = true; continue; }
                down += (rr < rho) ? s.jt[2 * n + down] : s.jt[3 * n + down];
            } else {
                --down;
            }
        }
    }
    if (j1 == -1 || best_dist > maxd2) return -1;
    if (j1 == 0 || j1 == n - 1) return -1;
    int j2up = -1, j2down = -1;
    for (int j = j1 + 1; j < n; j++) if (s.refr[j] >= 0) { j2up = j; break; }
    for (int j = j1 - 1; j >= 0; j--) if (s.refr[j] >= 0) { j2down = j; break; }
    if (j2up == -1 && j2down == -1) return -1;
    int j2;
    if (j2up == -1) j2 = j2down;
    else if (j2down == -1) j2 = j2up;
    else {
        double2 qu = s.refp[j2up], qd = s.refp[j2down];
        double dist_up = dist2(pwx, pwy, qu.x, qu.y);
        double dist_down = dist2(pwx, pwy, qd.x, qd.y);
        j2 = dist_up < dist_down ? j2up : j2down;
    }
    *d2out = best_dist;
    *j2out = j2;
    return j1;
}

// Per-thread correspondence state for its C consecutive rays.
template <int C>
struct Corr {
    int j1[C];
    int j2[C];
    double d2[C];
    unsigned valid;  // bit k: correspondence k valid
};

// find_correspondences_tricks over the thread's chunk, following last_best exactly inside the chunk.
// Returns the last_best leaving the chunk, or LB_PASS when nothing matched.
template <int C>
__device__ inline int search_chunk(const Smem &s, const LaserDev &L, int n, int i0, double C1, double min_theta,
                                   double maxd2, const double x_old[3], double cs, double sn, int lb_in, Corr<C> &c) {
    int lb = lb_in;
    int out = LB_PASS;
    c.valid = 0;
#pragma unroll
    for (int k = 0; k < C; k++) {
        int i = i0 + k;
        c.j1[k] = -1;
        c.j2[k] = -1;
        if (i >= n) continue;
        double2 p = s.sensp[i];
        if (isnan(p.x)) continue;  // invalid sens ray: null correspondence, last_best untouched
        // ld_compute_world_coords (laser_data.c)
        double wx = cs * p.x - sn * p.y + x_old[0];
        double wy = sn * p.x + cs * p.y + x_old[1];
        double rho = sqrt(wx * wx + wy * wy);
        double phi = atan2(wy, wx);
        double d2;
        int j2;
        int j1 = tricks_one(s, L, n, C1, min_theta, maxd2, wx, wy, rho, phi, lb, &d2, &j2);
        if (j1 >= 0) {
            c.j1[k] = j1;
            c.j2[k] = j2;
            c.d2[k] = d2;
            c.valid |= 1u << k;
            lb = j1;
            out = j1;
        }
    }
    return out;
}

// All-thread correspondence search with the speculate/verify fixpoint on last_best.  `guess` persists across
// iterations (correspondences barely move between iterations, so the first try is almost always right).
template <int C>
__device__ inline void find_correspondences(const Smem &s, const LaserDev &L, int n, int i0, double C1,
                                            double min_theta, double maxd2, const double x_old[3], double cs,
                                            double sn, int &guess, Corr<C> &c) {
    const int t = threadIdx.x;
    if (t == 0) guess = -1;
    int out = search_chunk<C>(s, L, n, i0, C1, min_theta, maxd2, x_old, cs, sn, guess, c);
    while (true) {
        s.outs[t] = out;
        __syncthreads();
        int actual = -1;
        for (int u = t - 1; u >= 0; u--) {
            int o = s.outs[u];
            if (o != LB_PASS) { actual = o; break; }
        }
        bool changed = (actual != guess);
        int any = __syncthreads_or(changed ? 1 : 0);
        if (!any) break;
        if (changed) {
            guess = actual;
            out = search_chunk<C>(s, L, n, i0, C1, min_theta, maxd2, x_old, cs, sn, guess, c);
        }
    }
}

// Exact order statistics of the block's valid keys (doubles >= 0 viewed as uint64): 8-bit MSD radix select for two
// ranks at once.  kill_outliers_trim sorts and reads dist2[order]; only those two values are needed.
template <int C>
__device__ inline void select2(const Smem &s, const double (&key)[C], unsigned valid, int rank0, int rank1,
                               double &v0, double &v1) {
    const int t = threadIdx.x;
    Ctl *ctl = s.ctl;
    if (t == 0) {
        ctl->sel_prefix[0] = ctl->sel_prefix[1] = 0;
        ctl->sel_rem[0] = rank0;
        ctl->sel_rem[1] = rank1;
    }
    for (int pass = 0; pass < 8; pass++) {
        const int shift = 56 - 8 * pass;
        for (int b = t; b < 512; b += blockDim.x) s.hist[b] = 0;
        __syncthreads();
        unsigned long long p0 = ctl->sel_prefix[0], p1 = ctl->sel_prefix[1];
#pragma unroll
        for (int k = 0; k < C; k++) {
            if (!(valid >> k & 1u)) continue;
            unsigned long long u = (unsigned long long)__double_as_longlong(key[k]);
            unsigned long long hi = pass == 0 ? 0ull : (u >> (shift + 8));
            int digit = (int)((u >> shift) & 255ull);
            if (hi == p0) atomicAdd(&s.hist[digit], 1);
            if (hi == p1) atomicAdd(&s.hist[256 + digit], 1);
        }
        __syncthreads();
        if (t < 64) {
            const int r = t >> 5, lane = t & 31;
            const int *h = s.hist + 256 * r;
            int loc[8];
            int sum = 0;
#pragma unroll
            for (int q = 0; q < 8; q++) { loc[q] = h[lane * 8 + q]; sum += loc[q]; }
            int incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            int excl = incl - sum;
            int rem = ctl->sel_rem[r];
            if (rem >= excl && rem < incl) {
                int acc = excl;
                int bsel = 0;
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    if (rem >= acc && rem < acc + loc[q]) { bsel = q; break; }
                    acc += loc[q];
                }
                ctl->sel_prefix[r] = (ctl->sel_prefix[r] << 8) | (unsigned long long)(lane * 8 + bsel);
                ctl->sel_rem[r] = rem - acc;
            }
        }
        __syncthreads();
    }
    v0 = __longlong_as_double((long long)ctl->sel_prefix[0]);
    v1 = __longlong_as_double((long long)ctl->sel_prefix[1]);
}

}  // namespace plicp

#include "plicp_kernel2.cuh"
