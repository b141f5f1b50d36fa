// laser_slam_b200/csrc/plicp2_search.cuh -- correspondence search of PL-ICP for one sens ray.
//
// find_correspondences_tricks (CSM icp_corr_tricks.c) is a SEQUENTIAL algorithm: ray i starts its walk at
// last_best+1, the j1 of the previous ray that got a valid correspondence, and its jump tables may skip a closer
// point, so the result depends on where the walk starts (the restatement differs from an exhaustive search on ~1 %
// of real scan pairs).  The kernel reproduces it exactly: every ray is searched with a SPECULATED last_best
// (the value the previous iteration saw, kept per ray in shared memory), then verified against its actual predecessor and re-searched until
// nothing changes -- the fixpoint is the sequential result (induction over i).
//
// The walk itself is restated without libm calls.  CSM stops a direction when
//     (rho * (dtheta > pi/2 ? 1 : sin(dtheta)))^2 > best_dist,      dtheta = |theta[j] - phi|
// and rho*|ref_j|*sin(dtheta) is the cross product of the world point with ref point j, so the test becomes
//     cross^2 > best_dist * |ref_j|^2         (or rho^2 > best_dist when dtheta > pi/2)
// which needs 4 multiplications instead of a double-precision sin().  `rr < rho` (the jump-table choice) is
// evaluated on the squares.  Both rewrites change the result only when a comparison is tied to ~1e-16 relative.
#pragma once
#include "plicp2_common.cuh"

namespace plicp2 {

struct Ray {
    double wx, wy;    // ld_compute_world_coords: R(x_old[2]) p + x_old[0:2]
    double rho2;      // |p_w|^2
    float phi;        // atan2(wy, wx), float accuracy (only used for the dtheta > pi/2 switch)
    int start_cell;   // (int)((phi - min_theta) * C1), exact
};

// world coordinates + start cell of sens ray with sensor-frame point p
__device__ __forceinline__ Ray make_ray(double2 p, double x0, double x1, double cs, double sn, double min_theta,
                                        double C1) {
    Ray r;
    r.wx = cs * p.x - sn * p.y + x0;
    r.wy = sn * p.x + cs * p.y + x1;
    r.rho2 = r.wx * r.wx + r.wy * r.wy;
    float phf = atan2f((float)r.wy, (float)r.wx);
    double cell = ((double)phf - min_theta) * C1;
    // atan2f is good to ~1e-6 rad; when that could move the truncation, decide with the double-precision atan2
    double fl = floor(cell);
    double fr = cell - fl;
    double thr = 2.5e-6 * C1 + 1e-9;
    if (fr < thr || fr > 1.0 - thr) cell = (atan2(r.wy, r.wx) - min_theta) * C1;
    r.phi = phf;
    r.start_cell = (int)cell;  // C truncation (toward zero), like the reference
    return r;
}

// Geometry a walk needs besides the ray: offsets of the ref tables in shared memory and the beam angles.
struct WalkCtx {
    unsigned refp, rr2, jt;  // shared-memory byte offsets
    int n;
    float fi_min, dfi;
    double maxd2;
};

// One ray's walk.  Returns j1 (a fully valid correspondence) or -1 and j2 in *j2out.
static __device__ __noinline__ int tricks_one(const WalkCtx c, const Ray r, int last_best, int *j2out) {
    const float HALF_PI_F = 1.57079632679489662f;
    const int n = c.n, to = n - 1;
    const double2 *refp = PL_SH(const double2, c.refp);
    const double *rr2a = PL_SH(const double, c.rr2);
    const short *jt = PL_SH(const short, c.jt);
    const double wx = r.wx, wy = r.wy, rho2 = r.rho2;
    const int start_cell = r.start_cell;
    int j1 = -1;
    double best = 42;
    int we_start_at = (last_best == -1) ? start_cell : last_best + 1;
    we_start_at = we_start_at > to ? to : we_start_at;
    we_start_at = we_start_at < 0 ? 0 : we_start_at;
    int up = we_start_at + 1, down = we_start_at;
    double last_up = 0, last_down = -1;
    bool up_stopped = false, down_stopped = false;
    while (!(up_stopped && down_stopped)) {
        const bool now_up = up_stopped ? false : down_stopped ? true : (last_up < last_down);
        const int idx = now_up ? up : down;
        if (now_up ? (idx > to) : (idx < 0)) {
            if (now_up) up_stopped = true; else down_stopped = true;
            continue;
        }
        // everything a step may need depends only on idx: issue all shared-memory loads up front so that their
        // latencies overlap instead of chaining behind the comparisons (the walk is latency-bound)
        const double rr2 = rr2a[idx];
        const double2 q = refp[idx];
        const short *jd = jt + (now_up ? 0 : 2 * n) + idx;
        const int hop_bigger = jd[0], hop_smaller = jd[n];
        int step = now_up ? 1 : -1;
        bool stop = false;
        if (rr2 >= 0) {
            const double d = dist2(wx, wy, q.x, q.y);
            if (now_up) last_up = d; else last_down = d;
            if (d < best || j1 == -1) { j1 = idx; best = d; }
            if (now_up ? (idx > start_cell) : (idx < start_cell)) {
                const float th = theta_f(c.fi_min, c.dfi, idx);
                const float dth = now_up ? th - r.phi : r.phi - th;
                const double cross = wx * q.y - wy * q.x;
                const bool far = dth > HALF_PI_F;
                const double lhs = far ? rho2 : cross * cross;
                const double rhs = far ? best : best * rr2;
                stop = lhs > rhs;
                step = rr2 < rho2 ? hop_bigger : hop_smaller;
            }
        }
        if (stop) {
            if (now_up) up_stopped = true; else down_stopped = true;
        } else {
            if (now_up) up = idx + step; else down = idx + step;
        }
    }
    if (j1 == -1 || best > c.maxd2) return -1;
    if (j1 == 0 || j1 == n - 1) return -1;
    int j2up = -1, j2down = -1;
    for (int j = j1 + 1; j < n; j++) if (rr2a[j] >= 0) { j2up = j; break; }
    for (int j = j1 - 1; j >= 0; j--) if (rr2a[j] >= 0) { j2down = j; break; }
    if (j2up == -1 && j2down == -1) return -1;
    int j2;
    if (j2up == -1) j2 = j2down;
    else if (j2down == -1) j2 = j2up;
    else {
        const double2 qu = refp[j2up], qd = refp[j2down];
        j2 = dist2(wx, wy, qu.x, qu.y) < dist2(wx, wy, qd.x, qd.y) ? j2up : j2down;
    }
    *j2out = j2;
    return j1;
}

// ld_create_jump_tables(laser_ref) (CSM laser_data.c): for every valid ref ray i
//   up_bigger[i]   = (first j > i that is invalid or has r[j] > r[i]) - i      (n - i when there is none)
//   up_smaller[i]  = (first j > i that is invalid or has r[j] < r[i]) - i
//   down_bigger / down_smaller the same towards j < i                          (-1 - i when there is none)
// A jump can be hundreds of rays long (a wall seen head-on: r falls to the foot of the perpendicular and rises again),
// so a linear scan per entry is O(n * jump).  Every entry is searched independently here over a 3-level structure:
// the rays, the extrema of aligned blocks of 8 rays and of 64 rays (JumpAux, filled while the scans are loaded).
// A search runs to the end of its 8-block, then over 8-blocks to the end of its 64-block, then over 64-blocks, and
// descends again: at most 7+7+nb64+8+8 steps (47 at 1080 rays), 1-3 for the common short jumps.  Entries of invalid
// rays are never read by the walk (it steps by +-1 there) and are set to +-1.
struct JumpAux {
    unsigned hi8, lo8, hi64, lo64;  // shared-memory byte offsets (inside the d2j area, free until iteration 0)
    int nb8, nb64;
};

__device__ __forceinline__ JumpAux make_jump_aux(const Lay &l, int n) {
    JumpAux a;
    a.nb8 = (n + 7) >> 3;
    a.nb64 = (a.nb8 + 7) >> 3;
    a.hi8 = l.d2j;
    a.lo8 = a.hi8 + 8u * a.nb8;
    a.hi64 = a.lo8 + 8u * a.nb8;
    a.lo64 = a.hi64 + 8u * a.nb64;
    return a;
}

// BIG: the search stops at an invalid ray (rr2 = -1) or a bigger one; else at an invalid or a smaller one
template <bool BIG>
__device__ __forceinline__ bool jt_stop(double k, double v) { return BIG ? (k < 0 || k > v) : (k < v); }
// block extrema: BIG uses max(valid ? rr2 : +inf), else min(rr2) (an invalid ray makes it -1 < v)
template <bool BIG>
__device__ __forceinline__ bool jt_blk(double m, double v) { return BIG ? (m > v) : (m < v); }

template <bool BIG>
__device__ __forceinline__ int jt_search_up(const double *rr2a, const double *m8, const double *m64, int n, int nb8,
                                            int nb64, int i, double v) {
    int j = i + 1;
    while (j < n && (j & 7)) { if (jt_stop<BIG>(rr2a[j], v)) return j; j++; }
    if (j >= n) return n;
    int b = j >> 3;
    bool hit = false;
    while (b < nb8 && (b & 7)) { if (jt_blk<BIG>(m8[b], v)) { hit = true; break; } b++; }
    if (!hit) {
        if (b >= nb8) return n;
        int c = b >> 3;
        while (c < nb64 && !jt_blk<BIG>(m64[c], v)) c++;
        if (c >= nb64) return n;
        b = c << 3;
        while (!jt_blk<BIG>(m8[b], v)) b++;  // the block extrema only cover existing rays: stops inside the data
    }
    j = b << 3;
    while (!jt_stop<BIG>(rr2a[j], v)) j++;
    return j;
}

template <bool BIG>
__device__ __forceinline__ int jt_search_down(const double *rr2a, const double *m8, const double *m64, int i, double v) {
    int j = i - 1;
    while (j >= 0 && (j & 7) != 7) { if (jt_stop<BIG>(rr2a[j], v)) return j; j--; }
    if (j < 0) return -1;
    int b = j >> 3;
    bool hit = false;
    while (b >= 0 && (b & 7) != 7) { if (jt_blk<BIG>(m8[b], v)) { hit = true; break; } b--; }
    if (!hit) {
        if (b < 0) return -1;
        int c = b >> 3;
        while (c >= 0 && !jt_blk<BIG>(m64[c], v)) c--;
        if (c < 0) return -1;
        b = (c << 3) + 7;
        while (!jt_blk<BIG>(m8[b], v)) b--;
    }
    j = (b << 3) + 7;  // every block at or below the start block is complete
    while (!jt_stop<BIG>(rr2a[j], v)) j--;
    return j;
}

// hi64 / lo64 from hi8 / lo8 (one thread per 64-block); the caller synchronises before and after
__device__ __forceinline__ void jump_aux_level2(const JumpAux &ja, int t) {
    if (t < ja.nb64) {
        const double *hi8 = PL_SH(const double, ja.hi8), *lo8 = PL_SH(const double, ja.lo8);
        double hi = -1.0, lo = INFINITY;
        const int e = (t << 3) + 8 < ja.nb8 ? (t << 3) + 8 : ja.nb8;
        for (int b = t << 3; b < e; b++) { hi = fmax(hi, hi8[b]); lo = fmin(lo, lo8[b]); }
        PL_SH(double, ja.hi64)[t] = hi;
        PL_SH(double, ja.lo64)[t] = lo;
    }
}

__device__ __forceinline__ void build_jump_tables(const Lay &l, const JumpAux &ja, int n, int t, int nthreads) {
    short *jt = PL_SH(short, l.jt);
    const double *rr2a = PL_SH(const double, l.rr2);
    const double *hi8 = PL_SH(const double, ja.hi8), *lo8 = PL_SH(const double, ja.lo8);
    const double *hi64 = PL_SH(const double, ja.hi64), *lo64 = PL_SH(const double, ja.lo64);
#pragma unroll 1
    for (int i = t; i < n; i += nthreads) {
        const double v = rr2a[i];  // squares are monotone in r >= 0; invalid = -1
        int ub = i + 1, us = i + 1, db = i - 1, ds = i - 1;
        if (v >= 0) {
            ub = jt_search_up<true>(rr2a, hi8, hi64, n, ja.nb8, ja.nb64, i, v);
            us = jt_search_up<false>(rr2a, lo8, lo64, n, ja.nb8, ja.nb64, i, v);
            db = jt_search_down<true>(rr2a, hi8, hi64, i, v);
            ds = jt_search_down<false>(rr2a, lo8, lo64, i, v);
        }
        jt[0 * n + i] = (short)(ub - i);
        jt[1 * n + i] = (short)(us - i);
        jt[2 * n + i] = (short)(db - i);
        jt[3 * n + i] = (short)(ds - i);
    }
}

}  // namespace plicp2
