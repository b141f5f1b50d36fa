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

}  // namespace plicp2
