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

// Geometry a walk needs besides the ray: 32-bit shared-window addresses of the ref tables and the beam angles.
struct WalkCtx {
    unsigned refp, rr2, jt;  // shared-memory addresses (cvta.to.shared)
    int n;
    float fi_min, dfi;
    double maxd2;
};

// The tables are constant while the walks run: plain (non-volatile) asm lets the compiler schedule the loads freely,
// and the 32-bit shared addresses keep the address arithmetic out of the 64-bit generic path.
__device__ __forceinline__ double lds_f64(unsigned a) {
    double v;
    asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ double2 lds_f64x2(unsigned a) {
    double2 v;
    asm("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
}
// (ptxas sinks this load to its conditional use whatever the source order; ld.volatile keeps it early but
// serialises the step -- measured 27 % slower)
__device__ __forceinline__ int lds_b32(unsigned a) {
    int v;
    asm("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}

#ifdef PLICP_WALK_STATS
static __device__ int g_walk_trips_dummy;
#define g_walk_trips (*trips_out)
#endif

constexpr int WALK_BUDGET = 12;

// One ray's walk.  Returns j1 (a fully valid correspondence) or -1 and j2 in *j2out.
// State: up / down = next index to probe in each direction; stopped bit 0 / 1 = that direction has ended.
// CSM initialises last_dist_up = 0 and last_dist_down = -1; the only use is "last_up < last_down", false for both
// (0, -1) and (0, 0), so 0 serves for both.
static __device__ __noinline__ int tricks_one(const WalkCtx c, const Ray r, int last_best, int *j2out
#ifdef PLICP_WALK_STATS
                                              , int *trips_out
#endif
) {
    const float HALF_PI_F = 1.57079632679489662f;
    const int n = c.n, to = n - 1;
    const unsigned a_rr2 = c.rr2, a_refp = c.refp;
    const unsigned a_jt_up = c.jt, a_jt_dn = c.jt + 4u * (unsigned)n;  // [2][n] pairs (bigger, smaller)
    const double wx = r.wx, wy = r.wy, rho2 = r.rho2;
    const int start_cell = r.start_cell;
    int j1 = -1;
    double best = INFINITY;  // CSM: best_dist = 42 with "j1 == -1 ||" in front of every comparison: the same
    int we_start_at = (last_best == -1) ? start_cell : last_best + 1;
    we_start_at = we_start_at > to ? to : we_start_at;
    we_start_at = we_start_at < 0 ? 0 : we_start_at;
    int up = we_start_at + 1, down = we_start_at;
    double last_up = 0, last_down = 0;
    int stopped = 0;
#ifdef PLICP_WALK_STATS
    int trips = 0;
#endif
    const float fi_min = c.fi_min, dfi = c.dfi, phi = r.phi;
    // The general loop: the next probe goes where the last distance was smaller (or where the walk is still alive).
    // A walk that is still running after WALK_BUDGET probes with one direction ended leaves for the loop below.
    int budget = WALK_BUDGET;
#pragma unroll 1
    while (stopped != 3) {
#ifdef PLICP_WALK_STATS
        trips++;
#endif
        if (--budget < 0 && stopped != 0) break;
        const bool now_up = (stopped == 2) | ((stopped == 0) & (last_up < last_down));
        const int idx = now_up ? up : down;
        const int bit = now_up ? 1 : 2;
        if ((unsigned)idx > (unsigned)to) {  // ran off either end of the scan
            stopped |= bit;
            continue;
        }
        const double rr2 = lds_f64(a_rr2 + 8u * idx);
        const double2 q = lds_f64x2(a_refp + 16u * idx);
        const int hops = lds_b32((now_up ? a_jt_up : a_jt_dn) + 4u * idx);  // both jumps of this direction
        const int sdir = now_up ? 1 : -1;
        int nidx = idx + sdir;
        if (rr2 >= 0) {
            const double d = dist2(wx, wy, q.x, q.y);
            last_up = now_up ? d : last_up;
            last_down = now_up ? last_down : d;
            const bool better = d < best;
            j1 = better ? idx : j1;
            best = better ? d : best;
            if ((idx - start_cell) * sdir > 0) {  // beyond the start cell in the walking direction
                const float dth = theta_f(fi_min, dfi, idx) - phi;  // phi - theta is the exact negation
                const float sdth = now_up ? dth : -dth;
                const double cross = wx * q.y - wy * q.x;
                // stop = dth > pi/2 ? rho2 > best : cross^2 > best * rr2, kept in predicate registers
                int stop;
                asm("{ .reg .pred pf, pa, pb;\n\t"
                    "setp.gt.f32 pf, %1, %2;\n\t"
                    "setp.gt.and.f64 pa, %3, %4, pf;\n\t"
                    "not.pred pf, pf;\n\t"
                    "setp.gt.and.f64 pb, %5, %6, pf;\n\t"
                    "or.pred pa, pa, pb;\n\t"
                    "selp.s32 %0, 1, 0, pa; }"
                    : "=r"(stop)
                    : "f"(sdth), "f"(HALF_PI_F), "d"(rho2), "d"(best), "d"(cross * cross), "d"(best * rr2));
                if (stop) { stopped |= bit; continue; }
                nidx = idx + (rr2 < rho2 ? (int)(short)hops : hops >> 16);
            }
        } else {
            nidx = idx + (hops >> 16);  // an invalid ref ray: its entries hold the skip over the whole invalid run
        }
        up = now_up ? nidx : up;
        down = now_up ? down : nidx;
    }
    // One direction has ended and the other runs on alone: the same loop body with the direction fixed (no selects, no
    // last distances).  The rare walks that cross most of the scan (about 1 in 600, 100+ probes) spend nearly all
    // their probes here with one lane of the warp alive, and a whole search phase waits for them: instructions per
    // probe are what counts.  Short walks never come here, so the warp does not serialise the two loops.
    if (stopped != 3) {
        const bool go_up = stopped == 2;
        const int sdir = go_up ? 1 : -1;
        const unsigned a_jt = go_up ? a_jt_up : a_jt_dn;
        int idx = go_up ? up : down;
#pragma unroll 1
        while ((unsigned)idx <= (unsigned)to) {
#ifdef PLICP_WALK_STATS
            trips += 0x10001;
#endif
            const double rr2 = lds_f64(a_rr2 + 8u * idx);
            const double2 q = lds_f64x2(a_refp + 16u * idx);
            const int hops = lds_b32(a_jt + 4u * idx);
            int hop = hops >> 16;  // an invalid ref ray: the skip over its whole run
            if (rr2 >= 0) {
                const double d = dist2(wx, wy, q.x, q.y);
                const bool better = d < best;
                j1 = better ? idx : j1;
                best = better ? d : best;
                hop = sdir;
                if ((idx - start_cell) * sdir > 0) {
                    const float dth = theta_f(fi_min, dfi, idx) - phi;
                    const float sdth = go_up ? dth : -dth;
                    if (sdth > HALF_PI_F) {
                        if (rho2 > best) break;
                    } else {
                        const double cross = wx * q.y - wy * q.x;
                        if (cross * cross > best * rr2) break;
                    }
                    hop = rr2 < rho2 ? (int)(short)hops : (hops >> 16);
                }
            }
            idx += hop;
        }
    }
#ifdef PLICP_WALK_STATS
    g_walk_trips = trips;
#endif
    if (j1 == -1 || best > c.maxd2) return -1;
    if (j1 == 0 || j1 == n - 1) return -1;
    // nearest valid neighbours of j1 (0 < j1 < n-1): an invalid neighbour's table entry skips its whole run
    int j2up = j1 + 1, j2down = j1 - 1;
    if (lds_f64(a_rr2 + 8u * j2up) < 0) j2up += lds_b32(a_jt_up + 4u * j2up) >> 16;
    if (lds_f64(a_rr2 + 8u * j2down) < 0) j2down += lds_b32(a_jt_dn + 4u * j2down) >> 16;
    if (j2up >= n) j2up = -1;
    if (j2down < 0) j2down = -1;
    if (j2up == -1 && j2down == -1) return -1;
    int j2;
    if (j2up == -1) j2 = j2down;
    else if (j2down == -1) j2 = j2up;
    else {
        const double2 qu = lds_f64x2(a_refp + 16u * j2up), qd = lds_f64x2(a_refp + 16u * j2down);
        j2 = dist2(wx, wy, qu.x, qu.y) < dist2(wx, wy, qd.x, qd.y) ? j2up : j2down;
    }
    *j2out = j2;
    return j1;
}

}  // namespace plicp2
