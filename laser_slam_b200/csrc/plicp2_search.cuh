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
//
// Contents: make_ray (world point + start cell of a sens ray), walk_finish (validity rules + j2), tricks_one (one
// ray's walk by its own lane), and -- for the long-scan instances (LIST, T >= 256) -- the warp-wide continuation of
// walks that run for hundreds of probes (load_window, chain_one, walk_warp_one, walk_warp_both).
#pragma once
#include "plicp2_common.cuh"

namespace PL_NAMESPACE {

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

// pose of the current iteration + what make_ray needs
struct IterCtx {
    double x0, x1, cs, sn;
    double min_theta, C1;
    const double2 *cs2;  // (cos, sin) of the beam angles (handle table, global memory)
};

// Geometry a walk needs besides the ray: 32-bit shared-window addresses of the ref tables and the beam angles.
struct WalkCtx {
    unsigned refp, rr2, jt;  // shared-memory addresses (cvta.to.shared)
    int n;
    float fi_min, dfi;
    double maxd2;
    int kq;  // a probe fewer than kq rays beyond the start cell has dtheta < pi/2 for certain (walk_kq)
};

// CSM's stop test switches form where dtheta = |theta[j] - phi| exceeds pi/2.  With start_cell = (int)((phi - theta_0) *
// n / (theta_{n-1} - theta_0)) exact (make_ray) and theta[j] = theta_0 + dfi * j, a probe `rel` rays beyond the start
// cell in its walking direction has dtheta <= dfi * (rel + 2) * (1 + 1e-6): below kq = pi/2 / dfi - 4 the switch cannot
// fire and the float angle arithmetic is skipped; at or above it the exact float test decides, as before.
__host__ __device__ inline int walk_kq(float dfi) {
    const float k = 1.5707963f / dfi - 4.f;
    return k > 0.f ? (k < 1e6f ? (int)k : 1000000) : 0;
}

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
// The three loads of a probe: squared reading, point, the direction's jump pair.  ptxas sinks a load behind any branch
// whose body alone uses it -- in the r01 build the point's load sat behind the validity branch on the reading, two
// shared-memory latencies in a row per probe (cuobjdump) -- so the callers' validity test reads the point as well
// (an invalid ref ray's point is NaN), and the three loads issue back to back.
__device__ __forceinline__ void lds_probe(unsigned a_rr2, unsigned a_q, unsigned a_hops, double &rr2, double2 &q, int &hops) {
    asm("ld.shared.f64 %0, [%4];\n\t"
        "ld.shared.v2.f64 {%1, %2}, [%5];\n\t"
        "ld.shared.b32 %3, [%6];"
        : "=d"(rr2), "=d"(q.x), "=d"(q.y), "=r"(hops)
        : "r"(a_rr2), "r"(a_q), "r"(a_hops));
}

#ifdef PLICP_WALK_STATS
static __device__ int g_walk_trips_dummy;
#define g_walk_trips (*trips_out)
#endif

// The end of every walk: the validity rules and j2.  Returns j1 (a fully valid correspondence) or -1, j2 in *j2out.
__device__ __forceinline__ int walk_finish(const WalkCtx &c, double wx, double wy, int j1, double best, int *j2out) {
    const int n = c.n;
    const unsigned a_rr2 = c.rr2, a_refp = c.refp;
    const unsigned a_jt_up = c.jt, a_jt_dn = c.jt + 4u * (unsigned)n;
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

constexpr int WALK_BUDGET = 12;    // probes of the general loop before a walk with one side ended leaves it
#ifndef PLICP_BOTH_AFTER
#define PLICP_BOTH_AFTER 100
#endif
#ifndef PLICP_SOLO_BUDGET
#define PLICP_SOLO_BUDGET 32
#endif
constexpr int BOTH_AFTER = PLICP_BOTH_AFTER;     // LIST: further probes with both sides alive before the walk goes on the list
constexpr int SOLO_BUDGET = PLICP_SOLO_BUDGET;    // LIST: probes of the one-direction loop before the walk goes on the list
constexpr int WALK_PENDING = -2;   // tricks_one returns this when the walk has been put on the list

// A walk that is still running after the budgets above is not finished by its own lane -- hundreds of dependent
// probes while the rest of the CTA waits at the next barrier -- but put on a list in shared memory in this state.
// After the barrier that ends the round of walks the warps take the list's entries in turn and finish each one
// warp-wide (process_pending in plicp2_kernel.cuh, walk_warp_one / walk_warp_both below).  The common path pays
// nothing for it: no warp-wide call protocol, one counter read per round.  A full list: the lane walks on.
struct WalkState {
    double best, last_up, last_down;
    int i;         // the sens ray
    int j1;
    int up, down;  // next index to probe in each direction (dir = +-1: `up` holds the one index)
    int dir;       // +1 / -1: one direction left; 2: both directions alive
    int pad_;      // steps the walk had executed when it was handed over
};
static_assert(sizeof(WalkState) == 48, "WalkState is a 48-byte shared-memory record");

// Where the list is: only the long-scan instances (LIST) have one, the others carry nothing around.
template <bool LIST>
struct PendRef {};
template <>
struct PendRef<true> {
    unsigned pend, npend;  // shared-memory addresses of the list and of its counter
    int pend_cap;          // entries the list can hold
};

// reserve a list entry and write the state; false when the list is full
__device__ __forceinline__ bool pend_push(const PendRef<false> &, const WalkState &) { return false; }
__device__ __forceinline__ bool pend_push(const PendRef<true> &c, const WalkState &st) {
    int slot;
    asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(slot) : "r"(c.npend) : "memory");
    if (slot >= c.pend_cap) return false;
    const unsigned a = c.pend + 48u * (unsigned)slot;
    asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(a), "d"(st.best), "d"(st.last_up) : "memory");
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(a + 16u), "d"(st.last_down) : "memory");
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(a + 24u), "r"(st.i), "r"(st.j1) : "memory");
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a + 32u), "r"(st.up), "r"(st.down), "r"(st.dir), "r"(st.pad_) : "memory");
    return true;
}

template <bool LIST>
static __device__ __noinline__ int tricks_one(const WalkCtx c, const Ray r, int last_best, int *j2out,
                                              const PendRef<LIST> pr, int ray_i
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
    const int kq = c.kq;
    // The general loop: the next probe goes where the last distance was smaller (or where the walk is still alive).
    // A walk that is still running after WALK_BUDGET probes with one direction ended leaves for the loop below.
    int budget = WALK_BUDGET;
    bool ask = LIST;  // may this walk be put on the list?
#pragma unroll 1
    while (stopped != 3) {
#ifdef PLICP_WALK_STATS
        trips++;
#endif
        if (--budget < 0) {
            if (stopped != 0) break;
            if (LIST && budget < -BOTH_AFTER && ask) {
                WalkState st;
                st.best = best; st.last_up = last_up; st.last_down = last_down;
                st.i = ray_i; st.j1 = j1; st.up = up; st.down = down; st.dir = 2;
                st.pad_ = WALK_BUDGET - budget;
                if (pend_push(pr, st)) {
#ifdef PLICP_WALK_STATS
                    g_walk_trips = trips;
#endif
                    return WALK_PENDING;
                }
                ask = false;
            }
        }
        const bool now_up = (stopped == 2) | ((stopped == 0) & (last_up < last_down));
        const int idx = now_up ? up : down;
        const int bit = now_up ? 1 : 2;
        if ((unsigned)idx > (unsigned)to) {  // ran off either end of the scan
            stopped |= bit;
            continue;
        }
        double rr2;
        double2 q;
        int hops;  // both jumps of this direction
        lds_probe(a_rr2 + 8u * idx, a_refp + 16u * idx, (now_up ? a_jt_up : a_jt_dn) + 4u * idx, rr2, q, hops);
        const int sdir = now_up ? 1 : -1;
        int nidx = idx + sdir;
        // an invalid ref ray has rr2 = -1 AND a NaN point: testing both makes the branch need the point, which keeps
        // ptxas from sinking the point's load behind this branch (the second of two latencies in a row)
        if ((rr2 >= 0) & (q.x == q.x)) {
            const double d = dist2(wx, wy, q.x, q.y);
            last_up = now_up ? d : last_up;
            last_down = now_up ? last_down : d;
            const bool better = d < best;
            j1 = better ? idx : j1;
            best = better ? d : best;
            const int rel = (idx - start_cell) * sdir;
            if (rel > 0) {  // beyond the start cell in the walking direction
                const double cross = wx * q.y - wy * q.x;
                // stop = dth > pi/2 ? rho2 > best : cross^2 > best * rr2
                bool stop = cross * cross > best * rr2;
                if (rel >= kq) {  // only here can dth exceed pi/2 (walk_kq): the exact float test, as the reference orders it
                    const float dth = theta_f(fi_min, dfi, idx) - phi;  // phi - theta is the exact negation
                    if ((now_up ? dth : -dth) > HALF_PI_F) stop = rho2 > best;
                }
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
    int solo = SOLO_BUDGET;
    if (stopped != 3) {
        const bool go_up = stopped == 2;
        const int sdir = go_up ? 1 : -1;
        const unsigned a_jt = go_up ? a_jt_up : a_jt_dn;
        int idx = go_up ? up : down;
#pragma unroll 1
        while ((unsigned)idx <= (unsigned)to) {
            --solo;  // (counted in every instance: the walk's step count, see the return value)
            if (LIST && solo < 0 && ask) {
                WalkState st;
                st.best = best; st.last_up = 0; st.last_down = 0;
                st.i = ray_i; st.j1 = j1; st.up = idx; st.down = 0; st.dir = sdir;
                st.pad_ = (WALK_BUDGET - budget) + (SOLO_BUDGET - solo);
                if (pend_push(pr, st)) {
#ifdef PLICP_WALK_STATS
                    g_walk_trips = trips;
#endif
                    return WALK_PENDING;
                }
                ask = false;
            }
#ifdef PLICP_WALK_STATS
            trips += 0x10001;
#endif
            double rr2;
            double2 q;
            int hops;
            lds_probe(a_rr2 + 8u * idx, a_refp + 16u * idx, a_jt + 4u * idx, rr2, q, hops);
            int hop = hops >> 16;  // an invalid ref ray: the skip over its whole run
            if ((rr2 >= 0) & (q.x == q.x)) {
                const double d = dist2(wx, wy, q.x, q.y);
                const bool better = d < best;
                j1 = better ? idx : j1;
                best = better ? d : best;
                hop = sdir;
                const int rel = (idx - start_cell) * sdir;
                if (rel > 0) {
                    const double cross = wx * q.y - wy * q.x;
                    bool stop = cross * cross > best * rr2;
                    if (rel >= kq) {
                        const float dth = theta_f(fi_min, dfi, idx) - phi;
                        if ((go_up ? dth : -dth) > HALF_PI_F) stop = rho2 > best;
                    }
                    if (stop) break;
                    hop = rr2 < rho2 ? (int)(short)hops : (hops >> 16);
                }
            }
            idx += hop;
        }
    }
#ifdef PLICP_WALK_STATS
    g_walk_trips = trips;
#endif
    // *j2out: j2 in the low half, the walk's executed steps (trips of the two loops: the budgets count them) above
    const int r1 = walk_finish(c, wx, wy, j1, best, j2out);
    *j2out = (*j2out & 0xffff) | (((WALK_BUDGET - budget) + (SOLO_BUDGET - solo)) << 16);
    return r1;
}

// ---------------------------------------------------------------------------------------------------------------
// LIST: the rest of a long walk, run by all 32 lanes of a warp for ONE ray (a list entry: all arguments uniform).
//
// The indices a direction probes do not depend on the distances found: from idx the walk goes to idx + hop(idx), hop
// being a unit step before the start cell, a jump-table entry chosen by `rr2 < rho2` after it, the skip over the run
// for an invalid ray.  Only where a direction STOPS depends on the best distance so far, and which direction goes
// next depends on the last distance of each.  So the warp looks at 32 consecutive indices of a direction at once
// (a window: lane k has index base + k*dir): every lane loads its ref point and computes its distance and its hop,
// pointer doubling over the hops marks the lanes the walk really visits, a prefix minimum over those gives each
// probe the best distance it would have seen, and ballots find the first lane where something happens.  All
// decisions are the sequential loop's, on the same operands: strict `<` for the best (the earliest probe keeps a tie),
// the stop test after the probe's own distance has been taken in.
struct RayU {  // lane L's ray, warp-uniform
    double wx, wy, rho2;
    float phi;
    int start_cell;
};

struct Win {
    double d;       // own lane: squared distance (INF unless visited and valid)
    double pm;      // own lane: minimum of d over the lanes up to and including this one
    double rr2, cross2;  // own lane: operands of the stop test
    int my;         // own lane: ref index
    int next;       // own lane: my + hop
    bool beyond;    // own lane: past the start cell in the walking direction (the stop test applies)
    bool wide;      // own lane: dtheta > pi/2 (the stop test is rho2 > best)
    unsigned visited, validm, oobm;  // uniform: lanes the walk visits / of those, valid rays / beyond the scan's end
    int exit_idx;   // uniform: where the walk goes after its last visited lane
};

// cb: the best distance so far (it only falls afterwards)
__device__ __forceinline__ Win load_window(const WalkCtx &c, const RayU &u, int base, int sdir, double cb) {
    const unsigned FULL = 0xffffffffu;
    const float HALF_PI_F = 1.57079632679489662f;
    const int lane = threadIdx.x & 31;
    const int to = c.n - 1;
    const unsigned a_jt = sdir > 0 ? c.jt : c.jt + 4u * (unsigned)c.n;
    Win w;
    w.my = base + sdir * lane;
    const bool inb = (unsigned)w.my <= (unsigned)to;
    double2 q = make_double2(0.0, 0.0);
    int hops = 0;
    w.rr2 = -1.0;
    if (inb) {
        w.rr2 = lds_f64(c.rr2 + 8u * w.my);
        q = lds_f64x2(c.refp + 16u * w.my);
        hops = lds_b32(a_jt + 4u * w.my);
    }
    const bool valid = inb && w.rr2 >= 0;
    w.beyond = (w.my - u.start_cell) * sdir > 0;
    int hop = hops >> 16;  // an invalid ref ray: the skip over its whole run
    if (valid) hop = w.beyond ? (w.rr2 < u.rho2 ? (int)(short)hops : (hops >> 16)) : sdir;
    w.next = w.my + hop;
    int nx = inb ? lane + hop * sdir : 32;
    nx = nx > 32 ? 32 : nx;
    unsigned R = 1u;
    const unsigned inbm = __ballot_sync(FULL, inb);
    if (__all_sync(FULL, !inb || nx == lane + 1)) {
        // unit steps only (always so before the start cell): every ray of the window is visited, and the first index
        // past the end of the scan if the window reaches it (the in-bounds lanes are a prefix)
        R = inbm | (inbm + 1u);
    } else {
#pragma unroll
        for (int k = 0; k < 5; k++) {
            const unsigned add = ((R >> lane) & 1u) && nx < 32 ? (1u << nx) : 0u;
            R |= __reduce_or_sync(FULL, add);
            const int nx2 = __shfl_sync(FULL, nx, nx & 31);
            nx = nx < 32 ? nx2 : 32;
        }
    }
    w.visited = R;
    const bool vis = (R >> lane) & 1u;
    w.validm = __ballot_sync(FULL, vis && valid);
    w.oobm = __ballot_sync(FULL, vis && !inb);
    w.d = (vis && valid) ? dist2(u.wx, u.wy, q.x, q.y) : INFINITY;
    w.pm = INFINITY;
    if (__any_sync(FULL, w.d < cb)) {  // else no probe of this window can ever lower the best: min(pm, best) = best
        w.pm = w.d;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double v = __shfl_up_sync(FULL, w.pm, o);
            if (lane >= o) w.pm = v < w.pm ? v : w.pm;
        }
    }
    const float dth = theta_f(c.fi_min, c.dfi, w.my) - u.phi;
    w.wide = (sdir > 0 ? dth : -dth) > HALF_PI_F;
    const double cross = u.wx * q.y - u.wy * q.x;
    w.cross2 = cross * cross;
    w.exit_idx = __shfl_sync(FULL, w.next, 31 - __clz(R));
    return w;
}

// own lane: does the stop test fire when the best distance is `best` at this probe?
__device__ __forceinline__ bool win_stop(const Win &w, const RayU &u, int lane, double best) {
    if (!((w.validm >> lane) & 1u) || !w.beyond) return false;
    const double bi = w.pm < best ? w.pm : best;
    return w.wide ? (u.rho2 > bi) : (w.cross2 > bi * w.rr2);
}

// take the lanes [from .. last] of a window into (cb, cj)
__device__ __forceinline__ void win_take(const Win &w, int lane, int from, int last, double &cb, int &cj) {
    const unsigned FULL = 0xffffffffu;
    const double nb = __shfl_sync(FULL, w.pm, last);
    const unsigned holders = __ballot_sync(FULL, w.d == nb && lane >= from && lane <= last);
    if (nb < cb && holders) {
        cb = nb;
        cj = __shfl_sync(FULL, w.my, __ffs(holders) - 1);
    }
}

// One direction alone, from idx, until it stops or leaves the scan.
__device__ __forceinline__ void chain_one(const WalkCtx &c, const RayU &u, int idx, int sdir, double &cb, int &cj, int &nwin) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int to = c.n - 1;
#pragma unroll 1
    while ((unsigned)idx <= (unsigned)to) {
        nwin++;
        const Win w = load_window(c, u, idx, sdir, cb);
        const unsigned ends = __ballot_sync(FULL, win_stop(w, u, lane, cb)) | w.oobm;
        const int last = ends ? __ffs(ends) - 1 : 31;
        win_take(w, lane, 0, last, cb, cj);
        if (ends) break;
        idx = w.exit_idx;
    }
}

// All arguments warp-uniform (a list entry and its ray); the result comes back in (cb, cj).
static __device__ __noinline__ int walk_warp_one(const WalkCtx c, const RayU u, int idx, int sdir, double *cbp, int *cjp) {
    double cb = *cbp;
    int cj = *cjp;
    int nwin = 0;
    chain_one(c, u, idx, sdir, cb, cj, nwin);
    *cbp = cb;
    *cjp = cj;
    return nwin;  // windows of 32 rays loaded (every lane evaluates its ray's distance)
}

// Both directions alive.  The sequential loop probes upwards while last_up < last_down and downwards otherwise, so
// it moves in RUNS: an upward run ends with the first valid probe whose distance is not below last_down, a downward
// run with the first whose distance is above last_up.  A run is a stretch of one window: its end is a ballot, the
// stop tests inside it see min(best before the run, prefix minimum), and the first of {run end, stop, end of the
// scan} decides how it goes on: with the other direction, with the other direction alone (chain_one), or with the
// next window of the same direction.
static __device__ __noinline__ int walk_warp_both(const WalkCtx c, const RayU u, int up, int down, double last_up,
                                                  double last_down, double *cbp, int *cjp) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    double cb = *cbp;
    int cj = *cjp;
    double last[2] = {last_down, last_up};  // [0] down, [1] up
    int nextidx[2] = {down, up};
    const int to = c.n - 1;
    Win W[2];
    int pos[2] = {32, 32};  // next lane to take in each window; 32: the window has to be loaded (from nextidx)
    int alone = -1;         // the direction that goes on alone once the other has ended
    int nwin = 0;
#pragma unroll 1
    for (;;) {
        const int s = last[1] < last[0] ? 1 : 0;  // 1: upwards
        const int sdir = s ? 1 : -1;
        if (pos[s] == 32) {
            if ((unsigned)nextidx[s] > (unsigned)to) { alone = s ^ 1; break; }  // ran off the scan: this side has ended
            if (s) W[1] = load_window(c, u, nextidx[1], +1, cb); else W[0] = load_window(c, u, nextidx[0], -1, cb);
            pos[s] = 0;
            nwin++;
        }
        const Win &w = s ? W[1] : W[0];
        const int from = pos[s];
        const unsigned ge = FULL << from;
        const double thr = last[s ^ 1];
        const bool mine_valid = (w.validm >> lane) & 1u;
        const unsigned endm = __ballot_sync(FULL, mine_valid && (s ? !(w.d < thr) : (thr < w.d))) & ge;
        const unsigned stopm = __ballot_sync(FULL, win_stop(w, u, lane, cb)) & ge;
        const unsigned oobm = w.oobm & ge;
        const int e_run = endm ? __ffs(endm) - 1 : 32, e_stop = stopm ? __ffs(stopm) - 1 : 32;
        const int e_oob = oobm ? __ffs(oobm) - 1 : 32;
        if (e_oob < e_run && e_oob < e_stop) {  // the visited lanes before the end of the scan, then this side has ended
            if (e_oob > from) win_take(w, lane, from, e_oob - 1, cb, cj);
            alone = s ^ 1;
            break;
        }
        if (e_stop <= e_run && e_stop < 32) {  // a stop test fired (its probe counts): this side has ended
            win_take(w, lane, from, e_stop, cb, cj);
            alone = s ^ 1;
            break;
        }
        if (e_run < 32) {  // the run ends inside the window: the other direction goes next
            win_take(w, lane, from, e_run, cb, cj);
            last[s] = __shfl_sync(FULL, w.d, e_run);
            const unsigned rest = w.visited & ~((2u << e_run) - 1u);
            if (rest) pos[s] = __ffs(rest) - 1;
            else { pos[s] = 32; nextidx[s] = w.exit_idx; }
            continue;
        }
        // the run takes the rest of the window and goes on in the next one
        win_take(w, lane, from, 31, cb, cj);
        const unsigned took = w.validm & ge;
        if (took) last[s] = __shfl_sync(FULL, w.d, 31 - __clz(took));
        pos[s] = 32;
        nextidx[s] = w.exit_idx;
        (void)sdir;
    }
    // one direction is left: from where it stands, alone
    {
        const int idx = pos[alone] < 32 ? __shfl_sync(FULL, (alone ? W[1].my : W[0].my), pos[alone]) : nextidx[alone];
        chain_one(c, u, idx, alone ? 1 : -1, cb, cj, nwin);
    }
    *cbp = cb;
    *cjp = cj;
    return nwin;
}

}  // namespace PL_NAMESPACE
