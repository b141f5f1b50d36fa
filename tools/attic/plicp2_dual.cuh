// NOT PART OF THE BUILD (tools/attic): the two-walks-per-thread experiment of round 2, kept for the record.  Measured
// SLOWER than one walk per call (profiles/r03l_dual_walks.txt: 1080 beams 17.6-18.1 ms against 16.0 ms per 20k pairs,
// 840 beams 13.8-13.9 against 11.8, 361 beams 21.3 against 17.3 ms per 60k pairs; records identical): the straight-line
// probe issues every instruction of both walks on every trip, and the walk phases are bound by issued instructions,
// not by the latency of a probe's chain.  To try it again: move this file to laser_slam_b200/csrc/, include it from
// plicp2_kernel.cuh and call search_two (git show of the experiment's commit message names the call sites).
//
// laser_slam_b200/csrc/plicp2_dual.cuh -- TWO correspondence walks per thread, interleaved (the 128-thread instance).
//
// One probe of a walk is a chain of ~25 dependent operations (shared-memory loads, FP64 arithmetic, selects): ~480
// cycles on an otherwise empty SM, 577 with four resident jobs (profiles/r02o_walk.txt), for 62 instructions.  The
// 192-thread instance hides that behind 24 warps per SM at 80 registers per thread.  The 128-thread instance has 16
// warps and 128 registers: room for a second, independent walk in the same thread.  tricks_two runs the walks of two
// rays in ONE loop whose body is two copies of the same straight-line probe (no branches on the common path: every
// decision is a select, a finished or absent walk is predicated off and its loads go to a safe address), so the
// compiler interleaves the two dependency chains and a trip costs little more than one probe's latency.  When every
// lane of the warp has finished one of the two, the loop continues with a single probe per trip.
//
// Each walk takes exactly the decisions of tricks_one (plicp2_search.cuh) on the same operands in the same order --
// best distance updated before the stop test, strict `<`, the cross-product stop test and the float test beyond kq,
// the jump chosen on the squared readings -- so the results are identical; tests/test_plicp_gpu.py compares the
// instance with the oracle and, through ZHSM_PLICP_T, with the other shapes.
#pragma once
#include "plicp2_search.cuh"

namespace PL_NAMESPACE {

struct Walk {
    double wx, wy, rho2;  // the ray (make_ray)
    double best, last_up, last_down;
    float phi;
    int start_cell;
    int j1;
    int up, down;   // next index to probe in each direction
    int stopped;    // bit 0: upwards ended, bit 1: downwards ended; 3: the walk is over (also: no walk)
    int steps;      // probes executed
};

__device__ __forceinline__ void walk_begin(Walk &w, const WalkCtx &c, const Ray &r, int last_best, bool have) {
    const int to = c.n - 1;
    w.wx = r.wx; w.wy = r.wy; w.rho2 = r.rho2; w.phi = r.phi; w.start_cell = r.start_cell;
    int we_start_at = (last_best == -1) ? r.start_cell : last_best + 1;
    we_start_at = we_start_at > to ? to : we_start_at;
    we_start_at = we_start_at < 0 ? 0 : we_start_at;
    w.up = we_start_at + 1; w.down = we_start_at;
    w.last_up = 0; w.last_down = 0;
    w.j1 = -1; w.best = INFINITY;
    w.stopped = have ? 0 : 3;
    w.steps = 0;
}

// One probe of walk w, straight-line.  A walk that is over changes nothing (its loads read index 0).
__device__ __forceinline__ void walk_probe(Walk &w, const WalkCtx &c) {
    const float HALF_PI_F = 1.57079632679489662f;
    const int to = c.n - 1;
    const bool active = w.stopped != 3;
    const bool now_up = (w.stopped == 2) | ((w.stopped == 0) & (w.last_up < w.last_down));
    const int idx = now_up ? w.up : w.down;
    const int bit = now_up ? 1 : 2;
    const bool inb = (unsigned)idx <= (unsigned)to;
    const int ia = (active & inb) ? idx : 0;
    double rr2;
    double2 q;
    int hops;
    lds_probe(c.rr2 + 8u * ia, c.refp + 16u * ia, (now_up ? c.jt : c.jt + 4u * (unsigned)c.n) + 4u * ia, rr2, q, hops);
    const int sdir = now_up ? 1 : -1;
    const bool valid = (rr2 >= 0) & (q.x == q.x);
    const bool take = active & inb & valid;
    const double d = dist2(w.wx, w.wy, q.x, q.y);
    w.last_up = (take & now_up) ? d : w.last_up;
    w.last_down = (take & !now_up) ? d : w.last_down;
    const bool better = take & (d < w.best);
    w.j1 = better ? idx : w.j1;
    w.best = better ? d : w.best;
    const int rel = (idx - w.start_cell) * sdir;
    const bool beyond = rel > 0;
    const double cross = w.wx * q.y - w.wy * q.x;
    bool stop = cross * cross > w.best * rr2;
    if (rel >= c.kq) {  // rare (walk_kq): only here can dtheta exceed pi/2 -- the exact float test, as the reference orders it
        const float dth = theta_f(c.fi_min, c.dfi, idx) - w.phi;
        if ((now_up ? dth : -dth) > HALF_PI_F) stop = w.rho2 > w.best;
    }
    stop = stop & take & beyond;
    const int jump = rr2 < w.rho2 ? (int)(short)hops : hops >> 16;
    const int hop = valid ? (beyond ? jump : sdir) : (hops >> 16);
    const int nidx = idx + hop;
    const bool ends = active & (!inb | stop);       // this direction has ended
    const bool moves = active & inb & !stop;        // ... or goes on to nidx
    w.stopped |= ends ? bit : 0;
    w.up = (moves & now_up) ? nidx : w.up;
    w.down = (moves & !now_up) ? nidx : w.down;
    w.steps += active ? 1 : 0;
}

// The walks of two rays (either may be absent: have = false), results as tricks_one returns them: j1 or -1, and j2 in
// the low half of *j2out with the walk's probes above.
static __device__ __noinline__ void tricks_two(const WalkCtx c, const Ray ra, const Ray rb, int last_best_a, int last_best_b,
                                               bool have_a, bool have_b, int *j1a, int *j2a, int *j1b, int *j2b) {
    constexpr unsigned FULL = 0xffffffffu;
    Walk a, b;
    walk_begin(a, c, ra, last_best_a, have_a);
    walk_begin(b, c, rb, last_best_b, have_b);
    // both walks per trip while some lane still has both (a lane with one walk left pays a predicated-off probe)
#pragma unroll 1
    while (__any_sync(FULL, (a.stopped != 3) & (b.stopped != 3))) {
        walk_probe(a, c);
        walk_probe(b, c);
    }
#pragma unroll 1
    while (a.stopped != 3) walk_probe(a, c);
#pragma unroll 1
    while (b.stopped != 3) walk_probe(b, c);
    int j2 = -1;
    int r1 = have_a ? walk_finish(c, a.wx, a.wy, a.j1, a.best, &j2) : -1;
    *j1a = r1;
    *j2a = (j2 & 0xffff) | (a.steps << 16);
    j2 = -1;
    r1 = have_b ? walk_finish(c, b.wx, b.wy, b.j1, b.best, &j2) : -1;
    *j1b = r1;
    *j2b = (j2 & 0xffff) | (b.steps << 16);
}

}  // namespace PL_NAMESPACE
