// laser_slam_b200/csrc/plicp2_flat.cuh -- a round of correspondence walks for the long-scan instances (LIST, T >= 256).
//
// tricks_one (plicp2_search.cuh) is called once per ray by every lane of a warp, and the warp leaves the call with its
// slowest lane.  On short scans with good first guesses the walks of a warp are alike (5.5 probes on average, 0.18 %
// of them longer than 32: profiles/r02y_walkstats.txt) and that costs little.  On 2048-beam multi-start jobs 11 % of
// the walks are longer than 32 probes and hold 46 % of all probes (profiles/r02y_c5_phases.txt): nearly every call has
// one, 15 of 32 lanes are alive in the loop on average, and the warps of a CTA end a round far apart.
//
// Here a round's walks are ONE pool of work items behind a counter in shared memory.  Every lane runs a small state
// machine -- idle / walking / ended -- whose walking state executes ONE probe of the same loop body as tricks_one per
// trip of a common loop.  When enough lanes of the warp have ended their walk or run idle (FLAT_R), or nobody walks,
// the warp takes a service trip: ended lanes do the walk's epilogue (walk_finish, the stores), lanes without a walk
// draw the next items from the counter and do the prologue (sens_point, make_ray).  Probes therefore run at the
// occupancy the threshold keeps, prologue and epilogue run FLAT_R lanes or more at a time, and a warp that is done
// early simply draws more of the pool: the round ends when the pool is empty, not when the unluckiest warp is done.
//
// Measured on one B200 (profiles/r02z_flat_variants_*.txt, r03a_flat_tail_variants_*.txt; 65 536 multi-start jobs at 2048 beams /
// 8 000 first-guess-0 pairs at 2048 beams): one tricks_one call per ray 85.5 k matches/s / 34.7 ms; the pool with
// FLAT_R 8: 95.4 k / 29.8, 16: 97.6 k / 28.3, 24: 98.7 k / 27.6, 32 (a service trip only when nobody walks): 104.2 k /
// 28.1 -- refilling inside a warp buys nothing over refilling the warp, the pool and the single loop do; list budgets
// 112 / 64 probes: 104.6 k / 28.1, with the tail budget 16: 107.8 k / 26.7; 224 / 96 / 16: 113.0 k / 27.6 (chosen);
// 448 / 192 / 16: 110.2 k / 30.6; without the list: 25.3 k / 168 ms.
//
// Decisions and operands are those of tricks_one, in the same order: the results are identical (the digest of
// tools/gpu_probe_c5.py over 65 536 jobs did not move; tests/test_plicp_gpu.py compares this path with the oracle on
// 2048-beam jobs with poor first guesses and, through the CTA-shape test, on a log).  Walks that outlive FLAT_LIST_* probes still go
// to the list of unfinished walks and are finished warp-wide after the round (finish_unfinished): with the other
// lanes kept busy they no longer cost lanes, only the round's tail.
#pragma once
#include "plicp2_search.cuh"

namespace PL_NAMESPACE {

#ifndef PLICP_FLAT_R
#define PLICP_FLAT_R 32            // lanes without a walk that trigger a service trip (32: only when no lane walks)
#endif
#ifndef PLICP_FLAT_LIST_BOTH
#define PLICP_FLAT_LIST_BOTH 224   // probes after which a walk with both directions alive may go on the list
#endif
#ifndef PLICP_FLAT_TAIL
#define PLICP_FLAT_TAIL 16         // ... once the round's pool is empty: what still walks then is the round's tail
#endif
#ifndef PLICP_FLAT_LIST_SOLO
#define PLICP_FLAT_LIST_SOLO 96    // ... with one direction left
#endif

struct BatchOut {
    int flags;  // bit 0: a walk changed its ray's j1; bit 1: a walk of this lane went on the list of unfinished walks
    int walks;  // walks this lane started
    int steps;  // probes this lane executed (a walk handed to the list carries its count in the entry)
};

// Work items [0, total): item e is sens ray queue[e] (from_queue: a re-search round) or sens ray e itself (the first
// pass; invalid sens rays are skipped).  ctl->work_next must be 0 and visible (a barrier) before the call; every
// thread of the job calls.
template <int T, bool LIST>
static __device__ __noinline__ BatchOut walk_batch(const Lay l, const WalkCtx c, const IterCtx it, const PendRef<LIST> pr,
                                                   int total, bool from_queue) {
    constexpr unsigned FULL = 0xffffffffu;
    const float HALF_PI_F = 1.57079632679489662f;
    const int lane = threadIdx.x & 31;
    const unsigned lane_lt = (1u << lane) - 1u;
    Ctl *ctl = PL_SH(Ctl, l.ctl);
    const float *sensr = PL_SH(const float, l.sensr);
    const short *queue = PL_SH(const short, l.d2j);
    const short *spec = PL_SH(const short, l.spec);
    short *cj1 = PL_SH(short, l.cj1), *cj2 = PL_SH(short, l.cj2);
    const int n = c.n, to = n - 1;
    const unsigned a_rr2 = c.rr2, a_refp = c.refp;
    const unsigned a_jt_up = c.jt, a_jt_dn = c.jt + 4u * (unsigned)n;
    const float fi_min = c.fi_min, dfi = c.dfi;
    const int kq = c.kq;

    // the lane's walk
    int state = 0;  // 0: no walk, 1: walking, 2: ended (epilogue pending)
    int ray_i = 0, j1 = -1, up = 0, down = 0, stopped = 0, probes = 0, start_cell = 0;
    int limit = PLICP_FLAT_LIST_BOTH;  // probes after which this walk may go on the list
    double wx = 0, wy = 0, rho2 = 0, best = INFINITY, last_up = 0, last_down = 0;
    float phi = 0.f;
    bool ask = LIST;   // may this walk still be put on the list?
    bool more = true;  // (warp-uniform) the pool may still hold items
    BatchOut out;
    out.flags = 0; out.walks = 0; out.steps = 0;

#pragma unroll 1
    for (;;) {
        // lanes a service trip would serve: while the pool lasts every lane that is not walking, afterwards the
        // lanes whose epilogue is pending (one ballot per trip until then)
        const unsigned m_walk = __ballot_sync(FULL, state == 1);
        int want = 32 - __popc(m_walk);
        if (!more) want = __popc(__ballot_sync(FULL, state == 2));
        if (m_walk == 0 && want == 0) break;
        if (want >= PLICP_FLAT_R || m_walk == 0) {
            // ---- service trip: epilogues, then new walks
            if (state == 2) {
                int j2 = -1;
                const int r1 = walk_finish(c, wx, wy, j1, best, &j2);
                if ((int)cj1[ray_i] != r1) out.flags |= 1;
                cj1[ray_i] = (short)r1;
                cj2[ray_i] = (short)(r1 >= 0 ? j2 : -1);
                out.steps += probes;
                state = 0;
            }
            if (more) {
                const unsigned m_get = __ballot_sync(FULL, state == 0);
                int base = 0;
                if (lane == 0) base = atomicAdd(&ctl->work_next, __popc(m_get));
                base = __shfl_sync(FULL, base, 0);
                more = base + __popc(m_get) < total;
                const int e = base + __popc(m_get & lane_lt);
                if (state == 0 && e < total) {
                    const int i = from_queue ? (int)queue[e] : e;
                    if (from_queue || sensr[i] >= 0) {
                        const Ray r = make_ray(sens_point(sensr, it.cs2, i), it.x0, it.x1, it.cs, it.sn, it.min_theta, it.C1);
                        wx = r.wx; wy = r.wy; rho2 = r.rho2; phi = r.phi; start_cell = r.start_cell;
                        const int last_best = (int)spec[i];
                        int we_start_at = (last_best == -1) ? start_cell : last_best + 1;
                        we_start_at = we_start_at > to ? to : we_start_at;
                        we_start_at = we_start_at < 0 ? 0 : we_start_at;
                        up = we_start_at + 1; down = we_start_at;
                        last_up = 0; last_down = 0;
                        j1 = -1; best = INFINITY;
                        stopped = 0; probes = 0; ask = LIST; limit = PLICP_FLAT_LIST_BOTH;
                        ray_i = i;
                        state = 1;
                        out.walks++;
                    }
                }
            }
            // the pool is empty: what still walks is the round's tail, and the list finishes long walks faster
            if (!more) limit = limit < PLICP_FLAT_TAIL ? limit : PLICP_FLAT_TAIL;
            continue;
        }
        if (state != 1) continue;
        // ---- one probe: the body of tricks_one's general loop (with one direction ended it walks the other)
        probes++;
        const bool now_up = (stopped == 2) | ((stopped == 0) & (last_up < last_down));
        const int idx = now_up ? up : down;
        const int bit = now_up ? 1 : 2;
        if ((unsigned)idx > (unsigned)to) {  // ran off either end of the scan
            stopped |= bit;
        } else {
            double rr2;
            double2 q;
            int hops;
            lds_probe(a_rr2 + 8u * idx, a_refp + 16u * idx, (now_up ? a_jt_up : a_jt_dn) + 4u * idx, rr2, q, hops);
            const int sdir = now_up ? 1 : -1;
            int nidx = idx + sdir;
            bool stop = false;
            if ((rr2 >= 0) & (q.x == q.x)) {
                const double d = dist2(wx, wy, q.x, q.y);
                last_up = now_up ? d : last_up;
                last_down = now_up ? last_down : d;
                const bool better = d < best;
                j1 = better ? idx : j1;
                best = better ? d : best;
                const int rel = (idx - start_cell) * sdir;
                if (rel > 0) {
                    const double cross = wx * q.y - wy * q.x;
                    stop = cross * cross > best * rr2;
                    if (rel >= kq) {
                        const float dth = theta_f(fi_min, dfi, idx) - phi;
                        if ((now_up ? dth : -dth) > HALF_PI_F) stop = rho2 > best;
                    }
                    nidx = idx + (rr2 < rho2 ? (int)(short)hops : hops >> 16);
                }
            } else {
                nidx = idx + (hops >> 16);  // an invalid ref ray: the skip over its whole run
            }
            if (stop) stopped |= bit;
            else {
                up = now_up ? nidx : up;
                down = now_up ? down : nidx;
            }
        }
        if (stopped) limit = limit < PLICP_FLAT_LIST_SOLO ? limit : PLICP_FLAT_LIST_SOLO;
        if (stopped == 3) state = 2;
        else if (probes >= limit && ask) {
            WalkState st;
            st.best = best; st.i = ray_i; st.j1 = j1; st.pad_ = probes;
            if (stopped == 0) { st.last_up = last_up; st.last_down = last_down; st.up = up; st.down = down; st.dir = 2; }
            else { st.last_up = 0; st.last_down = 0; st.up = stopped == 2 ? up : down; st.down = 0; st.dir = stopped == 2 ? 1 : -1; }
            if (pend_push(pr, st)) { state = 0; out.flags |= 2; }  // its probes so far travel with the entry
            else ask = false;                                      // a full list: the lane walks on
        }
    }
    return out;
}

}  // namespace PL_NAMESPACE
