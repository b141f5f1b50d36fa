// laser_slam_b200/csrc/plicp2_kernel.cuh -- the PL-ICP kernel: load, icp_loop x (1+6 restarts), result record.
// Replaces, per job, CCanonicalMatcher::FMatch(LDP,LDP) -> sm_icp (src/zhcsm/ZHCanonical_Matcher.cpp:251-285) with
// the parameters of sm_init (:83-157).  SURVEY 8a-2..8a-8.  FP64 throughout (CSM is double end to end), -fmad=false.
#pragma once
#include "plicp2_common.cuh"
#include "plicp2_cov.cuh"
#include "plicp2_search.cuh"
#include "plicp2_flat.cuh"
#include "plicp2_select.cuh"
#include "plicp2_solve.cuh"

namespace PL_NAMESPACE {

// Which instances carry the list of unfinished walks (plicp2_search.cuh: a walk that outlives its budgets is finished
// warp-wide after the round's barrier).  PLICP_LIST_MIN_T: tuning builds.
#ifndef PLICP_LIST_MIN_T
#define PLICP_LIST_MIN_T 256
#endif
template <int T>
constexpr bool kList = T >= PLICP_LIST_MIN_T;
// ... and run a round's walks as one pool of work items (plicp2_flat.cuh) instead of one tricks_one call per ray
#ifndef PLICP_FLAT
#define PLICP_FLAT 1
#endif
// The final-only instances (plicp2_final.cu): the last pass of a job whose restarts ran as a second batch
// (zhsm_api.cu: plicp_fanout).  A compile-time mode, so that the ordinary instances carry nothing of it.
#ifndef PLICP_FINAL
#define PLICP_FINAL 0
#endif
#ifndef PLICP_MAX_LOOPS
#define PLICP_MAX_LOOPS 7   // icp_loops a job may run: 1 + sm_icp's six restarts (tuning builds: 1 = no in-job restarts)
#endif
#ifndef PLICP_FLAT_ALL
#define PLICP_FLAT_ALL 0   // tuning builds: the short-scan instances as well (no list there: every walk ends in its lane)
#endif
template <int T>
constexpr bool kFlat = PLICP_FLAT && (kList<T> || PLICP_FLAT_ALL);
// (The other instances keep their rays fixed per thread: drawing groups of 32 rays from the same counter measured
// 16.38 against 16.34 ms per 20k pairs at 1080 beams and 11.80 against 11.60 ms per 40k pairs at 360 -- nothing.)

// Per-phase cycle counters of thread 0 (tuning builds only: -DPLICP_PHASE_TIMING; the counters then REPLACE
// x/cov/error/goodness/quality in the result record, see tools/gpu_probe_phases.py).
struct PhaseClock {
#ifdef PLICP_PHASE_TIMING
    long long last, acc[20];
    __device__ __forceinline__ void start() { for (int k = 0; k < 20; k++) acc[k] = 0; last = clock64(); }
    __device__ __forceinline__ void mark(int k) { const long long c = clock64(); acc[k] += c - last; last = c; }
    // one call of the walk routine by this thread's warp: cycles inside it (slot 1); slot 17 = calls (low 8 bits) +
    // the longest lane's step count summed (above); acc[18]'s record field is a float: steps summed over ALL lanes there
    __device__ __forceinline__ void walk(long long cycles, int max_steps) { acc[1] += cycles; acc[17] += ((long long)max_steps << 8) + 1; }
#else
    __device__ __forceinline__ void start() {}
    __device__ __forceinline__ void mark(int) {}
    __device__ __forceinline__ void walk(long long, int) {}
#endif
};

// diagnostics of the correspondence search (zhsm_result.reserved): walks by this thread, of which first-pass walks
// (= what the sequential algorithm executes); verify rounds that had work
struct WalkDiag {
    int walks, first, rounds;
    int steps;  // walk steps this thread executed (probes of its walks: the budgets of tricks_one count them)
};


#ifdef PLICP_WALK_STATS
// tuning build: per-thread walk statistics (trips, trips of walks > 32, count, trips of walks > 256, count, max)
#define g_stats st_
#define STATS_PARAM , int *st_
#define STATS_ARG , st_
#else
#define STATS_PARAM
#define STATS_ARG
#endif

// search sens ray i with its speculated last_best; returns false when the walk went on the list of unfinished walks
// (process_pending finishes it and writes cj1 / cj2 then)
template <bool LIST>
__device__ __forceinline__ bool search_ray(const Lay &l, const WalkCtx &w, const IterCtx &it,
                                           const PendRef<LIST> &pr, int i, int &steps, PhaseClock &pc STATS_PARAM) {
    const Ray ray = make_ray(sens_point(PL_SH(const float, l.sensr), it.cs2, i), it.x0, it.x1, it.cs, it.sn, it.min_theta, it.C1);
    int j2 = -1;
#ifdef PLICP_WALK_STATS
    int trips = 0;
    const int j1 = tricks_one<LIST>(w, ray, (int)PL_SH(const short, l.spec)[i], &j2, pr, i, &trips);
    const int fast = trips >> 16;
    trips &= 0xffff;
    g_stats[0] += trips;
    g_stats[1] += trips > 32 ? trips : 0;
    g_stats[2] += trips > 32;
    g_stats[3] += fast;
    g_stats[4] += trips > 32 ? fast : 0;
    g_stats[5] = trips > g_stats[5] ? trips : g_stats[5];
#else
#ifdef PLICP_PHASE_TIMING
    const long long c0_ = clock64();
#endif
    const int j1 = tricks_one<LIST>(w, ray, (int)PL_SH(const short, l.spec)[i], &j2, pr, i);
#ifdef PLICP_PHASE_TIMING
    pc.walk(clock64() - c0_, __reduce_max_sync(__activemask(), (unsigned)j2 >> 16));
#endif
#endif
    if constexpr (LIST) {
        if (j1 == WALK_PENDING) return false;  // its steps so far travel with the list entry
    }
    steps += (int)((unsigned)j2 >> 16);
    j2 = (int)(short)(j2 & 0xffff);
    PL_SH(short, l.cj1)[i] = (short)j1;
    PL_SH(short, l.cj2)[i] = (short)(j1 >= 0 ? j2 : -1);
    return true;
}

// The walks a round left unfinished (tricks_one put them on the list), finished warp-wide: entry e by warp e mod NW.
// Called by every thread after the barrier that ends the round; returns whether a finished walk changed its ray's j1.
// Ends with a barrier when there was anything on the list.
template <int T>
static __device__ __noinline__ int finish_unfinished(const Lay l, const WalkCtx w, const IterCtx it,
                                                     const PendRef<true> pr, int np);

template <int T>
__device__ __forceinline__ int process_pending(const Lay &l, const WalkCtx &w, const IterCtx &it,
                                               const PendRef<kList<T>> &pr) {
    if constexpr (!kList<T>) {
        return 0;  // the list exists in the long-scan instances only
    } else {
        const int np = PL_SH(Ctl, l.ctl)->npend;
        if (np == 0) return 0;
        return finish_unfinished<T>(l, w, it, pr, np);
    }  // rare: kept out of line, away from the callers' registers
}

template <int T>
static __device__ __noinline__ int finish_unfinished(const Lay l, const WalkCtx w, const IterCtx it,
                                                     const PendRef<true> pr, int np) {
    Ctl *ctl = PL_SH(Ctl, l.ctl);
    np = np < pr.pend_cap ? np : pr.pend_cap;
    const int lane = threadIdx.x & 31;
    int changed = 0;
    // the walks differ in length by two orders of magnitude: a warp takes the next entry when it is done with its last
#pragma unroll 1
    for (;;) {
        int e = 0;
        if (lane == 0) e = atomicAdd(&ctl->pend_next, 1);
        e = __shfl_sync(0xffffffffu, e, 0);
        if (e >= np) break;
        const unsigned a = pr.pend + 48u * (unsigned)e;
        const double2 bl = lds_f64x2(a);
        const double last_down = lds_f64(a + 16u);
        int i, j1, up, down, dir, pad;
        asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(i), "=r"(j1) : "r"(a + 24u));
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(up), "=r"(down), "=r"(dir), "=r"(pad) : "r"(a + 32u));
        const Ray ray = make_ray(sens_point(PL_SH(const float, l.sensr), it.cs2, i), it.x0, it.x1, it.cs, it.sn, it.min_theta, it.C1);
        RayU u;
        u.wx = ray.wx; u.wy = ray.wy; u.rho2 = ray.rho2; u.phi = ray.phi; u.start_cell = ray.start_cell;
        double cb = bl.x;
        int cj = j1;
        int nwin;
        if (dir == 2) nwin = walk_warp_both(w, u, up, down, bl.y, last_down, &cb, &cj);
        else nwin = walk_warp_one(w, u, up, dir, &cb, &cj);
        int j2 = -1;
        const int r1 = walk_finish(w, ray.wx, ray.wy, cj, cb, &j2);
        if (lane == 0) {
            atomicAdd(&ctl->steps_pend, pad + 32 * nwin);  // the lane's steps before the hand-over + 32 evaluations per window
            short *cj1 = PL_SH(short, l.cj1);
            changed |= (int)cj1[i] != r1;
            cj1[i] = (short)r1;
            PL_SH(short, l.cj2)[i] = (short)(r1 >= 0 ? j2 : -1);
        }
    }
    pl_sync<T>();
    if (pl_tid<T>() == 0) { ctl->npend = 0; ctl->pend_next = 0; }  // the next round of walks is at least one more barrier away
    return changed;
}

// The jump tables' way back into shared memory as ONE bulk copy (cp.async.bulk through the TMA unit, completion on an
// mbarrier) issued by thread 0 as soon as the radix select is done with the area they share -- the copy then runs behind
// the sums, the hash and the solve of the same iteration, and the next search only waits for the barrier's phase.
// PLICP_TMA_JT builds; off, the threads copy the tables themselves at the start of every search.  Needs 16-byte
// aligned tables: an even beam count.  All members are uniform over the CTA.
struct JtPrefetch {
    bool on;            // this job uses the bulk copy
    bool pending;       // a copy has been issued and not yet waited for
    unsigned phase;     // parity of the barrier phase the next wait observes
    unsigned mbar, dst, bytes;
    const void *src;
    __device__ __forceinline__ void issue() {  // call by every thread, after a barrier that ends all uses of the area
        if (!on) return;
        if (threadIdx.x == 0) bulk_load_1d(dst, src, bytes, mbar);
        pending = true;
    }
    __device__ __forceinline__ void wait() {
        if (!on || !pending) return;
        mbar_wait(mbar, phase);
        phase ^= 1u;
        pending = false;
    }
};

// find_correspondences_tricks for the whole scan: speculate / verify / re-search until the sequential result.
// First pass: thread t walks its rays k*T+t with the last_best the previous iteration saw (spec[]).
// Verify: last_best before ray i is j1 of the nearest earlier ray holding a correspondence -- found with warp ballots
// inside a group of 32 consecutive rays and a table of the groups' last results across groups.  A ray whose actual
// last_best differs from spec[] is appended to a CTA-wide queue (warp-aggregated atomicAdd), and the queue is walked
// densely, entry e by thread e mod T (re-searched rays are sparse, about 2 of 3; queueing them fills the lanes).
// The fixpoint is reached when no re-search changed a j1 (the usual case after one round).
// The queue and the group table live in the d2j area; d2j itself is initialised after the search.
// sens: bit k = ray k*T+t exists and is valid; returns the mask of those rays holding a correspondence.
template <int T>
__device__ __forceinline__ unsigned find_correspondences(const Lay &l, const WalkCtx &w, const IterCtx &it, int C,
                                                         unsigned sens, bool chunked, const uint2 *jt_src,
                                                         const PendRef<kList<T>> &pr, WalkDiag &dg,
                                                         PhaseClock &pc, JtPrefetch &tma, bool jt_in_place STATS_PARAM) {
    constexpr int NW = T / 32;
    const int t = pl_tid<T>(), lane = t & 31, warp = t >> 5;
    const int n = w.n;
    const unsigned lane_lt = (1u << lane) - 1u;
    Ctl *ctl = PL_SH(Ctl, l.ctl);
    if (t == 0) ctl->qn[0] = 0;
    if (jt_src) {
        // the radix select's histograms share the jump tables' shared memory (Lay::jt_shared): bring the tables back
        // (8 bytes per ray from this job's scratch, L2-resident)
        if (tma.on) {
            // ... which a bulk copy issued after the previous select has already done (JtPrefetch): wait for its phase
            tma.wait();
        } else if (!jt_in_place) {  // (else: copied behind the previous iteration's solve)
            uint2 *dst = PL_SH(uint2, l.jt);
            for (int q = t; q < n; q += T) dst[q] = __ldcs(jt_src + q);  // streaming: leave the L1 to the cs table
            pl_sync<T>();
        }
    }
    if constexpr (kFlat<T>) {
        // the pool instances have no chunk-order iterations (they lose there: plicp_chunk_iters), so the ray-by-ray
        // search is not compiled into them at all
        if (t == 0) ctl->work_next = 0;
        pl_sync<T>();
        const BatchOut bo = walk_batch<T, kList<T>>(l, w, it, pr, n, false);
        dg.walks += bo.walks;
        dg.first += bo.walks;
        dg.steps += bo.steps;
    } else if (!chunked) {
#pragma unroll 1
        for (int k = 0; k < C; k++)
            if (sens >> k & 1u) {
                dg.walks++;
                dg.first++;
                search_ray<kList<T>>(l, w, it, pr, k * T + t, dg.steps, pc STATS_ARG);
            }
    } else {
        // thread t walks the rays [t*C, (t+1)*C) in order and hands last_best from one to the next: only the first
        // ray of a chunk is speculated
        const float *sensr = PL_SH(const float, l.sensr);
        short *specw = PL_SH(short, l.spec);
        const short *cj1r = PL_SH(const short, l.cj1);
        int lb = 0;
        bool have = false;
#pragma unroll 1
        for (int q = 0; q < C; q++) {
            const int i = t * C + q;
            if (i >= n) break;
            if (sensr[i] < 0) continue;  // invalid sens ray: no walk, last_best unchanged
            if (have) specw[i] = (short)lb;
            dg.walks++;
            dg.first++;
            if (search_ray<kList<T>>(l, w, it, pr, i, dg.steps, pc STATS_ARG)) {
                const int c1 = cj1r[i];
                lb = c1 >= 0 ? c1 : (int)specw[i];
                have = true;
            } else {
                have = false;  // an unfinished walk: its successor is speculated like the first ray of a chunk
            }
        }
    }
    pl_sync<T>();
    process_pending<T>(l, w, it, pr);
    pc.mark(2);
    short *cj1 = PL_SH(short, l.cj1);
    short *spec = PL_SH(short, l.spec);
    short *queue = PL_SH(short, l.d2j);
    short *glast = queue + n;  // [C][NW]: j1 of the last ray with a correspondence in each group of 32 rays, or -1
    for (int round = 0;; round++) {
        int *qn = &ctl->qn[round & 1];
        if (t == 0) { ctl->qn[(round + 1) & 1] = 0; ctl->work_next = 0; }
#pragma unroll 1
        for (int k = 0; k < C; k++) {
            const int i = k * T + t;
            const int c1 = i < n ? (int)cj1[i] : -1;
            const unsigned vm = __ballot_sync(0xffffffffu, c1 >= 0);
            const int last = __shfl_sync(0xffffffffu, c1, vm ? 31 - __clz(vm) : 0);
            if (lane == 0) glast[k * NW + warp] = (short)(vm ? last : -1);
        }
        pl_sync<T>();
#pragma unroll 1
        for (int k = 0; k < C; k++) {
            const int i = k * T + t;
            const int c1 = i < n ? (int)cj1[i] : -1;
            const unsigned vm = __ballot_sync(0xffffffffu, c1 >= 0);
            const unsigned below = vm & lane_lt;
            int actual = __shfl_sync(0xffffffffu, c1, below ? 31 - __clz(below) : 0);
            bool need = false;
            if (sens >> k & 1u) {
                if (!below) {
                    actual = -1;
                    for (int g = k * NW + warp - 1; g >= 0; g--) {
                        const int v = glast[g];
                        if (v >= 0) { actual = v; break; }
                    }
                }
                if (actual != (int)spec[i]) { need = true; spec[i] = (short)actual; }
            }
            const unsigned m = __ballot_sync(0xffffffffu, need);
            if (m) {
                int base = 0;
                if (lane == 0) base = atomicAdd(qn, __popc(m));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (need) queue[base + __popc(m & lane_lt)] = (short)i;
            }
        }
        pl_sync<T>();
        const int total = *qn;
        if (total == 0) break;
        dg.rounds++;
        int changed = 0, unfinished = 0;
        if constexpr (kFlat<T>) {
            const BatchOut bo = walk_batch<T, kList<T>>(l, w, it, pr, total, true);
            changed = bo.flags & 1;
            unfinished = bo.flags >> 1;
            dg.walks += bo.walks;
            dg.steps += bo.steps;
        } else {
#pragma unroll 1
            for (int e = t; e < total; e += T) {
                const int i = (int)queue[e];
                const int before = cj1[i];
                dg.walks++;
                if (search_ray<kList<T>>(l, w, it, pr, i, dg.steps, pc STATS_ARG)) changed |= (int)cj1[i] != before;
                else unfinished = 1;  // on the list: process_pending says whether its j1 changed
            }
        }
        // the usual end: nothing changed and nothing is unfinished -- last_best of every ray is what it was walked with
        if (!pl_sync_or<T>(changed | unfinished)) break;
        if constexpr (kList<T>) {
            changed |= process_pending<T>(l, w, it, pr);
            if (!pl_sync_or<T>(changed)) break;
        }
    }
    pc.mark(3);
    unsigned valid = 0;
#pragma unroll 1
    for (int k = 0; k < C; k++)
        if ((sens >> k & 1u) && cj1[k * T + t] >= 0) valid |= 1u << k;
    return valid;
}

#ifndef PLICP_MINB_192
#define PLICP_MINB_192 4   // resident CTAs per SM the 192-thread instance is compiled for (80 registers); tuning builds: 3
#endif
template <int T>
__global__ void __launch_bounds__(T * PLICP_NS, (PLICP_NS > 1 ? 1 : T == 64 ? 8 : T == 160 ? 5 : T <= 192 ? PLICP_MINB_192 : T <= 288 ? 2 : 1))
plicp2_kernel(PlicpArgs a, const Lay lay0) {
    constexpr int NW = T / 32;
    const int n = a.L.n;
    const int C = (n + T - 1) / T;
    const int t = pl_tid<T>(), lane = t & 31, warp = t >> 5;
    const int job = PLICP_NS == 1 ? (int)blockIdx.x : (int)blockIdx.x * PLICP_NS + pl_slot<T>();
    // The shared-memory layout comes with the launch (plicp2_launch.cuh): its offsets are then operands from the
    // constant bank instead of arithmetic on n that the compiler repeats wherever registers are short (4 % of the
    // r02t capture's instructions).  lay0 is slot 0's layout; another slot's areas lie slot * slot_bytes further on.
    Lay l = lay0;
    if (PLICP_NS > 1) {
        const unsigned d = slot_bytes(n, T) * (unsigned)pl_slot<T>();
        l.d2j += d; l.red += d; l.ctl += d; l.hist += d; l.candi += d; l.cj1 += d; l.cj2 += d; l.spec += d;
    }
    Ctl *ctl = PL_SH(Ctl, l.ctl);
    const LaserDev L = a.L;
    const zhsm_plicp_params &P = a.P;
    double2 *refp = PL_SH(double2, l.refp);
    float *sensr = PL_SH(float, l.sensr);
    double *rr2a = PL_SH(double, l.rr2);
    unsigned long long *d2j = PL_SH(unsigned long long, l.d2j);
    short *cj1 = PL_SH(short, l.cj1), *cj2 = PL_SH(short, l.cj2);
    double *red = PL_SH(double, l.red);

    PhaseClock pc;
    pc.start();
    volatile int *msctl = PL_SH(volatile int, l.msctl);
    if (PLICP_NS > 1 && threadIdx.x == 0) { msctl[0] = 0; msctl[1] = 0; }  // ordered before any use by the CTA-wide barrier
    // ---------------- load + fromPMScan2LDP + ld_valid_fields + ld_invalid_if_outside + ld_compute_cartesian
    const int ref_i = a.ref_idx[job], sens_i = a.sens_idx[job];
    bool bad_job = (unsigned)ref_i >= (unsigned)a.S || (unsigned)sens_i >= (unsigned)a.S;
    if (PLICP_NS > 1) {
        // The CTA's jobs must share both scans (the caller's promise: ZHSM_MULTISTART); a job that does not is failed,
        // loudly, instead of being matched against the wrong scans.  Its threads still take part in the one CTA-wide
        // barrier below.  The shared arrays are filled from slot 0's scans, by every slot (the same values).
        const int ref0 = a.ref_idx[blockIdx.x * PLICP_NS], sens0 = a.sens_idx[blockIdx.x * PLICP_NS];
        bad_job = bad_job || ref_i != ref0 || sens_i != sens0;
        if ((unsigned)ref0 >= (unsigned)a.S || (unsigned)sens0 >= (unsigned)a.S) bad_job = true;  // (uniform over the CTA)
        if (bad_job) __syncthreads();
    }
    if (bad_job) {
        // device-resident job lists are not checked by the host: a bad index fails its own job, never the batch
        const int ob = PLICP_FINAL ? a.out_idx[job] : job;
        if (t == 0) {
            zhsm_result r = {};
            r.status = ZHSM_JOB_INVALID_SCAN;
            a.results[ob] = r;
        }
        if (a.corr_j1) for (int i = t; i < n; i += T) { a.corr_j1[(size_t)ob * n + i] = -1; a.corr_j2[(size_t)ob * n + i] = -1; }
        if (PLICP_NS > 1 && t == 0) atomicAdd((int *)&msctl[0], 1);  // this slot never reads the shared arrays
        return;
    }
    const float *ref_row = a.scans + (size_t)ref_i * n;
    const float *sens_row = a.scans + (size_t)sens_i * n;
    int bad_field = 0, nvf_ref = 0, nvf_sens = 0;
    unsigned sens = 0;
    // jump tables of this job's ref scan, built by jump_tables_kernel (plicp2_jump.cu) in the slot of the scan's owning
    // job: 8 bytes per ray
    const uint2 *jt_job = reinterpret_cast<const uint2 *>(a.jt + (size_t)a.owner[ref_i] * 4 * n);
    if (!l.jt_shared) {
        uint2 *dst = PL_SH(uint2, l.jt);
        for (int q = t; q < n; q += T) dst[q] = __ldcs(jt_job + q);
    }
    const uint2 *jt_src = l.jt_shared ? jt_job : nullptr;  // find_correspondences copies them in before every search
    JtPrefetch tma;
    tma.on = false; tma.pending = false; tma.phase = 0;
#ifdef PLICP_TMA_JT
    if (jt_src && (n & 1) == 0) {
        const unsigned sb = (unsigned)__cvta_generic_to_shared(smem_raw);
        tma.on = true;
        tma.mbar = sb + l.ctl + (unsigned)offsetof(Ctl, mbar);
        tma.dst = sb + l.jt;
        tma.bytes = 8u * (unsigned)n;
        tma.src = jt_job;
        if (t == 0) mbar_init(tma.mbar, 1);
        pl_sync<T>();   // the barrier object is initialised before anyone arms or polls it
        tma.issue();       // the first search's tables travel while the scans are loaded and converted
    }
#endif
#pragma unroll 1
    for (int k = 0; k < C; k++) {
        const int i = k * T + t;
        if (i >= n) break;
        const double2 cs = a.cs[i];
        double rr = reading_of(__ldcs(ref_row + i), L, P, &bad_field);
        nvf_ref += (rr != -1.0);
        rr2a[i] = rr >= 0 ? rr * rr : -1.0;
        refp[i] = rr >= 0 ? make_double2(cs.x * rr, cs.y * rr) : make_double2(nan(""), nan(""));
        const float fs = __ldcs(sens_row + i);
        double rs = reading_of(fs, L, P, &bad_field);
        nvf_sens += (rs != -1.0);
        sensr[i] = rs >= 0 ? fs : -1.f;
        if (rs >= 0) sens |= 1u << k;
        cj1[i] = -1;
        cj2[i] = -1;
        PL_SH(short, l.spec)[i] = -1;
    }
    if (PLICP_NS > 1) __syncthreads();  // the one CTA-wide barrier: the shared arrays are complete (every slot wrote them)
    if (t == 0) {
        ctl->topbyte = 0x3f;
        ctl->npend = 0;
        ctl->pend_next = 0;
        ctl->steps_pend = 0;
        for (int k = 0; k < 3; k++) ctl->x[0][k] = a.first_guess ? a.first_guess[3 * (size_t)job + k] : 0.0;
        ctl->cs[0] = cos(ctl->x[0][2]);
        ctl->sn[0] = sin(ctl->x[0][2]);
    }
    const int tot = block_sum<T>((nvf_ref << 16) | nvf_sens, ctl->ired, 0);
    const int tot_bad = pl_sync_or<T>(bad_field);
    const int tot_ref = tot >> 16, tot_sens = tot & 0xffff;
    const int out = PLICP_FINAL ? a.out_idx[job] : job;
    zhsm_result *res = a.results + out;
    int16_t *oj1 = a.corr_j1 ? a.corr_j1 + (size_t)out * n : nullptr;
    int16_t *oj2 = a.corr_j2 ? a.corr_j2 + (size_t)out * n : nullptr;
    const bool guess_nan = isnan(ctl->x[0][0]) || isnan(ctl->x[0][1]) || isnan(ctl->x[0][2]);
    const bool fields_bad = tot_bad || tot_ref < n * 0.10 || tot_sens < n * 0.10;
    if (fields_bad || guess_nan) {
        // ld_valid_fields failed (sm_icp returns valid=0 before iterating) or icp_loop refused a NaN first guess
        if (t == 0) {
            zhsm_result r = {};
            r.status = ZHSM_JOB_INVALID_SCAN;
            r.n_loops = fields_bad ? 0 : 1;
            *res = r;
        }
        if (oj1) for (int i = t; i < n; i += T) { oj1[i] = -1; oj2[i] = -1; }
        tma.wait();  // nothing may be in flight into this CTA's shared memory when it exits
        if (PLICP_NS > 1 && t == 0) atomicAdd((int *)&msctl[0], 1);
        return;
    }
    pc.mark(0);  // the block sums above were barriers: the tables are in place

    IterCtx it;
    bool jt_in_place = false;  // the jump tables are in shared memory (they share it with the select's histograms)
    it.cs2 = a.cs;
    it.min_theta = (double)theta_f(L.fi_min, L.dfi, 0);
    it.C1 = (double)n / ((double)theta_f(L.fi_min, L.dfi, n - 1) - it.min_theta);
    WalkCtx wc;
    {
        const unsigned sb = (unsigned)__cvta_generic_to_shared(smem_raw);
        wc.refp = sb + l.refp; wc.rr2 = sb + l.rr2; wc.jt = sb + l.jt;
    }
    wc.n = n; wc.fi_min = L.fi_min; wc.dfi = L.dfi;
    wc.kq = walk_kq(L.dfi);
    PendRef<kList<T>> pr;
    if constexpr (kList<T>) {
        // the list of unfinished walks lies in the d2j area above the search's queue and group table
        const unsigned sb = (unsigned)__cvta_generic_to_shared(smem_raw);
        const unsigned lo = (sb + l.d2j + 2u * (unsigned)(n + C * NW) + 15u) & ~15u, hi = sb + l.d2j + 8u * (unsigned)n;
        pr.pend = lo;
        pr.npend = sb + l.ctl + (unsigned)offsetof(Ctl, npend);
        const int room = hi > lo ? (int)((hi - lo) / 48u) : 0;
        pr.pend_cap = room < 1024 ? room : 1024;
    }
    wc.maxd2 = P.max_correspondence_dist * P.max_correspondence_dist;
    const int max_it = P.max_iterations < ZHSM_MAX_ITER ? P.max_iterations : ZHSM_MAX_ITER;
    const double fail_perc = 0.05;
    const unsigned long long BIG = (unsigned long long)__double_as_longlong(1000000.0);

    // sm_icp bookkeeping (icp.c), meaningful on thread 0
    double best_x[3] = {0, 0, 0}, x_first[3] = {0, 0, 0}, best_error = 0;
    int iterations = 0, nvalid0 = 0, n_loops = 0, restarted = 0, valid_out = 0, status = ZHSM_JOB_OK;
    unsigned valid = 0;
    WalkDiag dg = {0, 0, 0, 0};
#ifdef PLICP_WALK_STATS
    int st_arr[6] = {0, 0, 0, 0, 0, 0};
    int *st_ = st_arr;
#endif

    if constexpr (PLICP_FINAL != 0) {
        // The last pass of a job whose restarts ran as a second batch (zhsm_api.cu: plicp_fanout): the bookkeeping of
        // sm_icp comes with the launch, what is left is "recompute correspondences in case of restarts" and the record.
        if (t == 0) {
            const PlicpFin f = a.fin[job];
            for (int k = 0; k < 3; k++) best_x[k] = f.x[k];
            best_error = f.error;
            nvalid0 = f.nvalid;
            iterations = f.iterations;
            n_loops = f.n_loops;
            valid_out = 1;
            restarted = 1;
            dg.walks += f.diag[0]; dg.first += f.diag[1]; dg.rounds += f.diag[2]; dg.steps += f.diag[3];
            ctl->fin = 1 | 2;
        }
        pl_sync<T>();
    }
    // (the compiler unrolls this loop of 7: 600 KB of code, and 3.4 % FASTER on the 1080-beam batch than with
    // `#pragma unroll 1` -- the first copy, the only one most jobs run, is specialised for loop == 0; measured r02p)
    for (int loop = 0; loop < (PLICP_FINAL ? 0 : PLICP_MAX_LOOPS); loop++) {
        int iteration = 0, all_ok = 1, par = 0, newest = 0;
        for (iteration = 0; iteration < max_it; iteration++) {
            it.x0 = ctl->x[par][0];
            it.x1 = ctl->x[par][1];
            it.cs = ctl->cs[par];
            it.sn = ctl->sn[par];
            if (t == 0) ctl->hash_acc[iteration & 1] = 0;
            valid = find_correspondences<T>(l, wc, it, C, sens, iteration < a.chunk_iters, jt_src, pr, dg, pc, tma, jt_in_place STATS_ARG);
            if (P.outliers_remove_doubles) {
                for (int j = t; j < n; j += T) d2j[j] = BIG;  // the search used this area as its queue
                pl_sync<T>();
            }
            // kill_outliers_double (icp_outliers.c): min dist2_j1 per ref index.  dist2_j1 is recomputed
            // from the same operands the walk used (bit-identical), so it needs no storage.
            if (P.outliers_remove_doubles) {
#pragma unroll 1
                for (int k = 0; k < C; k++) {
                    if (!(valid >> k & 1u)) continue;
                    const int i = k * T + t;
                    const double2 p = sens_point(sensr, a.cs, i);
                    const double wx = it.cs * p.x - it.sn * p.y + it.x0;
                    const double wy = it.sn * p.x + it.cs * p.y + it.x1;
                    const int j1 = cj1[i];
                    const double2 q = refp[j1];
                    atomicMin(&d2j[j1], (unsigned long long)__double_as_longlong(dist2(wx, wy, q.x, q.y)));
                }
            }
            const int num_corr = block_sum<T>(__popc(valid), ctl->ired, 1);
            if (num_corr < fail_perc * n) { all_ok = 0; status = ZHSM_JOB_TOO_FEW_POINTS; break; }
            if (P.outliers_remove_doubles) {
#pragma unroll 1
                for (int k = 0; k < C; k++) {
                    if (!(valid >> k & 1u)) continue;
                    const int i = k * T + t;
                    const double2 p = sens_point(sensr, a.cs, i);
                    const double wx = it.cs * p.x - it.sn * p.y + it.x0;
                    const double wy = it.sn * p.x + it.cs * p.y + it.x1;
                    const int j1 = cj1[i];
                    const double2 q = refp[j1];
                    const double dj = __longlong_as_double((long long)d2j[j1]);
                    if (dist2(wx, wy, q.x, q.y) > (3.0 * 3.0) * dj) valid &= ~(1u << k);
                }
            }
            // all three histogram sets of the radix select; set 1 is filled by the key loop below (plicp2_select.cuh)
            for (int b = t; b < 768; b += T) PL_SH(int, l.hist)[b] = 0;
            if (t == 0) ctl->sel_below = 0;
            const int top_guess = ctl->topbyte;
            pc.mark(4);
            const int kcnt = block_sum<T>(__popc(valid), ctl->ired, 2);  // all reads of d2j are done
            // kill_outliers_trim: distance to the segment (j1,j2); keys go to shared memory, and their
            // second byte into the select's first histogram when the top byte is the guessed one
            {
                int *hist1 = PL_SH(int, l.hist) + 256;
                short *cand1 = PL_SH(short, l.candi) + 512;
                int nb = 0;
#pragma unroll 1
                for (int k = 0; k < C; k++) {
                    if (!(valid >> k & 1u)) continue;
                    const int i = k * T + t;
                    const double2 p = sens_point(sensr, a.cs, i);
                    const double wx = it.cs * p.x - it.sn * p.y + it.x0;
                    const double wy = it.sn * p.x + it.cs * p.y + it.x1;
                    const double2 q1 = refp[cj1[i]], q2 = refp[cj2[i]];
                    const unsigned long long u =
                        (unsigned long long)__double_as_longlong(dist_to_segment(q1.x, q1.y, q2.x, q2.y, wx, wy));
                    d2j[i] = u;
                    const int top = (int)(u >> 56);
                    if (top == top_guess) {
                        const int dg8 = (int)(u >> 48) & 255;
                        atomicAdd(&hist1[dg8], 0x10001);
                        cand1[dg8] = (short)i;
                    } else if (top < top_guess) {
                        nb++;
                    }
                }
                if (__any_sync(0xffffffffu, nb)) {
                    nb = __reduce_add_sync(0xffffffffu, nb);
                    if (lane == 0) atomicAdd(&ctl->sel_below, nb);
                }
            }
            pc.mark(16);
            int order = (int)floor(kcnt * P.outliers_maxPerc);
            order = order < kcnt - 1 ? order : kcnt - 1;
            order = order > 0 ? order : 0;
            int order2 = (int)floor(kcnt * P.outliers_adaptive_order);
            order2 = order2 < kcnt - 1 ? order2 : kcnt - 1;
            order2 = order2 > 0 ? order2 : 0;
            double lim1, d_o2;
            select2<T>(l, C, valid, order, order2, top_guess, &ctl->sel_below, lim1, d_o2, pc);
            if (t == 0) ctl->topbyte = (int)((unsigned long long)__double_as_longlong(lim1) >> 56);
            pc.mark(5);
            const double lim2 = P.outliers_adaptive_mult * d_o2;
            const double error_limit = lim1 < lim2 ? lim1 : lim2;
            // kill + per-thread partial sums: total_error, nvalid, GPC sums (compute_next_estimate + gpc_solve)
            double S[NSUM - 1];
#pragma unroll
            for (int q = 0; q < NSUM - 1; q++) S[q] = 0;
            int nkept = 0;
#pragma unroll 1
            for (int k = 0; k < C; k++) {
                if (!(sens >> k & 1u)) continue;
                const int i = k * T + t;
                bool keep = false;
                if (valid >> k & 1u) {
                    const double key = __longlong_as_double((long long)d2j[i]);
                    if (key > error_limit) {
                        valid &= ~(1u << k);
                    } else {
                        keep = true;
                        nkept++;
                        S[14] += key;
                        const double2 p = sens_point(sensr, a.cs, i);
                        const double2 q1 = refp[cj1[i]], q2 = refp[cj2[i]];
                        const double px = p.x, py = p.y, qx = q1.x, qy = q1.y;
                        const double diff0 = q1.x - q2.x, diff1 = q1.y - q2.y;
                        const double one_on_norm = 1 / sqrt(diff0 * diff0 + diff1 * diff1);
                        const double ca = +diff1 * one_on_norm, sa = -diff0 * one_on_norm;
                        const double C00 = ca * ca, C01 = ca * sa, C11 = sa * sa;
                        S[0] += C00;
                        S[1] += C01;
                        S[2] += +px * C00 + py * C01;
                        S[3] += -py * C00 + px * C01;
                        S[4] += C11;
                        S[5] += +px * C01 + py * C11;
                        S[6] += +px * C11 - py * C01;
                        S[7] += (px * px) * (+C00) + (px * py) * (+C01 + C01) + (py * py) * (+C11);
                        S[8] += (px * px) * (+C01) + (px * py) * (-C00 + C11) + (py * py) * (-C01);
                        S[9] += (px * px) * (+C11) + (px * py) * (-C01 - C01) + (py * py) * (+C00);
                        S[10] += C00 * qx + C01 * qy;
                        S[11] += C01 * qx + C11 * qy;
                        S[12] += qx * (C00 * px + C01 * py) + qy * (C01 * px + C11 * py);
                        S[13] += qx * (C00 * (-py) + C01 * px) + qy * (C01 * (-py) + C11 * px);
                    }
                }
                if (!keep) { cj1[i] = -1; cj2[i] = -1; }
            }
#pragma unroll
            for (int q = 0; q < NSUM - 1; q++) {
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) S[q] += __shfl_down_sync(0xffffffffu, S[q], o);
            }
            if (lane == 0) {
#pragma unroll
                for (int q = 0; q < NSUM - 1; q++) red[warp * NSUM + q] = S[q];
            }
            const int num_after = block_sum<T>(nkept, ctl->ired, 3);  // barrier: publishes red, cj1, cj2
            tma.issue();  // every thread is past the select: the histograms' area is free for the next search's tables
            pc.mark(6);
            // Warp 0 solves; the other warps, which would only wait for it, compute ld_corr_hash of this iteration's
            // correspondences (T chunks over T - 32 threads) and bring the jump tables back for the next search (the
            // select is done with the area they share; an iteration that turns out to be the last copies in vain).
            if (warp != 0) {
                unsigned hc = 0;
#pragma unroll 1
                for (int c = t - 32; c < T; c += T - 32) hc ^= corr_hash_chunk(l, n, c, a.hash_chunk, a.hashmat);
                hc = __reduce_xor_sync(0xffffffffu, hc);
                if (lane == 0 && hc) atomicXor(&ctl->hash_acc[iteration & 1], hc);
                if (jt_src && !tma.on) {
                    uint2 *dst = PL_SH(uint2, l.jt);
                    for (int q = t - 32; q < n; q += T - 32) dst[q] = __ldcs(jt_src + q);
                }
            }
            jt_in_place = true;
            pc.mark(7);
            if (warp == 0) {
                double v = 0;
                if (lane < NSUM - 1)
                    for (int w = 0; w < NW; w++) v += red[w * NSUM + lane];
                double Tt[NSUM - 1];
#pragma unroll
                for (int q = 0; q < NSUM - 1; q++) Tt[q] = __shfl_sync(0xffffffffu, v, q);
                int st = 0;
                double xn[3] = {0, 0, 0};
                pc.mark(11);
                if (num_after < fail_perc * n) st = ZHSM_JOB_TOO_FEW_POINTS;
                else if (!gpc_solve_warp(Tt, xn, &pc)) st = ZHSM_JOB_SINGULAR;
                pc.mark(15);
                if (lane == 0) {
                    ctl->solve_status = st;
                    ctl->error = Tt[14];
                    ctl->nvalid = num_after;
                    if (st == 0) {
                        const double xo[3] = {it.x0, it.x1, ctl->x[par][2]};
                        double delta[3];
                        pose_diff(xn, xo, it.cs, it.sn, delta);
                        const double da = sqrt(delta[0] * delta[0] + delta[1] * delta[1]);
                        const double db = fabs(delta[2]);
                        ctl->eps_ok = (da < P.epsilon_xy && db < P.epsilon_theta) ? 1 : 0;
                        ctl->x[par ^ 1][0] = xn[0];
                        ctl->x[par ^ 1][1] = xn[1];
                        ctl->x[par ^ 1][2] = xn[2];
                        ctl->cs[par ^ 1] = cos(xn[2]);
                        ctl->sn[par ^ 1] = sin(xn[2]);
                    }
                }
            }
            pl_sync<T>();
            pc.mark(8);
            // loop control of icp_loop, evaluated identically by every thread
            const int st = ctl->solve_status;
            int brk = 0;
            if (st) { brk = 2; status = st; }
            else {
                newest = par ^ 1;
                const unsigned h = ctl->hash_acc[iteration & 1] & 0x7FFFFFFFu;
                bool cyc = false;
                if (P.use_point_to_line_distance)
                    for (int q = 0; q < iteration; q++) cyc |= (ctl->hashes[q] == h);
                if (t == 0) ctl->hashes[iteration] = h;
                brk = (cyc || ctl->eps_ok) ? 1 : 0;
            }
            if (brk == 2) all_ok = 0;
            if (brk) break;
            par ^= 1;
        }
        // ---- after icp_loop: sm_icp bookkeeping on thread 0 (icp.c)
        if (t == 0) {
            const int my_iterations = iteration + 1;
            const double loop_error = ctl->error;
            const int loop_valid = ctl->nvalid;
            int next = 0;
            n_loops++;
            if (loop == 0) {
                iterations = my_iterations;
                if (!all_ok) { valid_out = 0; nvalid0 = 0; }
                else {
                    valid_out = 1;
                    status = ZHSM_JOB_OK;
                    best_error = loop_error;
                    nvalid0 = loop_valid;
                    for (int k = 0; k < 3; k++) { best_x[k] = ctl->x[newest][k]; x_first[k] = best_x[k]; }
                    if (P.restart && (loop_error / loop_valid) > P.restart_threshold_mean_error) { restarted = 1; next = 1; }
                }
            } else {
                if (!all_ok) { next = 0; status = ZHSM_JOB_OK; }  // a failing restart breaks the restart loop
                else {
                    iterations += my_iterations;
                    if (loop_error < best_error) {
                        best_error = loop_error;
                        for (int k = 0; k < 3; k++) best_x[k] = ctl->x[newest][k];
                    }
                    next = loop < PLICP_MAX_LOOPS - 1;
                }
            }
            if (next) {
                const int aidx = loop;  // perturbation of restart number loop+1
                const double dt = P.restart_dt, dth = P.restart_dtheta;
                const double pert[6][3] = {{dt, 0, 0}, {-dt, 0, 0}, {0, dt, 0}, {0, -dt, 0}, {0, 0, dth}, {0, 0, -dth}};
                for (int k = 0; k < 3; k++) ctl->x[0][k] = x_first[k] + pert[aidx][k];
                ctl->cs[0] = cos(ctl->x[0][2]);
                ctl->sn[0] = sin(ctl->x[0][2]);
            }
            ctl->next = next;
            ctl->fin = valid_out | (restarted << 1);
        }
        pl_sync<T>();
        if (!ctl->next) break;
    }
    pc.mark(9);
    // ---------------- final correspondences + result record
    const int fin = ctl->fin;
    const bool v_ok = fin & 1, v_restarted = fin & 2;
    pl_sync<T>();
    if (v_ok) {
        if (t == 0) {
            for (int k = 0; k < 3; k++) ctl->x[0][k] = best_x[k];
            ctl->cs[0] = cos(best_x[2]);
            ctl->sn[0] = sin(best_x[2]);
        }
        pl_sync<T>();
        it.x0 = ctl->x[0][0];
        it.x1 = ctl->x[0][1];
        it.cs = ctl->cs[0];
        it.sn = ctl->sn[0];
        // "recompute correspondences in case of restarts" (icp.c): world coords at best x + tricks, no outlier pass
        if (v_restarted) valid = find_correspondences<T>(l, wc, it, C, sens, a.chunk_iters > 0, jt_src, pr, dg, pc, tma, jt_in_place STATS_ARG);
    }
    tma.wait();  // a prefetch issued by the last iteration must have landed before the CTA gives up its shared memory
    if (t == 0) ctl->hash_acc[0] = 0;
#pragma unroll 1
    for (int k = 0; k < C; k++) {
        const int i = k * T + t;
        if (i >= n) break;
        const bool cv = v_ok && (valid >> k & 1u);
        if (!cv) { cj1[i] = -1; cj2[i] = -1; }
        if (oj1) { oj1[i] = cj1[i]; oj2[i] = cj2[i]; }
    }
    pl_sync<T>();
    {
        unsigned hc = corr_hash_chunk(l, n, t, a.hash_chunk, a.hashmat);
        hc = __reduce_xor_sync(0xffffffffu, hc);
        if (lane == 0 && hc) atomicXor(&ctl->hash_acc[0], hc);
    }
    pl_sync<T>();
    const int walks_all = block_sum<T>(dg.walks, ctl->ired, 0);

    const int walks_first = block_sum<T>(dg.first, ctl->ired, 1);
    const int steps_all = block_sum<T>(dg.steps, ctl->ired, 2) + ctl->steps_pend;
#ifdef PLICP_WALK_STATS
    int sts[5];
    for (int k = 0; k < 5; k++) sts[k] = block_sum<T>(st_arr[k], ctl->ired, k & 3), pl_sync<T>();
    int stmax = st_arr[5];
    for (int o = 16; o > 0; o >>= 1) stmax = max(stmax, __shfl_xor_sync(0xffffffffu, stmax, o));
    if (lane == 0) atomicMax(&ctl->qn[0], stmax);
    pl_sync<T>();
    stmax = ctl->qn[0];
#endif
    double cov[6] = {0, 0, 0, 0, 0, 0};
    if (PLICP_NS > 1) {  // this slot has read rr2 and the jump tables for the last time
        pl_sync<T>();
        if (t == 0) { __threadfence_block(); atomicAdd((int *)&msctl[0], 1); }
    }
    if (v_ok && P.do_compute_covariance) {
        if (PLICP_NS > 1) {
            // The covariance takes the SHARED rr2 and jt areas as scratch: wait until every slot of the CTA is done
            // with them, then take the turn.  The slots run independently, so the wait always ends.
            if (t == 0) {
                while (msctl[0] < PLICP_NS) __nanosleep(200);
                while (atomicCAS((int *)&msctl[1], 0, 1) != 0) __nanosleep(200);
                __threadfence_block();
            }
            pl_sync<T>();
        }
        covariance_exact<T>(l, n, C, a.cs, it.x0, it.x1, it.cs, it.sn, P.sigma, cov);
        if (PLICP_NS > 1) {
            pl_sync<T>();
            if (t == 0) { __threadfence_block(); atomicExch((int *)&msctl[1], 0); }
        }
    }
    if (t == 0) {
        zhsm_result r = {};
        r.valid = valid_out;
        r.status = valid_out ? ZHSM_JOB_OK : status;
        r.iterations = iterations;
        r.n_loops = n_loops;
        r.reserved[0] = walks_all;    // diagnostics: ray walks executed (speculation + re-searches)
        r.reserved[1] = walks_first;    // ... of which first-pass walks (= what the sequential algorithm does)
        r.reserved[2] = dg.rounds;      // verify rounds that had rays to re-search (summed over iterations)
        r.reserved[3] = steps_all;      // walk steps executed: distance evaluations of all walks (bench.py: roofline.executed)
        if (valid_out) {
            r.x[0] = best_x[0];
            r.x[1] = best_x[1];
            r.x[2] = best_x[2];
            r.error = best_error;
            r.nvalid = nvalid0;
            r.corr_hash = ctl->hash_acc[0] & 0x7FFFFFFFu;
            for (int q = 0; q < 6; q++) r.cov[q] = cov[q];
        }
#ifdef PLICP_WALK_STATS
        for (int k = 0; k < 5; k++) r.cov[k] = (double)sts[k];
        r.cov[5] = (double)stmax;
#endif
#ifdef PLICP_PHASE_TIMING
        pc.mark(10);
        pc.acc[5] += pc.acc[19];
        for (int k = 0; k < 3; k++) r.x[k] = (double)pc.acc[k];
        for (int k = 0; k < 6; k++) r.cov[k] = (double)pc.acc[3 + k];
        r.error = (double)pc.acc[9];
        r.goodness = (float)pc.acc[10];
        r.quality = (float)pc.acc[11];
        r.nvalid = (int)pc.acc[12];
        r.reserved[3] = (int)pc.acc[13];
        r.corr_hash = (unsigned)pc.acc[14];
        r.n_loops = (int)pc.acc[15];
        for (int k = 0; k < 3; k++) r.reserved[k] = (int)pc.acc[16 + k];
#endif
        *res = r;
    }
}

}  // namespace PL_NAMESPACE
