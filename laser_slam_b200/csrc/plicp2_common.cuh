// laser_slam_b200/csrc/plicp2_common.cuh -- shared-memory layout and small device helpers of the PL-ICP kernel.
//
// One CTA of T threads owns one job (ref scan, sens scan, first guess).  Thread t owns the sens rays
// i = k*T + t (k < C = ceil(n/T)): the 32 lanes of a warp work on 32 neighbouring rays, so their searches touch
// neighbouring ref points (conflict-free shared-memory reads) and take similar numbers of steps.
//
// Everything a job needs lives in dynamic shared memory, addressed as byte offsets from the extern array so that
// every access, also inside non-inlined device functions, compiles to LDS/STS (a generic pointer costs a long
// scoreboard stall per load).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stddef.h>
#include <stdint.h>

#include "zhsm_internal.h"

// Jobs per CTA.  1 (every instance but one): a CTA owns one job.  4 (plicp2_ms128.cu, the multi-start instance): a CTA
// owns four jobs ("slots") that share BOTH scans -- four first guesses of one scan pair -- as four groups of T threads
// over ONE staged copy of the ref scan's points, squared readings and jump tables and of the sens readings; everything
// else (correspondences, queues, histograms, sums, control block) is per slot, and the slots never wait for each
// other after the load: every barrier of the job is a named barrier of its own group (pl_sync).  Each instance file is
// its own namespace, so the single-job instances are untouched by the other's existence.
#ifndef PLICP_NS
#define PLICP_NS 1
#endif
#ifndef PL_NAMESPACE
#define PL_NAMESPACE plicp2
#endif

namespace PL_NAMESPACE {

constexpr int NSUM = 16;  // 14 GPC sums + total_error + pad

// thread within its job / the job's slot within the CTA / barriers and votes over the T threads of one job
template <int T>
__device__ __forceinline__ int pl_tid() { return PLICP_NS == 1 ? (int)threadIdx.x : (int)(threadIdx.x % T); }
template <int T>
__device__ __forceinline__ int pl_slot() { return PLICP_NS == 1 ? 0 : (int)(threadIdx.x / T); }
template <int T>
__device__ __forceinline__ void pl_sync() {
    if (PLICP_NS == 1) __syncthreads();
    else asm volatile("bar.sync %0, %1;" ::"r"(pl_slot<T>() + 1), "r"(T) : "memory");  // barrier 0 stays the CTA's
}
template <int T>
__device__ __forceinline__ int pl_sync_or(int v) {
    if (PLICP_NS == 1) return __syncthreads_or(v);
    int r;
    asm volatile("{\n\t"
                 ".reg .pred p, q;\n\t"
                 "setp.ne.s32 p, %1, 0;\n\t"
                 "bar.red.or.pred q, %2, %3, p;\n\t"
                 "selp.s32 %0, 1, 0, q;\n\t"
                 "}" : "=r"(r) : "r"(v), "r"(pl_slot<T>() + 1), "r"(T) : "memory");
    return r;
}

struct Ctl {
    unsigned long long mbar;  // mbarrier of the bulk (TMA) copy that brings the jump tables back (PLICP_TMA_JT builds)
    double x[2][3];     // pose ping-pong: x[p] is x_old of the iteration with parity p, x[p^1] receives x_new
    double cs[2], sn[2];  // cos/sin of x[p][2]
    double error;       // total_error of the last iteration
    unsigned hashes[ZHSM_MAX_ITER];
    unsigned hash_acc[2];
    int ired[4 * 16];   // block_sum scratch rows
    int eps_ok;         // |delta| below epsilon_xy / epsilon_theta
    int solve_status;   // 0 ok, else ZHSM_JOB_* failure of this iteration
    int nvalid;         // correspondences kept by the last iteration
    int next;           // run another icp_loop
    int fin;            // valid | restarted<<1
    int qn[2];          // re-search queue length (ping-pong over verify rounds)
    int topbyte;        // radix select: guessed top byte of the trim keys (that of the last result)
    int sel_below;      // radix select: keys whose top byte is below the guess
    int npend;          // walks on the list of unfinished walks (plicp2_search.cuh)
    int pend_next;      // next list entry to be taken by a warp
    int steps_pend;     // walk steps of the walks that went through the list (finish_unfinished)
    int work_next;      // next item of a round's pool of walks (plicp2_flat.cuh)
};

// byte offsets into the dynamic shared memory block
struct Lay {
    unsigned refp;   // double2[n] ref points ("world" frame = ref sensor frame); NaN when invalid
    unsigned sensr;  // float[n]  sens readings [cm] as they came, -1 when invalid: the point is rebuilt on use (sens_point)
    unsigned rr2;    // double[n]  squared ref reading [m^2], -1 when invalid
    unsigned d2j;    // u64[n]     min dist2 per ref index (kill_outliers_double); then the trim keys per sens ray
    unsigned red;    // double[nwarps][NSUM]
    unsigned ctl;    // Ctl
    unsigned hist;   // int[3][256]   radix-select histograms (both ranks packed 16:16); on the jump tables when they fit
    unsigned candi;  // short[3][2][256] radix-select: a ray index per bucket
    unsigned cj1;    // short[n] current correspondences (-1 none): fixpoint exchange, hash stream, output
    unsigned cj2;    // short[n]
    unsigned spec;   // short[n] last_best each ray was (or will be) searched with
    unsigned jt;     // short[2][n][2] pairs (up_bigger, up_smaller)[n], (down_bigger, down_smaller)[n]
    unsigned msctl;  // int[4], multi-slot layout only: [0] slots that are done with the shared arrays, [1] covariance turn lock
    unsigned total;
    bool jt_shared;  // hist/candi lie on top of jt: the jump tables are copied in again before every search
};

__host__ __device__ inline unsigned align16(unsigned b) { return (b + 15u) & ~15u; }

__host__ __device__ inline Lay make_layout(int n, int nthreads) {
    const unsigned nw = (unsigned)(nthreads + 31) / 32, un = (unsigned)n;
    Lay l;
    unsigned o = 0;
    l.refp = o;  o += 16u * un;
    l.sensr = o; o += align16(4u * un);
    l.rr2 = o;   o += 8u * un;
    l.d2j = o;   o += 8u * un;
    l.red = o;   o += 8u * nw * NSUM;
    l.ctl = o;   o += align16((unsigned)sizeof(Ctl));
    l.jt_shared = 8u * un >= 4u * 768 + 2u * 1536;
    if (!l.jt_shared) {
        l.hist = o;  o += 4u * 768;
        l.candi = o; o += 2u * 1536;
    }
    l.cj1 = o;   o += align16(2u * un);
    l.cj2 = o;   o += align16(2u * un);
    l.spec = o;  o += align16(2u * un);
    l.jt = o;    o += align16(8u * un);
    if (l.jt_shared) { l.hist = l.jt; l.candi = l.jt + 4u * 768; }
    l.msctl = 0;
    l.total = align16(o);
    return l;
}

// The multi-slot layout (PLICP_NS jobs per CTA): [refp | sensr | rr2 | jt] once, then per slot [d2j | red | ctl | hist |
// candi | cj1 | cj2 | spec].  The jump tables stay in place for the whole job (the histograms have their own bytes), so
// nothing is copied in again before a search.  92 n + 30 KB at four slots: 218 KB at 2048 beams.  The covariance (plicp2_cov.cuh)
// takes rr2 and jt as scratch: a slot computes it only when all slots are done iterating, one slot at a time (msctl).
__host__ __device__ inline unsigned slot_bytes(int n, int nthreads) {
    const unsigned nw = (unsigned)(nthreads + 31) / 32, un = (unsigned)n;
    return align16(8u * un) + 8u * nw * NSUM + align16((unsigned)sizeof(Ctl)) + 4u * 768 + 2u * 1536 + 3u * align16(2u * un);
}
__host__ __device__ inline Lay make_layout_slots(int n, int nthreads, int nslots, int slot) {
    const unsigned nw = (unsigned)(nthreads + 31) / 32, un = (unsigned)n;
    Lay l;
    unsigned o = 0;
    l.refp = o;  o += 16u * un;
    l.sensr = o; o += align16(4u * un);
    l.rr2 = o;   o += align16(8u * un);
    l.jt = o;    o += align16(8u * un);
    l.msctl = o; o += 16u;
    const unsigned per = slot_bytes(n, nthreads);
    l.total = align16(o + per * (unsigned)nslots);
    o += per * (unsigned)slot;
    l.d2j = o;   o += align16(8u * un);
    l.red = o;   o += 8u * nw * NSUM;
    l.ctl = o;   o += align16((unsigned)sizeof(Ctl));
    l.hist = o;  o += 4u * 768;
    l.candi = o; o += 2u * 1536;
    l.cj1 = o;   o += align16(2u * un);
    l.cj2 = o;   o += align16(2u * un);
    l.spec = o;
    l.jt_shared = false;
    return l;
}

__host__ __device__ inline size_t smem_bytes(int n, int nthreads) {
    return PLICP_NS == 1 ? make_layout(n, nthreads).total : make_layout_slots(n, nthreads, PLICP_NS, 0).total;
}

extern __shared__ __align__(16) unsigned char smem_raw[];
#define PL_SH(type, off) (reinterpret_cast<type *>(smem_raw + (off)))

// ---- 1-D bulk copy global -> shared through the TMA unit, completion on an mbarrier (sm_90+: cp.async.bulk) ----------
// One thread arms the barrier with the byte count and issues the copy; every thread that needs the data waits for the
// barrier's phase.  src, dst and bytes must be multiples of 16.
__device__ __forceinline__ void mbar_init(unsigned mbar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void bulk_load_1d(unsigned dst, const void *src, unsigned bytes, unsigned mbar) {
    // shared memory written before by ordinary stores / atomics: order those against the async proxy's writes
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned mbar, unsigned phase) {
    asm volatile("{\n\t"
                 ".reg .pred p;\n\t"
                 "MBAR_WAIT:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                 "@p bra MBAR_DONE;\n\t"
                 "bra MBAR_WAIT;\n\t"
                 "MBAR_DONE:\n\t"
                 "}" ::"r"(mbar), "r"(phase) : "memory");
}

// theta[j] exactly as fromPMScan2LDP builds it (src/zhcsm/ZHCanonical_Matcher.cpp:210): float fi_min + float dfi*j
__device__ __forceinline__ float theta_f(float fi_min, float dfi, int j) {
    return __fadd_rn(fi_min, __fmul_rn(dfi, (float)(unsigned)j));
}

__device__ __forceinline__ double dist2(double ax, double ay, double bx, double by) {
    double dx = ax - bx, dy = ay - by;
    return dx * dx + dy * dy;
}

// dist_to_segment_d (CSM math_utils.c), SURVEY 8a-6
__device__ __forceinline__ double dist_to_segment(double ax, double ay, double bx, double by, double x, double y) {
    double t0 = ax - bx;
    double t1 = ay - by;
    double one_on_r = 1 / sqrt(t0 * t0 + t1 * t1);
    double nx = t1 * one_on_r;
    double ny = -t0 * one_on_r;
    double c = nx, s = ny;
    double rho = c * ax + s * ay;
    double projx = c * rho + s * s * x - c * s * y;
    double projy = s * rho - c * s * x + c * c * y;
    double distance = fabs(rho - (c * x + s * y));
    if ((projx - ax) * (projx - bx) + (projy - ay) * (projy - by) < 0) return distance;
    double da = dist2(ax, ay, x, y), db = dist2(bx, by, x, y);
    return sqrt(da < db ? da : db);
}

// A sens point in the sens sensor frame, rebuilt from its float reading [cm] and the (cos, sin) table of the handle
// (global memory, L1-resident) exactly as the load phase of the reference builds it: m = (double)f / 100.0, then
// (cos * m, sin * m).  The quotient is formed without a division: q0 = f * 0.01, one FMA residual, one FMA correction
// is the correctly rounded f / 100 (Markstein; checked exhaustively for every float in [2^-20, 2^20)).  Keeping 4
// bytes per sens ray instead of a double2 is what lets one more CTA fit an SM.  Valid rays only.
__device__ __forceinline__ double2 sens_point(const float *sensr, const double2 *cs, int i) {
    const double r = (double)sensr[i];
    const double q0 = r * 0.01;
    const double m = fma(fma(-q0, 100.0, r), 0.01, q0);
    const double2 c = __ldg(cs + i);
    return make_double2(c.x * m, c.y * m);
}

// fromPMScan2LDP (:198-208) + ld_valid_fields' per-ray check + ld_invalid_if_outside.
// returns the reading in metres, -1 (invalid in the LDP) or -2 (valid for ld_valid_fields, then invalidated by
// ld_invalid_if_outside); *bad_field set when ld_valid_fields would reject the scan.
__device__ __forceinline__ double reading_of(float r_cm, const LaserDev &L, const zhsm_plicp_params &P, int *bad_field) {
    double r = (double)r_cm;
    if (r <= (double)L.min_range_cm || r >= (double)L.max_range_cm) return -1.0;
    double m = r / 100.0;
    if (!(0.0 < m && m < 100.0)) { *bad_field = 1; return -1.0; }
    if (m <= P.min_reading || m > P.max_reading) return -2.0;
    return m;
}

}  // namespace PL_NAMESPACE
