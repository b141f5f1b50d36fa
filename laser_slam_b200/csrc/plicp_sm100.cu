// laser_slam_b200/csrc/plicp_sm100.cu -- batched PL-ICP for sm_100a: one CTA per job (scan pair + first guess).
//
// Replaces, per job, CCanonicalMatcher::FMatch(LDP,LDP) -> sm_icp (src/zhcsm/ZHCanonical_Matcher.cpp:251-285)
// with the parameters of sm_init (:83-157).  Stages:
//   load       both range rows from HBM (the only global reads), fromPMScan2LDP validity (:193-227),
//              ld_invalid_if_outside, Cartesian points from host-libm cos/sin tables, jump tables
//   iteration  world coords -> find_correspondences_tricks -> kill_outliers_double -> kill_outliers_trim
//              -> GPC sums -> closed-form solve -> ld_corr_hash cycle test / epsilon test
//   restarts   up to 6 perturbed icp_loops, best error wins (strict <), all on device
//
// Mapping: each thread owns C CONSECUTIVE sens rays.  find_correspondences_tricks carries `last_best` from ray
// to ray; inside a thread's chunk that chain is followed exactly, and the value entering the chunk is
// speculated (previous iteration's value) and re-verified against the neighbouring thread until a fixpoint,
// so the sequential semantics are reproduced bit-for-bit while 8 warps search in parallel.
// FP64 throughout (CSM is double end to end); compiled with -fmad=false.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "plicp_solve.cuh"
#include "zhsm_internal.h"

namespace plicp {

constexpr int NSUM = 16;  // 14 GPC sums + total_error + (unused)
constexpr int LB_PASS = -2;  // "no valid match in this chunk": last_best passes through

struct Ctl {
    double x_old[3];
    double x_new[3];
    double cs, sn;       // cos/sin of x_old[2]
    double limit;        // trim threshold
    double sel_val[2];
    unsigned long long sel_prefix[2];
    int sel_rem[2];
    int brk;             // 0 continue, 1 loop ended ok, 2 loop failed
    int next;            // run another icp_loop
    int isum;
    int list_cnt[2];
};

struct Smem {
    double2 *refp;   // [n] ref points (sensor frame of ref = "world")
    double *refr;    // [n] ref readings [m], -1 when invalid
    short *jt;       // [4][n] up_bigger, up_smaller, down_bigger, down_smaller
    double2 *sensp;  // [n] sens points, sensor frame; x = NaN when invalid
    unsigned long long *d2j;  // [n] min dist2 per ref index (kill_outliers_double)
    int *str;        // [n] hash stream: j1 + 1000*j2 or -1
    double *red;     // [nwarps][NSUM]
    int *hist;       // [2][256]
    int *outs;       // [nthreads] last_best leaving each chunk
    int *ired;       // [nwarps]
    Ctl *ctl;
};

__host__ __device__ inline size_t smem_bytes(int n, int nthreads) {
    size_t nw = (nthreads + 31) / 32;
    size_t b = 0;
    b += sizeof(double2) * n;             // refp
    b += sizeof(double) * n;              // refr
    b += sizeof(double2) * n;             // sensp
    b += sizeof(unsigned long long) * n;  // d2j
    b += sizeof(double) * nw * NSUM;      // red
    b += sizeof(Ctl);
    b = (b + 15) & ~(size_t)15;
    b += sizeof(int) * n;                 // str
    b += sizeof(int) * 512;               // hist
    b += sizeof(int) * nthreads;          // outs
    b += sizeof(int) * nw;                // ired
    b += sizeof(short) * 4 * n;           // jt
    return (b + 15) & ~(size_t)15;
}

__device__ inline Smem carve(unsigned char *base, int n, int nthreads) {
    int nw = (nthreads + 31) / 32;
    Smem s;
    unsigned char *p = base;
    s.refp = (double2 *)p;  p += sizeof(double2) * n;
    s.sensp = (double2 *)p; p += sizeof(double2) * n;
    s.refr = (double *)p;   p += sizeof(double) * n;
    s.d2j = (unsigned long long *)p; p += sizeof(unsigned long long) * n;
    s.red = (double *)p;    p += sizeof(double) * nw * NSUM;
    s.ctl = (Ctl *)p;       p += sizeof(Ctl);
    p = (unsigned char *)(((uintptr_t)p + 15) & ~(uintptr_t)15);
    s.str = (int *)p;       p += sizeof(int) * n;
    s.hist = (int *)p;      p += sizeof(int) * 512;
    s.outs = (int *)p;      p += sizeof(int) * nthreads;
    s.ired = (int *)p;      p += sizeof(int) * nw;
    s.jt = (short *)p;
    return s;
}

__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// block-wide integer sum, result to every thread (2 barriers)
__device__ inline int block_sum_i(int v, const Smem &s) {
    int w = warp_sum_i(v);
    if ((threadIdx.x & 31) == 0) s.ired[threadIdx.x >> 5] = w;
    __syncthreads();
    int nw = (blockDim.x + 31) >> 5;
    int t = 0;
    for (int i = 0; i < nw; i++) t += s.ired[i];
    __syncthreads();
    return t;
}

// theta[j] exactly as fromPMScan2LDP builds it (:210): float fi_min + float dfi * (float)j, widened
__device__ __forceinline__ double theta_of(const LaserDev &L, int j) {
    return (double)__fadd_rn(L.fi_min, __fmul_rn(L.dfi, (float)(unsigned)j));
}

__device__ __forceinline__ double dist2(double ax, double ay, double bx, double by) {
    double dx = ax - bx, dy = ay - by;
    return dx * dx + dy * dy;
}

// dist_to_segment_d (math_utils.c), SURVEY 8a-6
__device__ inline double dist_to_segment(double ax, double ay, double bx, double by, double x, double y) {
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

// fromPMScan2LDP (:198-208) + ld_valid_fields' per-ray check + ld_invalid_if_outside.
// returns reading in metres or -1 (invalid); *bad_field set when ld_valid_fields would reject the scan.
__device__ __forceinline__ double reading_of(float r_cm, const LaserDev &L, const zhsm_plicp_params &P, int *bad_field) {
    double r = (double)r_cm;
    if (r <= (double)L.min_range_cm || r >= (double)L.max_range_cm) return -1.0;
    double m = r / 100.0;
    if (!(0.0 < m && m < 100.0)) { *bad_field = 1; return -1.0; }
    if (m <= P.min_reading || m > P.max_reading) return -2.0;  // valid for ld_valid_fields, then invalidated
    return m;
}

}  // namespace plicp

#include "plicp_kernel.cuh"
