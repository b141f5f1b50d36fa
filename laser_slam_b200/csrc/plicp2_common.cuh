// laser_slam_b200/csrc/plicp2_common.cuh -- shared-memory layout and small device helpers of the PL-ICP kernel.
//
// One CTA of T threads owns one job (ref scan, sens scan, first guess).  Thread t owns the sens rays
// i = k*T + t (k < C): the 32 lanes of a warp work on 32 neighbouring rays, so their searches touch
// neighbouring ref points (conflict-free shared-memory reads) and take similar numbers of steps.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "zhsm_internal.h"

namespace plicp2 {

constexpr int NSUM = 16;  // 14 GPC sums + total_error + pad

struct Ctl {
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
};

struct Smem {
    double2 *refp;             // [n] ref points ("world" frame = ref sensor frame); NaN when invalid
    double2 *sensp;            // [n] sens points, sens sensor frame; x = NaN when invalid
    double *rr2;               // [n] squared ref reading [m^2], -1 when invalid
    unsigned long long *d2j;   // [n] min dist2 per ref index (kill_outliers_double); then the trim keys per sens ray
    double *red;               // [nwarps][NSUM]
    Ctl *ctl;
    int *hist;                 // [3][256] radix-select histograms (both ranks packed 16:16)
    short *candi;              // [3][2][256] radix-select: a ray index per bucket
    short *cj1, *cj2;          // [n] current correspondences (-1 none): fixpoint exchange, hash stream, output
    short *jt;                 // [4][n] up_bigger, up_smaller, down_bigger, down_smaller
};

__host__ __device__ inline size_t align16(size_t b) { return (b + 15) & ~(size_t)15; }

__host__ __device__ inline size_t smem_bytes(int n, int nthreads) {
    size_t nw = (nthreads + 31) / 32;
    size_t b = 0;
    b += sizeof(double2) * n * 2;         // refp, sensp
    b += sizeof(double) * n * 2;          // rr2, d2j
    b += sizeof(double) * nw * NSUM;      // red
    b += align16(sizeof(Ctl));
    b += sizeof(int) * 768;               // hist
    b += sizeof(short) * 1536;            // candi
    b += align16(sizeof(short) * n) * 2;  // cj1, cj2
    b += align16(sizeof(short) * 4 * n);  // jt
    return align16(b);
}

__device__ inline Smem carve(unsigned char *p, int n, int nthreads) {
    int nw = (nthreads + 31) / 32;
    Smem s;
    s.refp = (double2 *)p;  p += sizeof(double2) * n;
    s.sensp = (double2 *)p; p += sizeof(double2) * n;
    s.rr2 = (double *)p;    p += sizeof(double) * n;
    s.d2j = (unsigned long long *)p; p += sizeof(unsigned long long) * n;
    s.red = (double *)p;    p += sizeof(double) * nw * NSUM;
    s.ctl = (Ctl *)p;       p += align16(sizeof(Ctl));
    s.hist = (int *)p;      p += sizeof(int) * 768;
    s.candi = (short *)p;   p += sizeof(short) * 1536;
    s.cj1 = (short *)p;     p += align16(sizeof(short) * n);
    s.cj2 = (short *)p;     p += align16(sizeof(short) * n);
    s.jt = (short *)p;
    return s;
}

// theta[j] exactly as fromPMScan2LDP builds it (src/zhcsm/ZHCanonical_Matcher.cpp:210): float fi_min + float dfi*j
__device__ __forceinline__ float theta_f(const LaserDev &L, int j) {
    return __fadd_rn(L.fi_min, __fmul_rn(L.dfi, (float)(unsigned)j));
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

}  // namespace plicp2
