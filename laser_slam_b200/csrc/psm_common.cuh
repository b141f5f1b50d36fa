// laser_slam_b200/csrc/psm_common.cuh -- shared-memory layout and the scan projection of the zhpsm matchers
// (CPolarMatch, src/zhpsm/ZHPolar_Match.cpp).  All arithmetic is float (PM_TYPE, PolarParameter.h:17) written with the
// reference's own expression shapes so that C++'s float/double promotions happen at the same places; compiled with
// -fmad=false; sinf/cosf/atan2f come from zh_libm.cuh (bit-identical to the host libm the oracle links).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "zh_libm.cuh"
#include "zhsm_internal.h"

namespace psm {

constexpr int T = 128;               // threads per CTA (one CTA per job)
constexpr double PI_BODY = 3.14159265358979323846;  // M_PI inside ZHPolar_Match.cpp's function bodies (math.h's)
constexpr double PI_HDR = 3.141592654;              // M_PI where PolarParameter.h's #define is active (norm_a, PM_D2R)
constexpr int PM_EMPTY = 16, PM_OCCLUDED = 8, PM_RANGE = 1;
constexpr unsigned long long EMPTY_KEY = ~0ull;

struct Sh {
    float *ref_r, *act_r, *co, *si, *fi;   // [n] ranges (cm) and the pm_co/pm_si/pm_fi tables
    float *pr, *pfi, *new_r;               // [n] projected range / bearing, interpolated range
    unsigned long long *key;               // [n] scatter-min keys of pm_scan_project
    float *nx, *ny, *refx, *refy, *actx, *acty, *dist, *ix, *iy, *fx, *fy;  // [n] pm_icp
    int *pairs, *lastpos;                  // [n] pm_icp: (i | j << 16) per match, last kept position per ref index
    float *tmpd;                           // [n] pm_icp: match distance per act index before compaction
    short *tmpj;                           // [n] pm_icp: matched ref index per act index (-1 none)
    short *act_seg;                        // [n]
    unsigned char *ref_bad, *act_bad, *new_bad;  // [n]
    float *err;                            // [n] orientation search error per shift
    int *scan;                             // [T + 32] block scan scratch
    float *ctl;                            // [32] scalars broadcast between phases
};

__host__ __device__ inline size_t sh_bytes(int n) {
    size_t a = (size_t)((n + 3) & ~3);
    return a * 4 * 19 + a * 8 + a * 4 * 2 + a * 4 + a * 2 * 2 + a * 3 + a * 4 + (T + 32) * 4 + 32 * 4 + 64;
}

__device__ inline Sh carve(unsigned char *p, int n) {
    const size_t a = (size_t)((n + 3) & ~3);
    Sh s;
    s.key = (unsigned long long *)p; p += a * 8;
    float **fl[] = {&s.ref_r, &s.act_r, &s.co, &s.si, &s.fi, &s.pr, &s.pfi, &s.new_r, &s.nx, &s.ny, &s.refx, &s.refy,
                    &s.actx, &s.acty, &s.dist, &s.ix, &s.iy, &s.fx, &s.fy};
    for (int k = 0; k < 19; k++) { *fl[k] = (float *)p; p += a * 4; }
    s.pairs = (int *)p; p += a * 4;
    s.lastpos = (int *)p; p += a * 4;
    s.tmpd = (float *)p; p += a * 4;
    s.err = (float *)p; p += a * 4;
    s.tmpj = (short *)p; p += a * 2;
    s.act_seg = (short *)p; p += a * 2;
    s.ref_bad = p; p += a;
    s.act_bad = p; p += a;
    s.new_bad = p; p += a;
    p = (unsigned char *)(((uintptr_t)p + 15) & ~(uintptr_t)15);
    s.scan = (int *)p; p += (T + 32) * 4;
    s.ctl = (float *)p;
    return s;
}

// order-preserving map float -> uint32 (all finite floats, negative ones included)
__device__ __forceinline__ unsigned order_key(float v) {
    const unsigned u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float order_val(unsigned k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// pm_scan_project (ZHPolar_Match.cpp:1356-1479): project the act scan at pose (ax, ay, ath) into the ref frame and
// resample it at the ref bearings.  The reference walks the point pairs in order and writes a bin iff the new range is
// strictly smaller, i.e. every bin ends with the lexicographic minimum of (range, pair index): a 64-bit shared-memory
// atomicMin on (ordered range bits << 32 | pair index).  Flags follow from the winning pair.  Whole-CTA call.
__device__ __forceinline__ void scan_project(const Sh &s, const LaserDev &L, float ax, float ay, float ath) {
    const int n = L.n, t = threadIdx.x;
    for (int i = t; i < n; i += T) {
        const float delta = ath + s.fi[i];
        const float x = s.act_r[i] * zhm::cosf_glibc(delta) + ax;
        const float y = s.act_r[i] * zhm::sinf_glibc(delta) + ay;
        s.pr[i] = sqrtf(x * x + y * y);
        float f = zhm::atan2f_glibc(y, x);
        if (x < 0 && y < 0) f = (float)((double)f + 2.0 * PI_BODY);
        s.pfi[i] = f;
        s.key[i] = EMPTY_KEY;
    }
    __syncthreads();
    for (int i = 1 + t; i < n; i += T) {
        if (!(s.act_seg[i] != 0 && s.act_seg[i] == s.act_seg[i - 1] && !s.act_bad[i] && !s.act_bad[i - 1])) continue;
        const float fi1 = s.pfi[i], fi0 = s.pfi[i - 1];
        if ((double)fabsf(fi1 - fi0) >= PI_BODY) continue;
        float a0, a1, r0, r1;
        int j0, j1;
        if (fi1 > fi0) {
            a0 = fi0; a1 = fi1;
            j0 = (int)ceilf((fi0 - L.fi_min) / L.dfi);
            j1 = (int)floorf((fi1 - L.fi_min) / L.dfi);
            r0 = s.pr[i - 1]; r1 = s.pr[i];
        } else {
            a0 = fi1; a1 = fi0;
            j0 = (int)ceilf((fi1 - L.fi_min) / L.dfi);
            j1 = (int)floorf((fi0 - L.fi_min) / L.dfi);
            r0 = s.pr[i]; r1 = s.pr[i - 1];
        }
        while (j0 <= j1) {
            const float ri = (r1 - r0) / (a1 - a0) * (((float)j0 * L.dfi + L.fi_min) - a0) + r0;
            if (j0 >= 0 && j0 < n && 10000.0f > ri)
                atomicMin(&s.key[j0], ((unsigned long long)order_key(ri) << 32) | (unsigned)i);
            j0++;
        }
    }
    __syncthreads();
    for (int j = t; j < n; j += T) {
        const unsigned long long k = s.key[j];
        if (k == EMPTY_KEY) {
            s.new_r[j] = 10000.0f;
            s.new_bad[j] = PM_EMPTY;
        } else {
            const int i = (int)(k & 0xffffffffull);
            s.new_r[j] = order_val((unsigned)(k >> 32));
            s.new_bad[j] = (s.pfi[i] > s.pfi[i - 1]) ? 0 : PM_OCCLUDED;
        }
    }
    __syncthreads();
}

// norm_a (ZHPolar_Match.h:77-87), including its "m*pi instead of m*2pi" slip
__device__ __forceinline__ float norm_a(float a) {
    const int m = (int)((double)a / (2.0 * PI_HDR));
    a = (float)((double)a - (double)(float)m * PI_HDR);
    if ((double)a < -PI_HDR) a = (float)((double)a + 2.0 * PI_HDR);
    if ((double)a >= PI_HDR) a = (float)((double)a - 2.0 * PI_HDR);
    return a;
}

}  // namespace psm
