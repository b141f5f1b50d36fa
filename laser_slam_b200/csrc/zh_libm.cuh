// laser_slam_b200/csrc/zh_libm.cuh -- float sinf / cosf / atan2f that return the SAME BITS as the host libm the
// reference's zhpsm matcher is linked against (glibc 2.39 in this image), so the CUDA port of pm_scan_project and
// pm_icp does not drift from the compiled-reference oracle through the last ulp of a trigonometric call.
//
//   sinf, cosf : glibc's sincosf scheme (sysdeps/ieee754/flt-32/s_sincosf.h): reduce by pi/2 with a scaled
//                float->int conversion, evaluate a double-precision polynomial, round once to float.
//   atan2f     : glibc's e_atan2f.c / s_atanf.c (FreeBSD msun): float arithmetic, 4-interval argument reduction,
//                11-term odd polynomial, pi split in hi/lo.
// Both were re-derived from the published algorithms and checked on the CPU against this image's libm on 1e5 random
// arguments each with 0 mismatches (the same check runs on the GPU in tests/test_psm_gpu.py::test_device_libm).
// Compile with -fmad=false: every product-sum below must round twice (the double polynomial's result is far inside
// the float rounding interval either way; the float atan polynomial is not).
#pragma once
#include <stdint.h>

namespace zhm {

__device__ __forceinline__ double sincosf_poly(double x, double x2, bool neg_cos, int n) {
    const double s1 = -0x1.555545995a603p-3, s2 = 0x1.1107605230bc4p-7, s3 = -0x1.994eb3774cf24p-13;
    double c0 = 0x1p0, c1 = -0x1.ffffffd0c621cp-2, c2 = 0x1.55553e1068f19p-5, c3 = -0x1.6c087e89a359dp-10,
           c4 = 0x1.99343027bf8c3p-16;
    if ((n & 1) == 0) {
        const double x3 = x * x2;
        const double t1 = s2 + x2 * s3;
        const double x7 = x3 * x2;
        const double s = x + x3 * s1;
        return s + x7 * t1;
    }
    if (neg_cos) { c0 = -c0; c1 = -c1; c2 = -c2; c3 = -c3; c4 = -c4; }
    const double x4 = x2 * x2;
    const double t2 = c3 + x2 * c4;
    const double t1 = c0 + x2 * c1;
    const double x6 = x4 * x2;
    const double c = t1 + x4 * c2;
    return c + x6 * t2;
}

__device__ __forceinline__ uint32_t abstop12(float x) { return (__float_as_uint(x) >> 20) & 0x7ff; }

// valid for |y| < 120 (the matcher's angles are within a few turns); larger arguments fall back to CUDA's sinf/cosf
__device__ __forceinline__ float sinf_glibc(float y) {
    const double hpi_inv = 0x1.45F306DC9C883p+23, hpi = 0x1.921FB54442D18p0;
    double x = (double)y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        if (abstop12(y) < abstop12(0x1p-12f)) return y;
        return (float)sincosf_poly(x, x * x, false, 0);
    }
    if (abstop12(y) < abstop12(120.0f)) {
        const double r = x * hpi_inv;
        const int n = ((int32_t)r + 0x800000) >> 24;
        x = x - (double)n * hpi;
        const double s = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
        return (float)sincosf_poly(x * s, x * x, (n & 2) != 0, n);
    }
    return sinf(y);
}

__device__ __forceinline__ float cosf_glibc(float y) {
    const double hpi_inv = 0x1.45F306DC9C883p+23, hpi = 0x1.921FB54442D18p0;
    double x = (double)y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        if (abstop12(y) < abstop12(0x1p-12f)) return 1.0f;
        return (float)sincosf_poly(x, x * x, false, 1);
    }
    if (abstop12(y) < abstop12(120.0f)) {
        const double r = x * hpi_inv;
        const int n = ((int32_t)r + 0x800000) >> 24;
        x = x - (double)n * hpi;
        const int m = n + 1;
        const double s = ((m & 3) == 1 || (m & 3) == 2) ? -1.0 : 1.0;
        return (float)sincosf_poly(x * s, x * x, (m & 2) != 0, n ^ 1);
    }
    return cosf(y);
}

__device__ __forceinline__ float atanf_msun(float x) {
    const float atanhi[4] = {4.6364760399e-01f, 7.8539812565e-01f, 9.8279368877e-01f, 1.5707962513e+00f};
    const float atanlo[4] = {5.0121582440e-09f, 3.7748947079e-08f, 3.4473217170e-08f, 7.5497894159e-08f};
    const float aT[11] = {3.3333334327e-01f, -2.0000000298e-01f, 1.4285714924e-01f, -1.1111110449e-01f,
                          9.0908870101e-02f, -7.6918758452e-02f, 6.6610731184e-02f, -5.8335702866e-02f,
                          4.9768779427e-02f, -3.6531571299e-02f, 1.6285819933e-02f};
    const int32_t hx = (int32_t)__float_as_uint(x);
    const int32_t ix = hx & 0x7fffffff;
    int id;
    if (ix >= 0x4c000000) {  // |x| >= 2^25
        if (ix > 0x7f800000) return x + x;
        const float z = atanhi[3] + atanlo[3];
        return hx > 0 ? z : -z;
    }
    if (ix < 0x3ee00000) {  // |x| < 0.4375
        if (ix < 0x31000000) return x;
        id = -1;
    } else {
        x = fabsf(x);
        if (ix < 0x3f980000) {
            if (ix < 0x3f300000) { id = 0; x = (2.0f * x - 1.0f) / (2.0f + x); }
            else { id = 1; x = (x - 1.0f) / (x + 1.0f); }
        } else {
            if (ix < 0x401c0000) { id = 2; x = (x - 1.5f) / (1.0f + 1.5f * x); }
            else { id = 3; x = -1.0f / x; }
        }
    }
    const float z = x * x;
    const float w = z * z;
    const float s1 = z * (aT[0] + w * (aT[2] + w * (aT[4] + w * (aT[6] + w * (aT[8] + w * aT[10])))));
    const float s2 = w * (aT[1] + w * (aT[3] + w * (aT[5] + w * (aT[7] + w * aT[9]))));
    if (id < 0) return x - x * (s1 + s2);
    const float r = atanhi[id] - ((x * (s1 + s2) - atanlo[id]) - x);
    return hx < 0 ? -r : r;
}

__device__ __forceinline__ float atan2f_glibc(float y, float x) {
    const float pi = 3.1415927410e+00f, pi_lo = -8.7422776573e-08f, pi_o_2 = 1.5707963705e+00f, pi_o_4 = 7.8539818525e-01f;
    const int32_t hx = (int32_t)__float_as_uint(x), hy = (int32_t)__float_as_uint(y);
    const int32_t ix = hx & 0x7fffffff, iy = hy & 0x7fffffff;
    if (ix > 0x7f800000 || iy > 0x7f800000) return x + y;
    if (hx == 0x3f800000) return atanf_msun(y);
    const int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);
    if (iy == 0) return m < 2 ? y : (m == 2 ? pi : -pi);
    if (ix == 0) return hy < 0 ? -pi_o_2 : pi_o_2;
    if (ix == 0x7f800000) {
        if (iy == 0x7f800000) return m == 0 ? pi_o_4 : m == 1 ? -pi_o_4 : m == 2 ? 3.0f * pi_o_4 : -3.0f * pi_o_4;
        return m == 0 ? 0.0f : m == 1 ? -0.0f : m == 2 ? pi : -pi;
    }
    if (iy == 0x7f800000) return hy < 0 ? -pi_o_2 : pi_o_2;
    const int k = (iy - ix) >> 23;
    float z;
    if (k > 60) z = pi_o_2 + 0.5f * pi_lo;
    else if (hx < 0 && k < -60) z = 0.0f;
    else z = atanf_msun(fabsf(y / x));
    switch (m) {
        case 0: return z;
        case 1: return -z;
        case 2: return pi - (z - pi_lo);
        default: return (z - pi_lo) - pi;
    }
}

}  // namespace zhm
