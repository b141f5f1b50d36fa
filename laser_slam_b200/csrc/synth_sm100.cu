// laser_slam_b200/csrc/synth_sm100.cu -- workload tooling that lives on the device:
//  (1) the synthetic scan-pair generator of SURVEY 8d configs C4/C5 (room + boxes, ray cast, noise, dropouts),
//      Philox counter = pair index, generated straight into HBM and excluded from timing;
//  (2) FFMA / DFMA peak micro-benchmarks: the measured CUDA-core roofline denominators bench.py reports against.
#include <cuda_runtime.h>
#include <curand_kernel.h>
#include <math.h>

#include "zhsm_internal.h"

struct SynthWorld {
    float W, H;          // room [0,W]x[0,H]
    int nbox;
    float bx0[12], by0[12], bx1[12], by1[12];
    float ax, ay, ath;   // pose A (ref)
    float bx, by, bth;   // pose B (sens)
};

__device__ inline bool synth_free(const SynthWorld &w, float x, float y, float clr) {
    if (x < clr || y < clr || x > w.W - clr || y > w.H - clr) return false;
    for (int b = 0; b < w.nbox; b++)
        if (x > w.bx0[b] - clr && x < w.bx1[b] + clr && y > w.by0[b] - clr && y < w.by1[b] + clr) return false;
    return true;
}

// distance along the ray (ox,oy)+t(dx,dy) to the inside of the room walls / outside of the boxes
__device__ inline float synth_cast(const SynthWorld &w, float ox, float oy, float dx, float dy) {
    float t = 1e30f;
    // room walls (we are inside): exit distance
    float tx = dx > 0 ? (w.W - ox) / dx : (dx < 0 ? (0.f - ox) / dx : 1e30f);
    float ty = dy > 0 ? (w.H - oy) / dy : (dy < 0 ? (0.f - oy) / dy : 1e30f);
    t = fminf(tx, ty);
    for (int b = 0; b < w.nbox; b++) {
        // slab test against an axis-aligned box (we are outside)
        float t0 = 0.f, t1 = t;
        bool hit = true;
        if (dx != 0.f) {
            float a = (w.bx0[b] - ox) / dx, c = (w.bx1[b] - ox) / dx;
            t0 = fmaxf(t0, fminf(a, c)); t1 = fminf(t1, fmaxf(a, c));
        } else if (ox < w.bx0[b] || ox > w.bx1[b]) hit = false;
        if (dy != 0.f) {
            float a = (w.by0[b] - oy) / dy, c = (w.by1[b] - oy) / dy;
            t0 = fmaxf(t0, fminf(a, c)); t1 = fminf(t1, fmaxf(a, c));
        } else if (oy < w.by0[b] || oy > w.by1[b]) hit = false;
        if (hit && t0 <= t1 && t0 > 0.f) t = fminf(t, t0);
    }
    return t;
}

// one block per pair; scans[2p] = ref at pose A, scans[2p+1] = sens at pose B = A (+) delta; truth[p] = delta
__global__ void synth_pairs_kernel(float *scans, double *truth, long long first_pair, int npairs, int n, float fi_min,
                                   float dfi, int max_range_cm, float dmax_xy, float dmax_th, float noise_m,
                                   float dropout, unsigned long long seed) {
    __shared__ SynthWorld w;
    const int p = blockIdx.x;
    if (p >= npairs) return;
    const unsigned long long pid = (unsigned long long)(first_pair + p);
    if (threadIdx.x == 0) {
        curandStatePhilox4_32_10_t st;
        curand_init(seed, pid, 0, &st);
        w.W = 6.f + 14.f * curand_uniform(&st);
        w.H = 6.f + 14.f * curand_uniform(&st);
        w.nbox = 4 + (int)(curand(&st) % 9u);
        for (int b = 0; b < w.nbox; b++) {
            float sx = 0.3f + 1.7f * curand_uniform(&st), sy = 0.3f + 1.7f * curand_uniform(&st);
            float cx = curand_uniform(&st) * (w.W - sx), cy = curand_uniform(&st) * (w.H - sy);
            w.bx0[b] = cx; w.bx1[b] = cx + sx; w.by0[b] = cy; w.by1[b] = cy + sy;
        }
        for (int b = w.nbox; b < 12; b++) { w.bx0[b] = w.bx1[b] = w.by0[b] = w.by1[b] = -1.f; }
        // pose A in free space with 0.5 m clearance; pose B must be free too (0.2 m)
        for (int tries = 0; tries < 256; tries++) {
            w.ax = curand_uniform(&st) * w.W; w.ay = curand_uniform(&st) * w.H;
            w.ath = (curand_uniform(&st) * 2.f - 1.f) * 3.14159265f;
            float ddx = (curand_uniform(&st) * 2.f - 1.f) * dmax_xy, ddy = (curand_uniform(&st) * 2.f - 1.f) * dmax_xy;
            float ddt = (curand_uniform(&st) * 2.f - 1.f) * dmax_th;
            w.bx = w.ax + cosf(w.ath) * ddx - sinf(w.ath) * ddy;
            w.by = w.ay + sinf(w.ath) * ddx + cosf(w.ath) * ddy;
            w.bth = w.ath + ddt;
            truth[3 * (size_t)p + 0] = ddx; truth[3 * (size_t)p + 1] = ddy; truth[3 * (size_t)p + 2] = ddt;
            if (synth_free(w, w.ax, w.ay, 0.5f) && synth_free(w, w.bx, w.by, 0.2f)) break;
            if (tries == 255) { w.nbox = 0; w.ax = w.W * 0.5f; w.ay = w.H * 0.5f; w.bx = w.ax + ddx; w.by = w.ay + ddy; w.ath = 0; w.bth = ddt;
                                truth[3 * (size_t)p + 0] = ddx; truth[3 * (size_t)p + 1] = ddy; }
        }
    }
    __syncthreads();
    for (int which = 0; which < 2; which++) {
        const float ox = which ? w.bx : w.ax, oy = which ? w.by : w.ay, oth = which ? w.bth : w.ath;
        float *row = scans + ((size_t)2 * p + which) * n;
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            curandStatePhilox4_32_10_t st;
            curand_init(seed ^ 0x9E3779B97F4A7C15ull, pid * 2 + which, (unsigned long long)i * 4, &st);
            float ang = oth + fi_min + dfi * (float)i;
            float t = synth_cast(w, ox, oy, cosf(ang), sinf(ang));
            float r = t + noise_m * curand_normal(&st);
            r = rintf(r * 1000.f) * 0.001f;  // 1 mm quantisation
            float cm = r * 100.f;
            bool drop = curand_uniform(&st) < dropout;
            if (drop || cm >= (float)max_range_cm || cm < 10.f) cm = (float)(max_range_cm + 1);
            row[i] = cm;
        }
    }
}

__global__ void ffma_peak_kernel(float *out, int iters) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const float b = 1.0000001f, c = 1e-7f;
    for (int i = 0; i < iters; i++) {
        a0 = fmaf(a0, b, c); a1 = fmaf(a1, b, c); a2 = fmaf(a2, b, c); a3 = fmaf(a3, b, c);
        a4 = fmaf(a4, b, c); a5 = fmaf(a5, b, c); a6 = fmaf(a6, b, c); a7 = fmaf(a7, b, c);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
__global__ void dfma_peak_kernel(double *out, int iters) {
    double a0 = threadIdx.x * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double b = 1.0000001, c = 1e-7;
    for (int i = 0; i < iters; i++) {
        a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
        a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

extern "C" zhsm_status zhsm_synth_pairs(zhsm_handle *h, float *d_scans, double *d_truth, long long first_pair, int npairs,
                                        float dmax_xy_m, float dmax_th_rad, float noise_m, float dropout,
                                        unsigned long long seed, void *cuda_stream) {
    if (!h || !d_scans || !d_truth || npairs < 0 || !h->have_laser) return ZHSM_ERR_ARG;
    if (npairs == 0) return ZHSM_OK;
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
    StreamScope scope(h, st);
    synth_pairs_kernel<<<npairs, 128, 0, st>>>(d_scans, d_truth, first_pair, npairs, h->laser.n, h->laser.fi_min,
                                               h->laser.dfi, h->laser.max_range_cm, dmax_xy_m, dmax_th_rad, noise_m,
                                               dropout, seed);
    ZHSM_CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    return ZHSM_OK;
}

// returns measured TFLOP/s (2 flops per FMA) of the CUDA-core pipe: which = 32 (FFMA) or 64 (DFMA)
extern "C" double zhsm_fp_peak_tflops(zhsm_handle *h, int which) {
    if (!h || (which != 32 && which != 64)) return -1;
    if (cudaSetDevice(h->device) != cudaSuccess) return -1;
    const int blocks = h->sm_count * 8, threads = 256, iters = which == 32 ? 1 << 16 : 1 << 14;
    void *out = nullptr;
    if (cudaMalloc(&out, (size_t)blocks * threads * 8) != cudaSuccess) return -1;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    double best = 0;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(e0, h->stream);
        if (which == 32) ffma_peak_kernel<<<blocks, threads, 0, h->stream>>>((float *)out, iters);
        else dfma_peak_kernel<<<blocks, threads, 0, h->stream>>>((double *)out, iters);
        cudaEventRecord(e1, h->stream);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        double tf = 2.0 * 8.0 * iters * (double)blocks * threads / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
        h->launches++;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(out);
    return best;
}
