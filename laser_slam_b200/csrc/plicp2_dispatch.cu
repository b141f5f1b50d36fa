// laser_slam_b200/csrc/plicp2_dispatch.cu -- host side of the PL-ICP kernel: CTA shape per beam count, the GF(2)
// matrices of the parallel ld_corr_hash, and the launch dispatch over the compiled instances.
#include <vector>

#include "zhsm_internal.h"

cudaError_t plicp_launch_t64(const PlicpArgs &a, cudaStream_t st);
cudaError_t plicp_launch_t128(const PlicpArgs &a, cudaStream_t st);
cudaError_t plicp_launch_t160(const PlicpArgs &a, cudaStream_t st);
cudaError_t plicp_launch_t192(const PlicpArgs &a, cudaStream_t st);
cudaError_t plicp_launch_t256(const PlicpArgs &a, cudaStream_t st);
cudaError_t plicp_launch_t512(const PlicpArgs &a, cudaStream_t st);
cudaError_t plicp_launch_t256_ray(const PlicpArgs &a, cudaStream_t st);
cudaError_t plicp_launch_final(const PlicpArgs &a, cudaStream_t st);
cudaError_t plicp_launch_ms128(const PlicpArgs &a, cudaStream_t st);

// CTA size per beam count.  What counts is jobs per SM (the shared-memory footprint, ~51 B per beam, decides: 4 at 1080
// beams) and then registers per thread, not warps per SM: a job has serial stretches that only other jobs on the same
// SM can fill, and threads beyond ~8 rays each mostly wait at the job's barriers.
//   n <= 1000: 128 threads, 128 registers, 4 CTAs/SM;
//   n <= 1472: 192 threads, 80 registers (4 CTAs/SM at 1080 beams, 3 up to 1472);
//   n <= 2304: 256 threads, 128 registers, 2 CTAs/SM (2048 beams), whose walks run as a pool and whose long walks
//   are finished warp-wide (plicp2_flat.cuh, plicp2_search.cuh);  else 512 threads (the same).
// knob_T (ZHSM_PLICP_T, read once in zhsm_create) overrides (tuning only).  Measured on one B200, ms per call
// (profiles/r03c_cta_shapes.txt), threads 64 / 128 / 192 / 256:
//   181 beams, 80k pairs: 21.0 / 24.7;  256, 80k: 16.8 / 19.9;  361, 60k: 16.1 / 17.6;  400, 40k: 11.7 / 13.0;
//   480, 40k: 14.1 / 14.2 / 15.3;  640, 40k: 20.2 / 17.8 / 19.3;  800, 40k: 26.4 / 22.8 / 23.6;  880, 20k: - / 12.4 / 12.8;
//   1000, 20k: - / 14.8 / 15.3;  1080, 20k: - / 16.33 / 16.34;  1200, 10k: - / - / 14.2 / 15.3;
//   1440, 10k: - / 20.7 / 19.6 / 19.5;  2048 beams, first guess 0, 8000 pairs: 256: 27.6, 512 (1 CTA/SM): slower.
// (Five 128-thread CTAs at 96 registers where the footprint allows, n <= 880: 3 % faster there, 5 % slower above.)
// The 64-thread instance (8 CTAs/SM) wins every UNIFORM batch up to 400 beams above by 9-19 % -- and loses on the real
// logs, whose jobs differ by an order of magnitude (sm_icp restarts on 10 % of intel-lab's pairs, 3.4-4 loops per pair
// on fr079 / mit-cscail): log x 16 in one batch, k pairs/s, threads 64 / 128 / 192: intel-lab (181 beams) 921 / 970 /
// 966, fr079 (361) 242 / 306 / 291, mit-cscail (361) 195 / 254 / 241; the log once: 380 / 454 / 472, 109 / 144 / 150,
// 58 / 85 / 86 (profiles/r03d_logs_by_threads.txt).  A long job wants its threads.  With the restarts as a second
// batch (zhsm_api.cu: plicp_fanout) the logs' jobs are one-loop jobs too, and the 64-thread shape is used for short scans
// in LARGE batches (plicp_wide_batch_T, zhsm_internal.h); the handle's shape stays 128 threads.
#ifndef PLICP_T64_MAX_BEAMS
#define PLICP_T64_MAX_BEAMS 0   // (off: see above)
#endif
void plicp_shape(int n, int knob_T, int *T, int *C, int *hash_chunk) {
    int t = n <= PLICP_T64_MAX_BEAMS ? 64 : n <= 1000 ? 128 : n <= 1472 ? 192 : n <= 2304 ? 256 : 512;
    if (knob_T == 64 || knob_T == 128 || knob_T == 160 || knob_T == 192 || knob_T == 256 || knob_T == 512) t = knob_T;
    while ((n + t - 1) / t > 32 && t < 512) t = t < 128 ? 128 : t < 256 ? 256 : 512;  // (a thread owns at most 32 rays)
    *T = t;
    *C = (n + t - 1) / t;
    int ch = (n + t - 1) / t;
    if (ch & 1) ch++;
    *hash_chunk = ch;
}

// ld_corr_hash steps without the data term (CSM laser_data.c): the linear parts over GF(2)
static inline uint32_t hash_lin_even(uint32_t h) { return h ^ (h << 7) ^ (h >> 3); }
static inline uint32_t hash_lin_odd(uint32_t h) { return h ^ (h << 11) ^ (h >> 5); }

// mat[t][b] = image of bit b under the linear part of steps i = min(n,(t+1)*ch) .. n-1 (identity when none remain)
void plicp_build_hash_matrices(int n, int T, int ch, std::vector<uint32_t> &mat) {
    mat.assign((size_t)T * 32, 0u);
    for (int t = 0; t < T; t++) {
        long long first = (long long)(t + 1) * ch;
        for (int b = 0; b < 32; b++) {
            uint32_t h = 1u << b;
            for (long long i = first; i < n; i++) h = (i & 1) ? hash_lin_odd(h) : hash_lin_even(h);
            mat[(size_t)t * 32 + b] = h;
        }
    }
}

size_t plicp_jump_bytes_per_job(int n);
cudaError_t plicp_jump_launch(const PlicpArgs &a, short *jt, int32_t *owner, cudaStream_t st, int *launched);

// a.jt must point at B * plicp_jump_bytes_per_job(n) bytes of scratch and a.owner at S ints: the owner / jump-table
// kernels fill them, the PL-ICP kernel reads them (same stream).  a.T / a.hash_chunk are the shape the handle's hash
// matrices were built for (zhsm_set_laser): the launch uses exactly that shape.
cudaError_t plicp_launch(const PlicpArgs &a, int sm_count, cudaStream_t st, int *launched) {
    (void)sm_count;
    const int T = a.T, C = (a.L.n + T - 1) / T;
    if ((T != 64 && T != 128 && T != 160 && T != 192 && T != 256 && T != 512) || C > 32 || (a.hash_chunk & 1) || a.hash_chunk < C)
        return cudaErrorInvalidValue;
    if (!a.jt || !a.owner) return cudaErrorInvalidValue;
    cudaError_t e = plicp_jump_launch(a, const_cast<short *>(a.jt), const_cast<int32_t *>(a.owner), st, launched);
    if (e != cudaSuccess) return e;
    if (a.ns == 4) {
        if (T != 128) return cudaErrorInvalidValue;
        e = plicp_launch_ms128(a, st);
        if (e == cudaSuccess && launched) *launched += 1;
        return e;
    }
    if (a.fin) e = plicp_launch_final(a, st);  // the final-only pass of plicp_fanout (plicp2_final.cu)
    else if (T == 256 && !a.pool) e = plicp_launch_t256_ray(a, st);  // tests: the 256-thread shape without the walk pool
    else
    e = T == 64 ? plicp_launch_t64(a, st) : T == 128 ? plicp_launch_t128(a, st) : T == 160 ? plicp_launch_t160(a, st) : T == 192 ? plicp_launch_t192(a, st) : T == 256 ? plicp_launch_t256(a, st) : plicp_launch_t512(a, st);
    if (e == cudaSuccess && launched) *launched += 1;
    return e;
}
