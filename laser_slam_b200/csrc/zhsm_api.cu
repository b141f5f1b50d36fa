// laser_slam_b200/csrc/zhsm_api.cu -- the C-ABI layer of libzhsm.so (include/zhsm.h) + small kernels.
// Host logic only: argument checks, staging of host buffers, table construction, launches.  No CPU fallback:
// every compute entry fails with ZHSM_ERR_CUDA when no usable device exists.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "zhsm_internal.h"

// ---------------------------------------------------------------------------------------------------------
// 5-tap in-place median filter: CCanonicalMatcher::median_filter (src/zhcsm/ZHCanonical_Matcher.cpp:159-189)
// == CPolarMatch::pm_median_filter (src/zhpsm/ZHPolar_Match.cpp:1610-1639).  The write-back makes taps i-2, i-1
// already-filtered values, i.e. a left-to-right recurrence along each scan: the parallelism is across scans.
// HBM-bound (8 B per beam).  A warp owns 32 scans and moves them through shared memory in tiles of 32 beams (+2 beams
// of look-ahead): every global access is a full coalesced row segment (lane = beam), while the recurrence runs with
// lane = scan on the transposed view (row pitch 35 floats: conflict-free).  The window lives in registers.
// ---------------------------------------------------------------------------------------------------------
constexpr int MF_W = 32, MF_PITCH = MF_W + 3, MF_WARPS = 4;

__device__ __forceinline__ float median5(float w0, float w1, float w2, float w3, float w4) {
    // bubble sort is a pure compare/swap sort: any exact median-of-5 gives the same value
    float a = w0, b = w1, c = w2, d = w3, t;
#define ZH_SW(x, y) if (x > y) { t = x; x = y; y = t; }
    ZH_SW(a, b) ZH_SW(c, d) ZH_SW(a, c) ZH_SW(b, d)   // a = min(a..d), d = max(a..d)
    ZH_SW(b, c)                                       // a <= b <= c <= d
#undef ZH_SW
    return w4 < b ? b : (w4 > c ? c : w4);            // median of {a,b,c,d,e} = median of {b, c, e}
}

__global__ void __launch_bounds__(32 * MF_WARPS) median_filter_kernel(float *scans, int S, int n) {
    __shared__ float tile_all[MF_WARPS][32 * MF_PITCH];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *tile = tile_all[warp];
    const long long base = ((long long)blockIdx.x * MF_WARPS + warp) * 32;  // first scan of this warp
    if (base >= S) return;
    const int ns = (int)(S - base < 32 ? S - base : 32);
    float m2 = 0, m1 = 0;  // filtered taps i-2, i-1 of scan `lane`
    for (int c0 = 0; c0 < n; c0 += MF_W) {
        // load: row segment [c0, c0 + 34) of each scan, indices clamped to n-1 (the reference clamps the taps)
        for (int sc = 0; sc < ns; sc++) {
            const float *row = scans + (size_t)(base + sc) * n;
            const int c = c0 + lane;
            tile[sc * MF_PITCH + lane] = row[c < n ? c : n - 1];
            if (lane < 2) { const int c2 = c0 + MF_W + lane; tile[sc * MF_PITCH + MF_W + lane] = row[c2 < n ? c2 : n - 1]; }
        }
        __syncwarp();
        if (lane < ns) {
            float *t = tile + lane * MF_PITCH;
            const int w = n - c0 < MF_W ? n - c0 : MF_W;
            if (c0 == 0) { m2 = t[0]; m1 = t[0]; }  // indices -2, -1 clamp to 0: the ORIGINAL r[0] until it is written
            for (int k = 0; k < w; k++) {
                const float med = median5(m2, m1, t[k], t[k + 1], t[k + 2]);
                t[k] = med;
                // taps for i+1 are f[i-1], f[i]; at i = 0 both clamped taps now see f[0]
                if (c0 + k == 0) { m2 = med; m1 = med; } else { m2 = m1; m1 = med; }
            }
        }
        __syncwarp();
        for (int sc = 0; sc < ns; sc++) {
            const int c = c0 + lane;
            if (c < n) scans[(size_t)(base + sc) * n + c] = tile[sc * MF_PITCH + lane];
        }
        __syncwarp();
    }
}

cudaError_t median_filter_launch(float *scans, int S, int n, cudaStream_t st) {
    const int per_block = 32 * MF_WARPS;
    median_filter_kernel<<<(S + per_block - 1) / per_block, 32 * MF_WARPS, 0, st>>>(scans, S, n);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------
// handle
// ---------------------------------------------------------------------------------------------------------
zhsm_status zhsm_stage(zhsm_handle *h, int slot, size_t bytes, void **out) {
    if (h->stage_bytes[slot] < bytes) {
        if (h->d_stage[slot]) cudaFree(h->d_stage[slot]);
        h->d_stage[slot] = nullptr;
        h->stage_bytes[slot] = 0;
        size_t cap = bytes + bytes / 4 + 256;
        if (cudaMalloc(&h->d_stage[slot], cap) != cudaSuccess) {
            cudaGetLastError();
            h->err = "cudaMalloc(staging) failed";
            return ZHSM_ERR_NOMEM;
        }
        h->stage_bytes[slot] = cap;
    }
    *out = h->d_stage[slot];
    return ZHSM_OK;
}

extern "C" zhsm_status zhsm_create(int device, zhsm_handle **out) {
    if (!out) return ZHSM_ERR_ARG;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0 || device < 0 || device >= count) {
        cudaGetLastError();
        return ZHSM_ERR_CUDA;  // no CPU fallback by design
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return ZHSM_ERR_CUDA;
    if (prop.major < 10) return ZHSM_ERR_UNSUPPORTED;  // sm_100a cubins only
    if (cudaSetDevice(device) != cudaSuccess) return ZHSM_ERR_CUDA;
    zhsm_handle *h = new zhsm_handle();
    h->device = device;
    h->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&h->ev0) != cudaSuccess || cudaEventCreate(&h->ev1) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->ev_busy, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->ev_stat, cudaEventDisableTiming) != cudaSuccess ||
        cudaMallocHost(&h->stat_pin, sizeof(int32_t)) != cudaSuccess || cudaMalloc(&h->d_stat, sizeof(int32_t)) != cudaSuccess) {
        zhsm_destroy(h);
        return ZHSM_ERR_CUDA;
    }
    // tuning knobs: read here once, clamped; the launch path never looks at the environment
    if (const char *e = getenv("ZHSM_PLICP_T")) { int v = atoi(e); if (v == 64 || v == 128 || v == 160 || v == 192 || v == 256 || v == 512) h->knob_T = v; }
    if (const char *e = getenv("ZHSM_PLICP_CHUNK")) { int v = atoi(e); if (v >= 1 && v <= (1 << 22)) h->knob_chunk = v; }
    if (const char *e = getenv("ZHSM_PLICP_CHUNK_ITERS")) { int v = atoi(e); if (v >= 0 && v <= ZHSM_MAX_ITER) h->knob_chunk_iters = v; }
    if (const char *e = getenv("ZHSM_PLICP_POOL")) h->knob_pool = atoi(e) != 0;
    if (const char *e = getenv("ZHSM_PLICP_FANOUT")) { int v = atoi(e); if (v >= 0) h->knob_fanout = v; }
    if (const char *e = getenv("ZHSM_PLICP_SMEM_PAD")) { int v = atoi(e); if (v >= 0 && v <= 160 * 1024) h->knob_smem_pad = (v & ~15) | (v & 1); }
    *out = h;
    return ZHSM_OK;
}

extern "C" void zhsm_destroy(zhsm_handle *h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (int i = 0; i < ZHSM_STAGE_SLOTS; i++) if (h->d_stage[i]) cudaFree(h->d_stage[i]);
    if (h->d_fi) cudaFree(h->d_fi);
    if (h->d_si) cudaFree(h->d_si);
    if (h->d_co) cudaFree(h->d_co);
    if (h->d_cs) cudaFree(h->d_cs);
    if (h->d_hashmat) cudaFree(h->d_hashmat);
    if (h->d_hashmat_ms) cudaFree(h->d_hashmat_ms);
    if (h->d_hashmat_64) cudaFree(h->d_hashmat_64);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->ev_busy) cudaEventDestroy(h->ev_busy);
    if (h->ev_stat) cudaEventDestroy(h->ev_stat);
    if (h->stat_pin) cudaFreeHost(h->stat_pin);
    if (h->d_stat) cudaFree(h->d_stat);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

extern "C" const char *zhsm_last_error(zhsm_handle *h) { return h ? h->err.c_str() : "null handle"; }

extern "C" uint64_t zhsm_launch_count(zhsm_handle *h) { return h ? h->launches : 0; }

extern "C" float zhsm_last_kernel_ms(zhsm_handle *h) {
    if (!h || !h->ev_valid) return -1.f;
    float ms = -1.f;
    if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) { cudaGetLastError(); return -1.f; }
    return ms;
}

extern "C" zhsm_status zhsm_sync(zhsm_handle *h, void *cuda_stream) {
    if (!h) return ZHSM_ERR_ARG;
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    return ZHSM_OK;
}

extern "C" zhsm_status zhsm_set_laser(zhsm_handle *h, const zhsm_laser *laser) {
    if (!h || !laser) return ZHSM_ERR_ARG;
    if (laser->n < 10 || laser->n > ZHSM_MAX_BEAMS) { h->err = "laser.n out of range [10,4096]"; return ZHSM_ERR_ARG; }
    if (h->have_laser && memcmp(&h->laser, laser, sizeof(*laser)) == 0) return ZHSM_OK;  // tables already built
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    // the tables are replaced: wait for everything that may still read the old ones (any stream this handle was used on)
    if (h->busy_valid) ZHSM_CUDA_TRY(h, cudaEventSynchronize(h->ev_busy));
    h->have_laser = false;
    const int n = laser->n;
    std::vector<double2> cs(n);
    for (int i = 0; i < n; i++) {
        // theta[i] exactly as fromPMScan2LDP (:210): float arithmetic, then widened; cos/sin by the host libm
        float thf = laser->fi_min + laser->dfi * (float)(unsigned)i;
        double th = (double)thf;
        cs[i].x = cos(th);
        cs[i].y = sin(th);
    }
    if (h->d_cs) { cudaFree(h->d_cs); h->d_cs = nullptr; }
    ZHSM_CUDA_TRY(h, cudaMalloc(&h->d_cs, sizeof(double2) * n));
    ZHSM_CUDA_TRY(h, cudaMemcpy(h->d_cs, cs.data(), sizeof(double2) * n, cudaMemcpyHostToDevice));
    // pm_init (ZHPolar_Match.cpp:68-78): pm_fi[i] = (float)i*dfi + fi_min, pm_si = sinf, pm_co = cosf by the host libm
    {
        std::vector<float> fi(n), si(n), co(n);
        for (int i = 0; i < n; i++) {
            fi[i] = ((float)i) * laser->dfi + laser->fi_min;
            si[i] = sinf(fi[i]);
            co[i] = cosf(fi[i]);
        }
        float **dst[3] = {&h->d_fi, &h->d_si, &h->d_co};
        const float *src[3] = {fi.data(), si.data(), co.data()};
        for (int k = 0; k < 3; k++) {
            if (*dst[k]) { cudaFree(*dst[k]); *dst[k] = nullptr; }
            ZHSM_CUDA_TRY(h, cudaMalloc(dst[k], sizeof(float) * n));
            ZHSM_CUDA_TRY(h, cudaMemcpy(*dst[k], src[k], sizeof(float) * n, cudaMemcpyHostToDevice));
        }
    }
    // GF(2) matrices of the parallel ld_corr_hash for this beam count / CTA shape
    {
        int T, C, ch;
        plicp_shape(n, h->knob_T, &T, &C, &ch);
        if (C > 32) { h->err = "no PL-ICP kernel instance for this beam count"; return ZHSM_ERR_UNSUPPORTED; }
        std::vector<uint32_t> mat;
        plicp_build_hash_matrices(n, T, ch, mat);
        if (h->d_hashmat) { cudaFree(h->d_hashmat); h->d_hashmat = nullptr; }
        ZHSM_CUDA_TRY(h, cudaMalloc(&h->d_hashmat, sizeof(uint32_t) * mat.size()));
        ZHSM_CUDA_TRY(h, cudaMemcpy(h->d_hashmat, mat.data(), sizeof(uint32_t) * mat.size(), cudaMemcpyHostToDevice));
        h->hash_chunk = ch;
        h->hash_T = T;  // the launch uses exactly the shape these matrices were built for
        h->hash_C = C;
        // ... and for the multi-start instance (128 threads per job)
        int ch_ms = (n + 127) / 128;
        if (ch_ms & 1) ch_ms++;
        plicp_build_hash_matrices(n, 128, ch_ms, mat);
        if (h->d_hashmat_ms) { cudaFree(h->d_hashmat_ms); h->d_hashmat_ms = nullptr; }
        ZHSM_CUDA_TRY(h, cudaMalloc(&h->d_hashmat_ms, sizeof(uint32_t) * mat.size()));
        ZHSM_CUDA_TRY(h, cudaMemcpy(h->d_hashmat_ms, mat.data(), sizeof(uint32_t) * mat.size(), cudaMemcpyHostToDevice));
        h->hash_chunk_ms = ch_ms;
        // ... and for the 64-thread shape of short scans (plicp_wide_batch_T)
        if (h->d_hashmat_64) { cudaFree(h->d_hashmat_64); h->d_hashmat_64 = nullptr; }
        if (n <= 256) {
            int ch_64 = (n + 63) / 64;
            if (ch_64 & 1) ch_64++;
            plicp_build_hash_matrices(n, 64, ch_64, mat);
            ZHSM_CUDA_TRY(h, cudaMalloc(&h->d_hashmat_64, sizeof(uint32_t) * mat.size()));
            ZHSM_CUDA_TRY(h, cudaMemcpy(h->d_hashmat_64, mat.data(), sizeof(uint32_t) * mat.size(), cudaMemcpyHostToDevice));
            h->hash_chunk_64 = ch_64;
        }
    }
    h->laser = *laser;
    h->have_laser = true;
    h->restart_frac = -1.f;  // another sensor, another workload: what the previous batches did says nothing (plicp_use_fanout)
    h->stat_pending = false;
    return ZHSM_OK;
}

// Host emulation of the kernel's chunked ld_corr_hash (test hook: the same matrices, chunking and XOR combine as
// plicp2_select.cuh::corr_hash_chunk, no device needed).  seq = the sequential CSM hash, par = the chunked one.
extern "C" int zhsm_debug_corr_hash(const int32_t *stream, int n, uint32_t *seq, uint32_t *par) {
    if (!stream || n < 1 || n > ZHSM_MAX_BEAMS || !seq || !par) return 1;
    int T, C, ch;
    plicp_shape(n, 0, &T, &C, &ch);
    if (C > 32) return 2;
    std::vector<uint32_t> mat;
    plicp_build_hash_matrices(n, T, ch, mat);
    uint32_t h = 0;
    for (int i = 0; i < n; i++) {
        uint32_t v = (uint32_t)stream[i];
        h ^= ((i & 1) == 0) ? ((h << 7) ^ v ^ (h >> 3)) : (~((h << 11) ^ v ^ (h >> 5)));
    }
    *seq = h & 0x7FFFFFFFu;
    uint32_t acc = 0;
    for (int t = 0; t < T; t++) {
        int lo = t * ch, hi = lo + ch < n ? lo + ch : n;
        if (lo >= n) break;
        uint32_t c = 0;
        for (int i = lo; i < hi; i++) {
            uint32_t v = (uint32_t)stream[i];
            c ^= ((i & 1) == 0) ? ((c << 7) ^ v ^ (c >> 3)) : (~((c << 11) ^ v ^ (c >> 5)));
        }
        uint32_t out = 0;
        for (int b = 0; b < 32; b++) if (c >> b & 1u) out ^= mat[(size_t)t * 32 + b];
        acc ^= out;
    }
    *par = acc & 0x7FFFFFFFu;
    return 0;
}

// CCanonicalMatcher::sm_init (src/zhcsm/ZHCanonical_Matcher.cpp:83-157)
extern "C" void zhsm_plicp_default_params(zhsm_plicp_params *p, int pm_min_range_cm, int pm_max_range_cm) {
    memset(p, 0, sizeof(*p));
    p->min_reading = pm_min_range_cm / 100.;
    p->max_reading = pm_max_range_cm / 100.;
    p->max_angular_correction_deg = 45;
    p->max_linear_correction = 0.5;
    p->max_iterations = 10;
    p->epsilon_xy = 0.001;
    p->epsilon_theta = 0.001;
    p->max_correspondence_dist = 2;
    p->sigma = 0.04;
    p->use_corr_tricks = 1;
    p->restart = 1;
    p->restart_threshold_mean_error = 0.01;
    p->restart_dt = 0.01;
    p->restart_dtheta = 0.0261799;
    p->use_point_to_line_distance = 1;
    p->outliers_maxPerc = 0.95;
    p->outliers_adaptive_order = 0.7;
    p->outliers_adaptive_mult = 2.0;
    p->outliers_remove_doubles = 1;
    p->do_compute_covariance = 0;
}

extern "C" zhsm_status zhsm_median_filter_batch(zhsm_handle *h, float *scans, int S, unsigned flags, void *cuda_stream) {
    if (!h || !scans || S < 0) return ZHSM_ERR_ARG;
    if (!h->have_laser) { h->err = "zhsm_set_laser not called"; return ZHSM_ERR_ARG; }
    if (S == 0) return ZHSM_OK;
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
    StreamScope scope(h, st);
    const int n = h->laser.n;
    float *d = scans;
    size_t bytes = sizeof(float) * (size_t)S * n;
    if (!(flags & ZHSM_SCANS_ON_DEVICE)) {
        void *p;
        zhsm_status s = zhsm_stage(h, 0, bytes, &p);
        if (s != ZHSM_OK) return s;
        d = (float *)p;
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(d, scans, bytes, cudaMemcpyHostToDevice, st));
    }
    ZHSM_CUDA_TRY(h, median_filter_launch(d, S, n, st));
    h->launches++;
    if (!(flags & ZHSM_SCANS_ON_DEVICE)) {
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(scans, d, bytes, cudaMemcpyDeviceToHost, st));
        ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    }
    return ZHSM_OK;
}

// ---------------------------------------------------------------------------------------------------------
// sm_icp's restart loop as a second batch (plicp_fanout)
// ---------------------------------------------------------------------------------------------------------
// sm_icp (CSM icp.c) restarts a match whose mean error is above restart_threshold_mean_error from six
// perturbed poses, one icp_loop after the other, and keeps the best.  Each of those loops depends on its start pose
// only, so a SMALL batch -- a log run once, the batch of one behind the CSM-shaped sm_icp() -- need not wait for its
// few seven-loop jobs: pass A runs every job with restarts off (one icp_loop each), the jobs that sm_icp would restart
// are listed, pass B runs their 6 R perturbed loops as ordinary one-loop jobs, a small kernel does sm_icp's
// bookkeeping (the first failing restart ends the loop, strict < keeps the earliest best, iterations add up), and
// pass C -- the PL-ICP kernel in its final-only mode (PlicpArgs::fin) -- recomputes the correspondences at the best
// pose and writes the record.  Same records as the in-job restart loop (tests/test_plicp_gpu.py compares the bytes).
// Large batches keep the in-job loop: there the long jobs are a matter of throughput, which is the same either way.
// (also takes the listed jobs' scan indices along: after the host has read the count, nothing reads the caller's job
// arrays any more -- with ZHSM_JOBS_ON_DEVICE they are only promised to live as long as an ordinary pass needs them)
static __global__ void restart_select_kernel(const zhsm_result *res, const int32_t *ref_idx, const int32_t *sens_idx, int B,
                                             double threshold, int32_t *list, int32_t *ref_f, int32_t *sens_f, int32_t *count) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= B) return;
    const zhsm_result r = res[j];
    // icp.c: restart && (best_error / nvalid) > restart_threshold_mean_error, after a valid first loop
    if (r.valid == 1 && (r.error / r.nvalid) > threshold) {
        const int slot = atomicAdd(count, 1);
        list[slot] = j;
        ref_f[slot] = ref_idx[j];
        sens_f[slot] = sens_idx[j];
    }
}

static __global__ void restart_jobs_kernel(const zhsm_result *res, const int32_t *ref_f, const int32_t *sens_f,
                                           const int32_t *list, int R, double dt, double dth, int32_t *ref_b,
                                           int32_t *sens_b, double *guess_b) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= 6 * R) return;
    const int j = list[s / 6], k = s % 6;
    ref_b[s] = ref_f[s / 6];
    sens_b[s] = sens_f[s / 6];
    // icp.c's perturbations in its order: +-dt in x, +-dt in y, +-dtheta
    const double pert[6][3] = {{dt, 0, 0}, {-dt, 0, 0}, {0, dt, 0}, {0, -dt, 0}, {0, 0, dth}, {0, 0, -dth}};
    for (int q = 0; q < 3; q++) guess_b[3 * (size_t)s + q] = res[j].x[q] + pert[k][q];
}

static __global__ void restart_combine_kernel(const zhsm_result *res, const zhsm_result *res_b, const int32_t *list, int R,
                                              PlicpFin *fin) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const int j = list[r];
    const zhsm_result a = res[j];
    PlicpFin f;
    for (int q = 0; q < 3; q++) f.x[q] = a.x[q];
    f.error = a.error;
    f.nvalid = a.nvalid;
    f.iterations = a.iterations;
    f.n_loops = 1;
    for (int q = 0; q < 4; q++) f.diag[q] = a.reserved[q];
    f.pad_ = 0;
    for (int k = 0; k < 6; k++) {
        const zhsm_result b = res_b[6 * (size_t)r + k];
        f.n_loops++;
        for (int q = 0; q < 4; q++) f.diag[q] += b.reserved[q];
        if (b.valid != 1) break;  // a failing restart ends the restart loop
        f.iterations += b.iterations;
        if (b.error < f.error) {
            f.error = b.error;
            for (int q = 0; q < 3; q++) f.x[q] = b.x[q];
        }
    }
    fin[r] = f;
}

#ifndef ZHSM_FANOUT_MIN_FRAC
#define ZHSM_FANOUT_MIN_FRAC 0.02f   // restarting share of the previous batch from which a larger batch fans out (tuning builds)
#endif
static __global__ void restart_count_kernel(const zhsm_result *res, int B, int32_t *count) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const bool r = j < B && res[j].n_loops > 1;
    const unsigned m = __ballot_sync(0xffffffffu, r);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(count, __popc(m));
}

// Fan out or restart in the job?  The fan-out costs two more passes (each ends with the tail of its last jobs) and a
// round trip to the host: a few per cent of a mid-size batch that hardly restarts (the synthetic pairs: 1 job in
// 20 000; 16.4 -> 17.0 ms per 20k pairs, and 7 % of the two-handle end-to-end rate, whose chunks overlap) and 1.6x the
// rate of one that does (the logs: 10-50 % of the pairs).  So: batches of 512 to 4096 jobs always (a log run once: the
// seven-loop jobs are its tail), larger ones by what the handle's previous batch did, and less than a wave of jobs
// never: there every job runs at once anyway and three passes cost more launch latency than six loops take (a batch
// of one behind sm_icp(), intel-lab, host buffers: 187 us without a restart, 947 us with one in the job, 1122 us as
// three passes; tools/gpu_probe_latency.py).
static bool plicp_use_fanout(zhsm_handle *h, const zhsm_plicp_params &P, int B, int ns) {
    if (h->stat_pending && cudaEventQuery(h->ev_stat) == cudaSuccess) {
        h->restart_frac = h->stat_B > 0 ? (float)*h->stat_pin / (float)h->stat_B : 0.f;
        h->stat_pending = false;
    } else if (h->stat_pending) cudaGetLastError();  // (cudaErrorNotReady is not an error)
    if (!P.restart || ns == 4 || B > h->knob_fanout) return false;
    if (B < 512) return false;  // less than a wave of jobs: the passes' launch latencies outweigh six loops in a row
    return B <= 4096 || h->restart_frac < 0.f || h->restart_frac >= ZHSM_FANOUT_MIN_FRAC;
}

// after an in-job pass: count the restarted jobs, the number travels back behind the stream
static void plicp_issue_restart_count(zhsm_handle *h, const zhsm_result *d_results, int B, cudaStream_t st) {
    if (h->stat_pending) return;
    if (cudaMemsetAsync(h->d_stat, 0, sizeof(int32_t), st) != cudaSuccess) { cudaGetLastError(); return; }
    restart_count_kernel<<<(B + 255) / 256, 256, 0, st>>>(d_results, B, h->d_stat);
    if (cudaMemcpyAsync(h->stat_pin, h->d_stat, sizeof(int32_t), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaEventRecord(h->ev_stat, st) != cudaSuccess) { cudaGetLastError(); return; }
    h->stat_B = B;
    h->stat_pending = true;
}

// one pass of the PL-ICP kernel over `count` jobs in launches of at most `chunk`
static cudaError_t plicp_pass(zhsm_handle *h, const PlicpArgs &a, int count, int chunk, int n, cudaStream_t st, int *launched) {
    for (int off = 0; off < count; off += chunk) {
        PlicpArgs c = a;
        c.B = count - off < chunk ? count - off : chunk;
        c.ref_idx = a.ref_idx + off;
        c.sens_idx = a.sens_idx + off;
        if (a.first_guess) c.first_guess = a.first_guess + 3 * (size_t)off;
        if (a.fin) c.fin = a.fin + off;
        if (a.out_idx) c.out_idx = a.out_idx + off;  // (records and correspondences are addressed through it)
        else {
            c.results = a.results + off;
            if (a.corr_j1) { c.corr_j1 = a.corr_j1 + (size_t)off * n; c.corr_j2 = a.corr_j2 + (size_t)off * n; }
        }
        cudaError_t e = plicp_launch(c, h->sm_count, st, launched);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// `a` is ready for an ordinary pass over a.B jobs (everything on the device).  Blocks until pass A is done (the number
// of restarting jobs is needed on the host).
static zhsm_status plicp_fanout(zhsm_handle *h, const PlicpArgs &a, int chunk, cudaStream_t st, int *launched) {
    const int B = a.B, n = a.L.n;
    void *p;
    zhsm_status s;
    if ((s = zhsm_stage(h, 24, sizeof(int32_t) * (3 * (size_t)B + 4), &p)) != ZHSM_OK) return s;
    int32_t *count = (int32_t *)p, *list = count + 4, *ref_f = list + B, *sens_f = ref_f + B;
    ZHSM_CUDA_TRY(h, cudaMemsetAsync(count, 0, sizeof(int32_t), st));
    PlicpArgs A = a;
    A.P.restart = 0;
    ZHSM_CUDA_TRY(h, plicp_pass(h, A, B, chunk, n, st, launched));
    restart_select_kernel<<<(B + 255) / 256, 256, 0, st>>>(a.results, a.ref_idx, a.sens_idx, B, a.P.restart_threshold_mean_error, list,
                                                          ref_f, sens_f, count);
    ZHSM_CUDA_TRY(h, cudaGetLastError());
    int32_t R = 0;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(&R, count, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    *launched += 1;
    h->restart_frac = (float)R / (float)B;
    if (R <= 0) return ZHSM_OK;
    const size_t R6 = 6 * (size_t)R;
    if ((s = zhsm_stage(h, 25, (sizeof(int32_t) * 2 + sizeof(double) * 3) * R6 + 64, &p)) != ZHSM_OK) return s;
    double *guess_b = (double *)p;
    int32_t *ref_b = (int32_t *)(guess_b + 3 * R6), *sens_b = ref_b + R6;
    if ((s = zhsm_stage(h, 26, sizeof(zhsm_result) * R6, &p)) != ZHSM_OK) return s;
    zhsm_result *res_b = (zhsm_result *)p;
    if ((s = zhsm_stage(h, 27, sizeof(PlicpFin) * (size_t)R, &p)) != ZHSM_OK) return s;
    PlicpFin *fin = (PlicpFin *)p;
    restart_jobs_kernel<<<(int)((R6 + 255) / 256), 256, 0, st>>>(a.results, ref_f, sens_f, list, R, a.P.restart_dt,
                                                                a.P.restart_dtheta, ref_b, sens_b, guess_b);
    ZHSM_CUDA_TRY(h, cudaGetLastError());
    PlicpArgs Bq = a;
    Bq.P.restart = 0;
    Bq.P.do_compute_covariance = 0;
    Bq.ref_idx = ref_b; Bq.sens_idx = sens_b; Bq.first_guess = guess_b;
    Bq.results = res_b; Bq.corr_j1 = nullptr; Bq.corr_j2 = nullptr;
    ZHSM_CUDA_TRY(h, plicp_pass(h, Bq, (int)R6, chunk, n, st, launched));
    restart_combine_kernel<<<(R + 255) / 256, 256, 0, st>>>(a.results, res_b, list, R, fin);
    ZHSM_CUDA_TRY(h, cudaGetLastError());
    PlicpArgs Cq = a;
    Cq.ref_idx = ref_f; Cq.sens_idx = sens_f; Cq.first_guess = nullptr;
    Cq.fin = fin; Cq.out_idx = list;
    ZHSM_CUDA_TRY(h, plicp_pass(h, Cq, R, chunk, n, st, launched));
    *launched += 2;
    return ZHSM_OK;
}

// ---------------------------------------------------------------------------------------------------------
// batched PL-ICP
// ---------------------------------------------------------------------------------------------------------
extern "C" zhsm_status zhsm_plicp_batch(zhsm_handle *h, const float *scans, int S, const int32_t *ref_idx,
                                        const int32_t *sens_idx, const double *first_guess, int B,
                                        const zhsm_plicp_params *params, zhsm_result *results, int16_t *corr_j1,
                                        int16_t *corr_j2, unsigned flags, void *cuda_stream) {
    if (!h) return ZHSM_ERR_ARG;
    if (!scans || !ref_idx || !sens_idx || !results || S <= 0 || B < 0 || ((corr_j1 == nullptr) != (corr_j2 == nullptr))) {
        h->err = "zhsm_plicp_batch: bad argument";
        return ZHSM_ERR_ARG;
    }
    if (!h->have_laser) { h->err = "zhsm_set_laser not called"; return ZHSM_ERR_ARG; }
    if (B == 0) return ZHSM_OK;
    zhsm_plicp_params P;
    if (params) P = *params; else zhsm_plicp_default_params(&P, h->laser.min_range_cm, h->laser.max_range_cm);
    if (P.use_corr_tricks != 1 || P.use_point_to_line_distance != 1) {
        h->err = "only use_corr_tricks=1 / use_point_to_line_distance=1 (the reference's settings) are implemented";
        return ZHSM_ERR_UNSUPPORTED;
    }
    if (P.max_iterations < 1 || P.max_iterations > ZHSM_MAX_ITER) { h->err = "max_iterations out of [1,64]"; return ZHSM_ERR_ARG; }
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
    StreamScope scope(h, st);
    const int n = h->laser.n;

    PlicpArgs a;
    memset(&a, 0, sizeof(a));
    a.B = B;
    a.S = S;
    a.T = h->hash_T;
    a.smem_pad = h->knob_smem_pad;
    a.P = P;
    a.L.n = n; a.L.fi_min = h->laser.fi_min; a.L.dfi = h->laser.dfi;
    a.L.min_range_cm = h->laser.min_range_cm; a.L.max_range_cm = h->laser.max_range_cm;
    a.L.min_valid_points = h->laser.min_valid_points; a.L.scan_window = h->laser.scan_window;
    a.L.flags = h->laser.flags;
    a.cs = h->d_cs;
    a.hashmat = h->d_hashmat;
    a.hash_chunk = h->hash_chunk;
    a.chunk_iters = plicp_chunk_iters(h->knob_chunk_iters, h->hash_T);
    a.pool = h->knob_pool;
    if (plicp_wide_batch_T(n, B, h->knob_T) == 64 && h->d_hashmat_64) {
        a.T = 64;
        a.hashmat = h->d_hashmat_64;
        a.hash_chunk = h->hash_chunk_64;
        a.chunk_iters = plicp_chunk_iters(h->knob_chunk_iters, 64);
    }
    // ZHSM_MULTISTART: the caller promises jobs in runs of four that share both scans (pair-major multi-start).  Used
    // where it pays: long scans, whose single-job CTAs fit only twice per SM.  The kernel re-checks the promise per job.
    if ((flags & ZHSM_MULTISTART) && B % 4 == 0 && n > 1472 && plicp_ms128_smem_bytes(n) <= 227 * 1024 && (n + 127) / 128 <= 32 &&
        h->knob_T == 0) {
        a.ns = 4;
        a.T = 128;
        a.hashmat = h->d_hashmat_ms;
        a.hash_chunk = h->hash_chunk_ms;
    }

    // ld_valid_fields' per-scan-geometry checks (laser_data.c): nrays, fov, theta[] consistent by construction
    {
        float thf0 = h->laser.fi_min, thf1 = h->laser.fi_min + h->laser.dfi * (float)(unsigned)(n - 1);
        double fov = (double)thf1 - (double)thf0;
        const double PI = 3.14159265358979323846;
        if (n < 10 || n > 10000 || !(fov >= 20.0 * PI / 180.0 && fov <= 2.01 * PI)) {
            h->err = "geometry rejected by ld_valid_fields (fov must be in [20deg, 2.01*pi], 10<=n<=10000)";
            return ZHSM_ERR_ARG;
        }
    }
    void *p;
    zhsm_status s;
    const size_t scan_bytes = sizeof(float) * (size_t)S * n;
    if (flags & ZHSM_SCANS_ON_DEVICE) a.scans = scans;
    else {
        if ((s = zhsm_stage(h, 0, scan_bytes, &p)) != ZHSM_OK) return s;
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(p, scans, scan_bytes, cudaMemcpyHostToDevice, st));
        a.scans = (const float *)p;
    }
    if (flags & ZHSM_JOBS_ON_DEVICE) { a.ref_idx = ref_idx; a.sens_idx = sens_idx; a.first_guess = first_guess; }
    else {
        for (int k = 0; k < B; k++)
            if (ref_idx[k] < 0 || ref_idx[k] >= S || sens_idx[k] < 0 || sens_idx[k] >= S) {
                h->err = "zhsm_plicp_batch: scan index out of range";
                return ZHSM_ERR_ARG;
            }
        if ((s = zhsm_stage(h, 1, sizeof(int32_t) * 2 * (size_t)B, &p)) != ZHSM_OK) return s;
        int32_t *di = (int32_t *)p;
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(di, ref_idx, sizeof(int32_t) * B, cudaMemcpyHostToDevice, st));
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(di + B, sens_idx, sizeof(int32_t) * B, cudaMemcpyHostToDevice, st));
        a.ref_idx = di; a.sens_idx = di + B;
        if (first_guess) {
            if ((s = zhsm_stage(h, 2, sizeof(double) * 3 * (size_t)B, &p)) != ZHSM_OK) return s;
            ZHSM_CUDA_TRY(h, cudaMemcpyAsync(p, first_guess, sizeof(double) * 3 * B, cudaMemcpyHostToDevice, st));
            a.first_guess = (const double *)p;
        }
    }
    const size_t corr_bytes = sizeof(int16_t) * (size_t)B * n;
    if (flags & ZHSM_RESULTS_ON_DEVICE) { a.results = results; a.corr_j1 = corr_j1; a.corr_j2 = corr_j2; }
    else {
        if ((s = zhsm_stage(h, 3, sizeof(zhsm_result) * (size_t)B, &p)) != ZHSM_OK) return s;
        a.results = (zhsm_result *)p;
        if (corr_j1) {
            if ((s = zhsm_stage(h, 4, 2 * corr_bytes, &p)) != ZHSM_OK) return s;
            a.corr_j1 = (int16_t *)p;
            a.corr_j2 = a.corr_j1 + (size_t)B * n;
        }
    }
    if (flags & ZHSM_FILTER_SCANS) {
        ZHSM_CUDA_TRY(h, median_filter_launch(const_cast<float *>(a.scans), S, n, st));
        h->launches++;
        if ((flags & ZHSM_WRITEBACK_SCANS) && !(flags & ZHSM_SCANS_ON_DEVICE))
            ZHSM_CUDA_TRY(h, cudaMemcpyAsync(const_cast<float *>(scans), a.scans, scan_bytes, cudaMemcpyDeviceToHost, st));
    }
    // Jump tables of the jobs' ref scans go through a scratch buffer (8 bytes per ray and job); large batches are
    // cut into chunks so that the scratch stays bounded (ZHSM_PLICP_CHUNK jobs, default 131072 = 1.1 GB at 1080 beams).
    int chunk = h->knob_chunk;
    if (chunk > B) chunk = B;
    if ((s = zhsm_stage(h, 15, plicp_jump_bytes_per_job(n) * (size_t)chunk, &p)) != ZHSM_OK) return s;
    a.jt = (const short *)p;
    // per scan: the job of the running launch that owns (builds) its jump tables -- jobs sharing a ref scan share them
    if ((s = zhsm_stage(h, 16, sizeof(int32_t) * (size_t)S, &p)) != ZHSM_OK) return s;
    a.owner = (const int32_t *)p;
    ZHSM_CUDA_TRY(h, cudaEventRecord(h->ev0, st));
    int launched = 0;
    if (plicp_use_fanout(h, P, B, a.ns)) {
        if ((s = plicp_fanout(h, a, chunk, st, &launched)) != ZHSM_OK) return s;
    } else {
        ZHSM_CUDA_TRY(h, plicp_pass(h, a, B, chunk, n, st, &launched));
        if (P.restart && B >= 512) { plicp_issue_restart_count(h, a.results, B, st); launched++; }  // (smaller batches never fan out)
    }
    h->launches += launched;
    ZHSM_CUDA_TRY(h, cudaEventRecord(h->ev1, st));
    h->ev_valid = true;
    if (!(flags & ZHSM_RESULTS_ON_DEVICE)) {
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(results, a.results, sizeof(zhsm_result) * (size_t)B, cudaMemcpyDeviceToHost, st));
        if (corr_j1) {
            ZHSM_CUDA_TRY(h, cudaMemcpyAsync(corr_j1, a.corr_j1, corr_bytes, cudaMemcpyDeviceToHost, st));
            ZHSM_CUDA_TRY(h, cudaMemcpyAsync(corr_j2, a.corr_j2, corr_bytes, cudaMemcpyDeviceToHost, st));
        }
        ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));  // host results are ready on return
    }
    return ZHSM_OK;
}
