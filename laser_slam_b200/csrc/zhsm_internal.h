// laser_slam_b200/csrc/zhsm_internal.h -- internals shared by the kernels and the C-ABI layer.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/zhsm.h"

#define ZHSM_MAX_BEAMS 4096
#define ZHSM_MAX_ITER 64
#define ZHSM_STAGE_SLOTS 28

struct zhsm_handle {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    uint64_t launches = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool ev_valid = false;
    // The staging slots and scratch buffers belong to the handle, not to a call: every entry point orders its stream
    // behind the last call that used them (StreamScope below), so calls on different streams never overlap on them.
    cudaEvent_t ev_busy = nullptr;
    bool busy_valid = false;
    // tuning knobs, read from the environment ONCE in zhsm_create (never on the launch path) and clamped
    int knob_T = 0;            // ZHSM_PLICP_T: 0 = by beam count, else 128 / 192 / 256 / 512
    int knob_chunk = 131072;   // ZHSM_PLICP_CHUNK: jobs per launch (bounds the jump-table scratch)
    int knob_chunk_iters = -1; // ZHSM_PLICP_CHUNK_ITERS: first iterations searched in chunk order; -1: by CTA shape (plicp_chunk_iters)
    int knob_smem_pad = 0;     // ZHSM_PLICP_SMEM_PAD: extra dynamic shared memory (fewer CTAs per SM)
    int knob_fanout = 1 << 30; // ZHSM_PLICP_FANOUT: batches of up to this many jobs run sm_icp's restarts as a second batch (0: never)
    // How often the last PL-ICP batch restarted (jobs with n_loops > 1 / jobs), -1: not known yet.  The fan-out measures
    // it on the host; the in-job pass counts on the device and the number comes back asynchronously (pinned stat_pin,
    // ev_stat) and is picked up by the next call.  See plicp_use_fanout (zhsm_api.cu).
    float restart_frac = -1.f;
    int32_t *stat_pin = nullptr;   // pinned: restarted jobs of the batch that issued the count
    int32_t *d_stat = nullptr;     // its device counter
    int stat_B = 0;                // that batch's size
    cudaEvent_t ev_stat = nullptr;
    bool stat_pending = false;
    int knob_pool = 1;         // ZHSM_PLICP_POOL=0: the 256-thread shape runs its instance without the walk pool (plicp2_t256r.cu; tests)
    int hash_T = 0, hash_C = 0;  // CTA shape the hash matrices were built for (zhsm_set_laser)
    uint32_t *d_hashmat_ms = nullptr;  // the same for the multi-start instance's shape (128 threads per job)
    int hash_chunk_ms = 0;
    uint32_t *d_hashmat_64 = nullptr;  // ... and for the 64-thread shape (short scans in large batches: plicp_wide_batch_T)
    int hash_chunk_64 = 0;

    bool have_laser = false;
    zhsm_laser laser{};
    double2 *d_cs = nullptr;      // [n] (cos theta_i, sin theta_i), host libm, double
    uint32_t *d_hashmat = nullptr; // GF(2) matrices for the parallel ld_corr_hash, see plicp_sm100.cu
    int hash_chunk = 0, hash_chunks = 0;

    float *d_fi = nullptr, *d_si = nullptr, *d_co = nullptr;  // pm_fi / pm_si / pm_co (pm_init), host libm, float

    // staging for host-pointer calls
    void *d_stage[ZHSM_STAGE_SLOTS] = {nullptr};
    size_t stage_bytes[ZHSM_STAGE_SLOTS] = {0};
};

// Orders `st` behind the handle's previous call (any stream) on entry and marks the handle busy until everything this
// call queued on `st` is done.  One per entry point that touches staging / scratch buffers.
struct StreamScope {
    zhsm_handle *h;
    cudaStream_t st;
    StreamScope(zhsm_handle *h_, cudaStream_t st_) : h(h_), st(st_) {
        if (h->busy_valid) cudaStreamWaitEvent(st, h->ev_busy, 0);
    }
    ~StreamScope() {
        if (cudaEventRecord(h->ev_busy, st) == cudaSuccess) h->busy_valid = true;
        else cudaGetLastError();
    }
};

// device-side geometry passed by value to kernels
struct LaserDev {
    int n;
    float fi_min, dfi;
    int min_range_cm, max_range_cm;
    int min_valid_points, scan_window;
    int flags;  // ZHSM_LASER_*
};

// sm_icp's restart loop as a second batch (zhsm_api.cu: plicp_fanout): what the LAST pass of a restarted job starts from
struct PlicpFin {
    double x[3];        // best pose over loop 0 and the restarts
    double error;       // its total error
    int32_t nvalid;     // loop 0's correspondences (sm_result.nvalid)
    int32_t iterations; // summed over the loops that count
    int32_t n_loops;    // icp_loop executions
    int32_t diag[4];    // walks, first-pass walks, verify rounds, walk steps of the passes before (zhsm_result.reserved)
    int32_t pad_;
};

struct PlicpArgs {
    const float *scans;      // [S][n]
    const int32_t *ref_idx;  // [B]
    const int32_t *sens_idx; // [B]
    const double *first_guess; // [B][3] or null
    zhsm_result *results;    // [B]
    int16_t *corr_j1, *corr_j2; // [B][n] or null
    const double2 *cs;       // [n]
    const short *jt;         // [B][2][n][2] jump tables (plicp2_jump.cu): slot k holds the tables of job k's ref scan
                             // when job k is the OWNER of that scan (lowest job of the launch that uses it)
    const int32_t *owner;    // [S] per scan: the owning job of this launch (jobs sharing a ref scan share its tables)
    int S;                   // scans in the pool (device-side range check of the job indices)
    int ns;                  // 4: the multi-start instance (four jobs per CTA that share both scans, 128 threads each); else 1
    int T;                   // CTA shape the handle's hash matrices were built for
    int smem_pad;            // tuning: extra dynamic shared memory per CTA
    const uint32_t *hashmat; // [T][32] GF(2) matrices of the parallel ld_corr_hash (plicp2_select.cuh)
    int hash_chunk;          // stream elements hashed per thread (even)
    int chunk_iters;         // first iterations of an icp_loop searched in chunk order (find_correspondences)
    int pool;                // 0: the 256-thread shape is launched from plicp2_t256r.cu (no walk pool); tests only
    const PlicpFin *fin;     // not null: the jobs skip the icp_loops and run the final stage only (correspondences at the
                             // given pose, hash, covariance, record) with this bookkeeping -- the last pass of plicp_fanout
    const int32_t *out_idx;  // not null: job k writes record (and correspondences) out_idx[k] instead of k
    int B;
    LaserDev L;
    zhsm_plicp_params P;
};

struct PsmArgs {
    const float *scans;          // [S][n] preprocessed ranges (cm)
    const unsigned char *bad;    // [S][n]
    const short *seg;            // [S][n]
    const int32_t *ref_idx, *act_idx;  // [B]
    const float *init_pose;      // [B][3] act pose (cm, cm, rad) or null
    zhsm_result *results;        // [B]
    float *proj_r;               // scan_project mode: [B][n]
    unsigned char *proj_bad;     // scan_project mode: [B][n]
    const float *fi, *si, *co;   // [n] host-libm tables (pm_init, ZHPolar_Match.cpp:68-78)
    int B, algo;
    int S;                       // scans in the pool (device-side range check of the job indices)
    LaserDev L;
};
cudaError_t psm_launch(const PsmArgs &a, int mode, cudaStream_t st);  // mode 0 matcher, 1 scan_project
cudaError_t far_segment_launch(const float *scans, unsigned char *bad, short *seg, int S, int n, int max_range_cm,
                               cudaStream_t st);
struct IcpArgs {
    const float *pts;            // cloud pool, x-block then y-block per cloud
    const int32_t *off, *cnt;    // [C] device copies of the pool's table of contents
    const int32_t *ref_idx, *cur_idx;  // [B]
    const double *init;          // [B][3] or null
    zhsm_result *results;        // [B]
    int C;                       // clouds in the pool (device-side range check of the job indices)
    int B, cap;                  // cap = capacity (points) of the shared-memory arrays = largest cloud
    zhsm_icp_options opt;
};
cudaError_t icp_launch(const IcpArgs &a, cudaStream_t st);
size_t icp_smem_bytes(int cap);
cudaError_t best_of_starts_launch(const zhsm_result *results, int P, int G, int criterion, int32_t *best,
                                  zhsm_result *best_results, cudaStream_t st);
cudaError_t corridor_launch(const float *scans, const unsigned char *bad, const short *seg, const float *fi,
                            const float *si, const float *co, int S, int n, signed char *out, cudaStream_t st);
zhsm_status zhsm_stage(zhsm_handle *h, int slot, size_t bytes, void **out);

cudaError_t plicp_launch(const PlicpArgs &a, int sm_count, cudaStream_t st, int *launched);
// First iterations of an icp_loop searched in chunk order (thread t walks rays [t*C, (t+1)*C) and hands last_best on:
// fewer re-searched walks, lanes that wait for each other more).  Measured on one B200 (profiles/r02x_chunk_iters.txt):
// 1080 beams, 192 threads: 0: 16.7+, 2: 16.33, 3: 16.29, 6: 16.56 ms per 20k pairs; 2048 beams, multi-start, 256
// threads (long walks go to the list): 0: 87.3 k, 2: 84.4 k, 6: 77.1 k matches/s.
// (the instances that run their walks as a pool, T >= 256, ignore the value: they have no chunk order)
inline int plicp_chunk_iters(int knob, int T) { return knob >= 0 ? knob : (T >= 256 ? 0 : 2); }
// Short scans in LARGE batches run 64 threads per job (8 jobs per SM instead of 4): measured on one B200
// (profiles/r03c_cta_shapes.txt, r03s_logs_by_threads_fanout.txt), 64 against 128 threads: uniform batches of 60-80 k
// pairs +17 % at 181 beams, +19 % at 256, +9 % at 361; intel-lab x 16 (42 736 jobs, restarts as a second batch) 1.93 M
// against 1.58 M pairs/s, fr079 x 16 +8 %, mit-cscail x 16 -2 %; the logs ONCE (1-3 k jobs: a few waves, where a job's
// own speed counts) 7-27 % slower.  0: the handle's shape.
inline int plicp_wide_batch_T(int n, int B, int knob_T) { return (knob_T == 0 && n <= 256 && B >= 16384) ? 64 : 0; }
size_t plicp_ms128_smem_bytes(int n);  // dynamic shared memory of the multi-start instance (4 jobs per CTA)
size_t plicp_jump_bytes_per_job(int n);
// CTA shape the PL-ICP kernel uses for n beams: T threads, C rays per thread, hash chunk (even) per thread
void plicp_shape(int n, int knob_T, int *T, int *C, int *hash_chunk);
void plicp_build_hash_matrices(int n, int T, int ch, std::vector<uint32_t> &mat);
cudaError_t median_filter_launch(float *scans, int S, int n, cudaStream_t st);

#define ZHSM_CUDA_TRY(h, expr)                                                              \
    do {                                                                                    \
        cudaError_t _e = (expr);                                                            \
        if (_e != cudaSuccess) {                                                            \
            (h)->err = std::string(#expr) + ": " + cudaGetErrorString(_e);                  \
            return ZHSM_ERR_CUDA;                                                           \
        }                                                                                   \
    } while (0)
