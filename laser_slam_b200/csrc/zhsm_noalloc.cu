// laser_slam_b200/csrc/zhsm_noalloc.cu -- zhsm_plicp_batch_noalloc (include/zhsm.h): batched PL-ICP for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// No counterpart in the reference.
//
// zhsm_plicp_batch keeps its scratch (the jump tables of the jobs' ref scans, 8 bytes per ray and job, and the per-scan
// owner table; plicp2_jump.cu) in handle-owned buffers that grow with cudaMalloc.  Here the same kernels get their
// scratch from ONE file-scope device array of fixed size: nothing is allocated, inputs and outputs are the caller's
// device buffers.  Every GPU that loads libzhsm.so holds the array (8 MiB + 32 KiB) until the process exits.
//
// The array is shared by every call on a device and keeps the last call's contents.  Nothing in it has to start at
// zero: owner_reset_kernel re-initialises exactly the owner entries a launch touches before they are claimed, and an
// owner writes all n entries of its jump-table slot before any job reads it (same stream).  Calls that OVERLAP on the
// GPU (two streams, two handles on one device) race on it: issue them on one stream, or order them with events.
#include <string.h>

#include <mutex>

#include "zhsm_internal.h"

constexpr size_t JT_BYTES = (size_t)ZHSM_NOALLOC_MAX_RAYS * 8;         // short[4] per ray
constexpr size_t OWNER_BYTES = (size_t)ZHSM_NOALLOC_MAX_SCANS * 4;     // int32 per scan
constexpr int MAX_DEVICES = 64;

// In host code this name is a host-side stand-in, NOT a device address: it goes through cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[JT_BYTES + OWNER_BYTES];

namespace {

std::mutex g_mu;
void *g_scratch_ptr[MAX_DEVICES] = {nullptr};  // per device: the array's address in that device's memory

// once per device, with that device current
cudaError_t scratch_address(int device, unsigned char **out) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_ptr[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_ptr[device] = p;
    }
    *out = (unsigned char *)g_scratch_ptr[device];
    return cudaSuccess;
}

}  // namespace

extern "C" zhsm_status zhsm_plicp_batch_noalloc(zhsm_handle *h, const float *d_scans, int S, const int32_t *d_ref_idx,
                                                const int32_t *d_sens_idx, const double *d_first_guess, int B,
                                                const zhsm_plicp_params *params, zhsm_result *d_results,
                                                int16_t *d_corr_j1, int16_t *d_corr_j2, void *cuda_stream) {
    if (!h) return ZHSM_ERR_ARG;
    if (!d_scans || !d_ref_idx || !d_sens_idx || !d_results || S <= 0 || B < 0 ||
        ((d_corr_j1 == nullptr) != (d_corr_j2 == nullptr))) {
        h->err = "zhsm_plicp_batch_noalloc: bad argument";
        return ZHSM_ERR_ARG;
    }
    if (!h->have_laser) { h->err = "zhsm_set_laser not called"; return ZHSM_ERR_ARG; }
    if (B == 0) return ZHSM_OK;
    const int n = h->laser.n;
    if ((unsigned long long)B * (unsigned long long)n > (unsigned long long)ZHSM_NOALLOC_MAX_RAYS) {
        h->err = "zhsm_plicp_batch_noalloc: B * beams exceeds ZHSM_NOALLOC_MAX_RAYS (1048576): the static scratch holds "
                 "8 bytes per ray and job; split the batch or use zhsm_plicp_batch";
        return ZHSM_ERR_ARG;
    }
    if (S > ZHSM_NOALLOC_MAX_SCANS) {
        h->err = "zhsm_plicp_batch_noalloc: S exceeds ZHSM_NOALLOC_MAX_SCANS (8192): pass a smaller scan pool or use "
                 "zhsm_plicp_batch";
        return ZHSM_ERR_ARG;
    }
    if (h->device < 0 || h->device >= MAX_DEVICES) { h->err = "zhsm_plicp_batch_noalloc: device index above 63"; return ZHSM_ERR_ARG; }
    zhsm_plicp_params P;
    if (params) P = *params; else zhsm_plicp_default_params(&P, h->laser.min_range_cm, h->laser.max_range_cm);
    if (P.use_corr_tricks != 1 || P.use_point_to_line_distance != 1) {
        h->err = "only use_corr_tricks=1 / use_point_to_line_distance=1 (the reference's settings) are implemented";
        return ZHSM_ERR_UNSUPPORTED;
    }
    if (P.max_iterations < 1 || P.max_iterations > ZHSM_MAX_ITER) { h->err = "max_iterations out of [1,64]"; return ZHSM_ERR_ARG; }
    {   // ld_valid_fields' per-geometry checks, as in zhsm_plicp_batch
        const float thf0 = h->laser.fi_min, thf1 = h->laser.fi_min + h->laser.dfi * (float)(unsigned)(n - 1);
        const double fov = (double)thf1 - (double)thf0, PI = 3.14159265358979323846;
        if (n < 10 || n > 10000 || !(fov >= 20.0 * PI / 180.0 && fov <= 2.01 * PI)) {
            h->err = "geometry rejected by ld_valid_fields (fov must be in [20deg, 2.01*pi], 10<=n<=10000)";
            return ZHSM_ERR_ARG;
        }
    }
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    unsigned char *scratch = nullptr;
    ZHSM_CUDA_TRY(h, scratch_address(h->device, &scratch));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
    StreamScope scope(h, st);

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
    a.scans = d_scans;
    a.ref_idx = d_ref_idx;
    a.sens_idx = d_sens_idx;
    a.first_guess = d_first_guess;
    a.results = d_results;
    a.corr_j1 = d_corr_j1;
    a.corr_j2 = d_corr_j2;
    a.jt = reinterpret_cast<const short *>(scratch);                       // [B][2][n][2] shorts, B*n*8 <= JT_BYTES
    a.owner = reinterpret_cast<const int32_t *>(scratch + JT_BYTES);       // [S] ints, S*4 <= OWNER_BYTES
    ZHSM_CUDA_TRY(h, cudaEventRecord(h->ev0, st));
    int launched = 0;
    ZHSM_CUDA_TRY(h, plicp_launch(a, h->sm_count, st, &launched));
    h->launches += launched;
    ZHSM_CUDA_TRY(h, cudaEventRecord(h->ev1, st));
    h->ev_valid = true;
    return ZHSM_OK;  // asynchronous: the records are ready when `st` has drained (zhsm_sync)
}
