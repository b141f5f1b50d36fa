// laser_slam_b200/csrc/plicp2_final.cu -- the FINAL-ONLY instances of the PL-ICP kernel, every CTA shape in one
// translation unit: load, "recompute correspondences in case of restarts" at the given pose, ld_corr_hash, covariance,
// result record -- no icp_loop.  The last pass of a job whose restarts ran as a second batch (zhsm_api.cu:
// plicp_fanout): sm_icp's bookkeeping (PlicpArgs::fin) and the record to write (PlicpArgs::out_idx) come with the launch.
#define PLICP_FINAL 1
#define PL_NAMESPACE plicp2fin
#include "plicp2_kernel.cuh"
#include "plicp2_launch.cuh"

static cudaError_t fin64(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(64)
static cudaError_t fin128(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(128)
static cudaError_t fin160(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(160)
static cudaError_t fin192(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(192)
static cudaError_t fin256(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(256)
static cudaError_t fin512(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(512)

cudaError_t plicp_launch_final(const PlicpArgs &a, cudaStream_t st) {
    if (!a.fin || !a.out_idx) return cudaErrorInvalidValue;
    switch (a.T) {
        case 64: return fin64(a, st);
        case 128: return fin128(a, st);
        case 160: return fin160(a, st);
        case 192: return fin192(a, st);
        case 256: return fin256(a, st);
        case 512: return fin512(a, st);
    }
    return cudaErrorInvalidValue;
}
