// laser_slam_b200/csrc/plicp2_t64.cu -- PL-ICP kernel instance with 64-thread CTAs (one CTA per scan pair): eight
// resident CTAs per SM at 128 registers.  Short scans (up to 256 beams) in large batches (plicp_wide_batch_T,
// zhsm_internal.h); a few waves of jobs are faster with 128 threads each.
#include "plicp2_kernel.cuh"
#include "plicp2_launch.cuh"

cudaError_t plicp_launch_t64(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(64)
