// laser_slam_b200/csrc/plicp2_t64.cu -- PL-ICP kernel instance with 64-thread CTAs (one CTA per scan pair): eight
// resident CTAs per SM at 128 registers.  Tuning only (ZHSM_PLICP_T=64): faster on uniform batches of short scans, slower
// on the real logs (plicp2_dispatch.cu).
#include "plicp2_kernel.cuh"
#include "plicp2_launch.cuh"

cudaError_t plicp_launch_t64(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(64)
