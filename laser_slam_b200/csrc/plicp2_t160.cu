// laser_slam_b200/csrc/plicp2_t160.cu -- PL-ICP kernel instance with 160-thread CTAs (one CTA per scan pair), compiled
// for five resident CTAs per SM (80 registers): scans whose shared-memory footprint lets five jobs share an SM.
#include "plicp2_kernel.cuh"
#include "plicp2_launch.cuh"

cudaError_t plicp_launch_t160(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(160)
