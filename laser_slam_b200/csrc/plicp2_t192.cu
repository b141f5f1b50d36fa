// laser_slam_b200/csrc/plicp2_t192.cu -- PL-ICP kernel instance with 192-thread CTAs (one CTA per scan pair).
#include "plicp2_kernel.cuh"
#include "plicp2_launch.cuh"

cudaError_t plicp_launch_t192(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(192)
