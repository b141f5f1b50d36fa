// laser_slam_b200/csrc/plicp2_t64.cu -- PL-ICP kernel instance with 64-thread CTAs (one CTA per scan pair): scans of up
// to 400 beams (the 181- and 361-beam SICK logs), eight resident CTAs per SM at 128 registers.
#include "plicp2_kernel.cuh"
#include "plicp2_launch.cuh"

cudaError_t plicp_launch_t64(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(64)
