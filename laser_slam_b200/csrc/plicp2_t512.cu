// laser_slam_b200/csrc/plicp2_t512.cu -- PL-ICP kernel instances for scans of 2305..4096 beams (512-thread CTAs).
#include "plicp2_kernel.cuh"
#include "plicp2_launch.cuh"

cudaError_t plicp_launch_t512(const PlicpArgs &a, int C, cudaStream_t st) {
    switch (C) {
        PLICP2_CASE(512, 6)
        PLICP2_CASE(512, 8)
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}
