// laser_slam_b200/csrc/plicp2_t128.cu -- PL-ICP kernel instances for scans of up to 1152 beams (128-thread CTAs).
#include "plicp2_kernel.cuh"
#include "plicp2_launch.cuh"

cudaError_t plicp_launch_t128(const PlicpArgs &a, int C, cudaStream_t st) {
    switch (C) {
        PLICP2_CASE(128, 2)
        PLICP2_CASE(128, 3)
        PLICP2_CASE(128, 5)
        PLICP2_CASE(128, 7)
        PLICP2_CASE(128, 9)
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}
