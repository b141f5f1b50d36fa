// laser_slam_b200/csrc/plicp2_t256.cu -- PL-ICP kernel instances for scans of 1153..2304 beams (256-thread CTAs).
#include "plicp2_kernel.cuh"
#include "plicp2_launch.cuh"

cudaError_t plicp_launch_t256(const PlicpArgs &a, int C, cudaStream_t st) {
    switch (C) {
        PLICP2_CASE(256, 5)
        PLICP2_CASE(256, 7)
        PLICP2_CASE(256, 9)
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}
