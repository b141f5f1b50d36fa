// laser_slam_b200/csrc/plicp2_t256r.cu -- the 256-thread PL-ICP instance WITHOUT the walk pool: a round's walks are one
// tricks_one call per ray, as in the short-scan instances, long walks still go to the list.  For tests only
// (ZHSM_PLICP_POOL=0, read in zhsm_create): tests/test_plicp_gpu.py compares its records with the pool's, byte for byte.
// A run-time switch inside one instance was measured to cost the pool 2.3 % (both paths inlined into the search), so
// the other path lives in its own translation unit and namespace.
#define PLICP_FLAT 0
#define PL_NAMESPACE plicp2ray
#include "plicp2_kernel.cuh"
#include "plicp2_launch.cuh"

cudaError_t plicp_launch_t256_ray(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(256)
