// laser_slam_b200/csrc/plicp2_ms128.cu -- the MULTI-START instance of the PL-ICP kernel: one CTA of 4 x 128 threads owns
// four jobs that share both scans (four first guesses of one scan pair: loop-closure multi-start, SURVEY 8d C5 / 8e
// "scans staged in shared memory once per CTA"), over one staged copy of the ref scan's points, squared readings and
// jump tables and of the sens readings.  At 2048 beams: 74 KB shared + 36 KB per job = 218 KB, four jobs per SM where
// two single-job CTAs (105 KB, 256 threads each) fit; the same 512 threads x 128 registers per SM.  The jobs never wait
// for each other after the load (named barriers per job).  Its own namespace: the single-job instances do not change.
#define PLICP_NS 4
#define PL_NAMESPACE plicp2ms
#define PLICP_LIST_MIN_T 128  // long walks are the rule on this workload: finished warp-wide (plicp2_search.cuh)
#include "plicp2_kernel.cuh"
#include "plicp2_launch.cuh"

cudaError_t plicp_launch_ms128(const PlicpArgs &a, cudaStream_t st) PLICP2_LAUNCH(128)

size_t plicp_ms128_smem_bytes(int n) { return plicp2ms::smem_bytes(n, 128); }
