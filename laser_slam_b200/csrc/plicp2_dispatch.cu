// laser_slam_b200/csrc/plicp2_dispatch.cu -- host side of the PL-ICP kernel: CTA shape per beam count, the GF(2)
// matrices of the parallel ld_corr_hash, and the launch dispatch over the compiled instances.
#include <vector>

#include "zhsm_internal.h"

cudaError_t plicp_launch_t128(const PlicpArgs &a, int C, cudaStream_t st);
cudaError_t plicp_launch_t256(const PlicpArgs &a, int C, cudaStream_t st);
cudaError_t plicp_launch_t512(const PlicpArgs &a, int C, cudaStream_t st);

static int round_up_to(int c, const int *opts, int nopts) {
    for (int i = 0; i < nopts; i++) if (c <= opts[i]) return opts[i];
    return -1;
}

void plicp_shape(int n, int *T, int *C, int *hash_chunk) {
    static const int c128[] = {2, 3, 5, 7, 9}, c256[] = {5, 7, 9}, c512[] = {6, 8};
    int t, c;
    if (n <= 128 * 9) { t = 128; c = round_up_to((n + t - 1) / t, c128, 5); }
    else if (n <= 256 * 9) { t = 256; c = round_up_to((n + t - 1) / t, c256, 3); }
    else { t = 512; c = round_up_to((n + t - 1) / t, c512, 2); }
    *T = t;
    *C = c;
    int ch = (n + t - 1) / t;
    if (ch & 1) ch++;
    *hash_chunk = ch;
}

// ld_corr_hash steps without the data term (CSM laser_data.c): the linear parts over GF(2)
static inline uint32_t hash_lin_even(uint32_t h) { return h ^ (h << 7) ^ (h >> 3); }
static inline uint32_t hash_lin_odd(uint32_t h) { return h ^ (h << 11) ^ (h >> 5); }

// mat[t][b] = image of bit b under the linear part of steps i = min(n,(t+1)*ch) .. n-1 (identity when none remain)
void plicp_build_hash_matrices(int n, int T, int ch, std::vector<uint32_t> &mat) {
    mat.assign((size_t)T * 32, 0u);
    for (int t = 0; t < T; t++) {
        long long first = (long long)(t + 1) * ch;
        for (int b = 0; b < 32; b++) {
            uint32_t h = 1u << b;
            for (long long i = first; i < n; i++) h = (i & 1) ? hash_lin_odd(h) : hash_lin_even(h);
            mat[(size_t)t * 32 + b] = h;
        }
    }
}

cudaError_t plicp_launch(const PlicpArgs &a, int sm_count, cudaStream_t st, int *launched) {
    (void)sm_count;
    int T, C, ch;
    plicp_shape(a.L.n, &T, &C, &ch);
    if (C < 0 || ch != a.hash_chunk) return cudaErrorInvalidValue;
    cudaError_t e = T == 128 ? plicp_launch_t128(a, C, st) : T == 256 ? plicp_launch_t256(a, C, st) : plicp_launch_t512(a, C, st);
    if (e == cudaSuccess && launched) *launched += 1;
    return e;
}
