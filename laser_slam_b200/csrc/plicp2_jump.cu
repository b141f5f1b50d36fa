// laser_slam_b200/csrc/plicp2_jump.cu -- ld_create_jump_tables(laser_ref) (CSM laser_data.c)
// as its own kernel, run once per job ahead of the PL-ICP kernel.
//
// For every valid ref ray i
//   up_bigger[i]   = (first j > i that is invalid or has r[j] > r[i]) - i      (n - i when there is none)
//   up_smaller[i]  = (first j > i that is invalid or has r[j] < r[i]) - i
//   down_bigger / down_smaller the same towards j < i                          (-1 - i when there is none)
// A jump can be hundreds of rays long (a wall seen head-on: r falls to the foot of the perpendicular and rises again),
// so a linear scan per entry is O(n * jump), and CSM's own loop is sequential.  Here every entry is searched
// independently over a 3-level structure: the rays, the extrema of aligned blocks of 8 rays and of 64 rays.  A search
// runs to the end of its 8-block, then over 8-blocks to the end of its 64-block, then over 64-blocks, and descends
// again: at most 7+7+nb64+8+8 probes (47 at 1080 rays), 1-3 for the common short jumps.
//
// Why a separate kernel: the searches are chains of dependent shared-memory loads with divergent trip counts.
// Inside the PL-ICP kernel (24 warps per SM, limited by its shared memory) that phase took 10 % of
// the job; here a CTA needs 5 KB, 64 warps per SM hide the latency, and the tables cost 8 bytes per
// ray of HBM traffic each way (17 KB per job: nothing against 7.7 TB/s).
//
// The comparisons are made on the raw float readings (cm): r = (double)f / 100 is strictly monotone in f, so
// r[j] <= r[i]  <=>  f[j] <= f[i], and validity is reading_of() of the PL-ICP kernel.
//
// Entries of INVALID rays: CSM never reads them -- its walk skips an invalid ref ray with ++up / --down and nothing
// else changes, so the next trip takes the same direction again until a valid ray or the end of the scan.  Both
// entries of an invalid ray therefore hold the length of that whole skip (to the next valid ray; n - i / -1 - i when
// there is none), and the kernel's walk crosses a run of invalid rays in one probe.
#include "plicp2_common.cuh"

namespace plicp2 {

// MODE 0 (bigger): the search stops at an invalid ray (key -1) or a bigger one; MODE 1 (smaller): at an invalid or a
// smaller one; MODE 2 (next valid, for invalid rays): at any valid ray
template <int MODE>
__device__ __forceinline__ bool jt_stop(float k, float v) {
    return MODE == 0 ? (k < 0 || k > v) : MODE == 1 ? (k < v) : (k >= 0);
}
// block extrema: MODE 0 uses max(valid ? f : +inf), MODE 1 min(key) (an invalid ray makes it -1 < v), MODE 2 max(key)
template <int MODE>
__device__ __forceinline__ bool jt_blk(float m, float v) { return MODE == 0 ? (m > v) : MODE == 1 ? (m < v) : (m >= 0); }

template <int BIG>
__device__ __forceinline__ int jt_search_up(const float *key, const float *m8, const float *m64, int n, int nb8,
                                            int nb64, int j, float v) {
    while (j < n && (j & 7)) { if (jt_stop<BIG>(key[j], v)) return j; j++; }
    if (j >= n) return n;
    int b = j >> 3;
    bool hit = false;
    while (b < nb8 && (b & 7)) { if (jt_blk<BIG>(m8[b], v)) { hit = true; break; } b++; }
    if (!hit) {
        if (b >= nb8) return n;
        int c = b >> 3;
        while (c < nb64 && !jt_blk<BIG>(m64[c], v)) c++;
        if (c >= nb64) return n;
        b = c << 3;
        while (!jt_blk<BIG>(m8[b], v)) b++;  // the block extrema only cover existing rays: stops inside the data
    }
    j = b << 3;
    while (!jt_stop<BIG>(key[j], v)) j++;
    return j;
}

template <int BIG>
__device__ __forceinline__ int jt_search_down(const float *key, const float *m8, const float *m64, int j, float v) {
    while (j >= 0 && (j & 7) != 7) { if (jt_stop<BIG>(key[j], v)) return j; j--; }
    if (j < 0) return -1;
    int b = j >> 3;
    bool hit = false;
    while (b >= 0 && (b & 7) != 7) { if (jt_blk<BIG>(m8[b], v)) { hit = true; break; } b--; }
    if (!hit) {
        if (b < 0) return -1;
        int c = b >> 3;
        while (c >= 0 && !jt_blk<BIG>(m64[c], v)) c--;
        if (c < 0) return -1;
        b = (c << 3) + 7;
        while (!jt_blk<BIG>(m8[b], v)) b--;
    }
    j = (b << 3) + 7;  // every block at or below the start block is complete
    while (!jt_stop<BIG>(key[j], v)) j--;
    return j;
}

constexpr int JT_THREADS = 256;

// Jobs that share a ref scan (the G starts of a multi-start pair, many-to-one loop-closure candidates) share its
// tables: the lowest job of the launch that uses a scan OWNS it, builds the tables into its own scratch slot, and the
// others read that slot.  owner[] is per scan; only the entries this launch touches are reset and claimed.
__global__ void owner_reset_kernel(const int32_t *ref_idx, int B, int S, int32_t *owner) {
    const int job = blockIdx.x * blockDim.x + threadIdx.x;
    if (job >= B) return;
    const int r = ref_idx[job];
    if ((unsigned)r < (unsigned)S) owner[r] = 0x7fffffff;
}
__global__ void owner_claim_kernel(const int32_t *ref_idx, int B, int S, int32_t *owner) {
    const int job = blockIdx.x * blockDim.x + threadIdx.x;
    if (job >= B) return;
    const int r = ref_idx[job];
    if ((unsigned)r < (unsigned)S) atomicMin(&owner[r], job);
}

// jt: [B][2][n][2] shorts: per job the pairs (up_bigger, up_smaller)[n] then (down_bigger, down_smaller)[n] of its
// ref scan -- the walk fetches both jumps of a direction with one 32-bit load
__global__ void __launch_bounds__(JT_THREADS) jump_tables_kernel(PlicpArgs a, short *jt) {
    extern __shared__ float jsm[];
    const int n = a.L.n, t = threadIdx.x, lane = t & 31;
    const int nb8 = (n + 7) >> 3, nb64 = (nb8 + 7) >> 3;
    float *key = jsm;                       // [n] reading in cm, -1 when invalid in the LDP
    float *hi8 = key + ((n + 31) & ~31);    // [nb8] max(valid ? key : +inf)
    float *lo8 = hi8 + nb8;                 // [nb8] min(key)
    float *hi64 = lo8 + nb8, *lo64 = hi64 + nb64;
    float *any8 = lo64 + nb64, *any64 = any8 + nb8;  // max(key): >= 0 when the block holds a valid ray
    const int job = blockIdx.x;
    const int ref = a.ref_idx[job];
    if ((unsigned)ref >= (unsigned)a.S || a.owner[ref] != job) return;  // bad index (the PL-ICP kernel reports it) / not the owner
    const float *row = a.scans + (size_t)ref * n;
    const int rounds = (n + JT_THREADS - 1) / JT_THREADS;
    for (int k = 0; k < rounds; k++) {
        const int i = k * JT_THREADS + t;
        float hi = 0.f, lo = INFINITY, any = -1.f;  // neutral for rays past the end of the scan
        if (i < n) {
            int bad = 0;
            const float f = __ldg(row + i);
            const float kf = reading_of(f, a.L, a.P, &bad) >= 0 ? f : -1.f;
            key[i] = kf;
            hi = kf >= 0 ? kf : INFINITY;
            lo = kf;
            any = kf;
        }
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            any = fmaxf(any, __shfl_xor_sync(0xffffffffu, any, o));
        }
        if (i < n && (lane & 7) == 0) { hi8[i >> 3] = hi; lo8[i >> 3] = lo; any8[i >> 3] = any; }
    }
    __syncthreads();
    if (t < nb64) {
        float hi = 0.f, lo = INFINITY, any = -1.f;
        const int e = (t << 3) + 8 < nb8 ? (t << 3) + 8 : nb8;
        for (int b = t << 3; b < e; b++) { hi = fmaxf(hi, hi8[b]); lo = fminf(lo, lo8[b]); any = fmaxf(any, any8[b]); }
        hi64[t] = hi;
        lo64[t] = lo;
        any64[t] = any;
    }
    __syncthreads();
    short2 *out2 = reinterpret_cast<short2 *>(jt + (size_t)job * 4 * n);
    for (int i = t; i < n; i += JT_THREADS) {
        const float v = key[i];
        int ub, us, db, ds;
        if (v >= 0) {
            ub = jt_search_up<0>(key, hi8, hi64, n, nb8, nb64, i + 1, v);
            us = jt_search_up<1>(key, lo8, lo64, n, nb8, nb64, i + 1, v);
            db = jt_search_down<0>(key, hi8, hi64, i - 1, v);
            ds = jt_search_down<1>(key, lo8, lo64, i - 1, v);
        } else {
            ub = us = jt_search_up<2>(key, any8, any64, n, nb8, nb64, i + 1, v);
            db = ds = jt_search_down<2>(key, any8, any64, i - 1, v);
        }
        out2[i] = make_short2((short)(ub - i), (short)(us - i));
        out2[n + i] = make_short2((short)(db - i), (short)(ds - i));
    }
}

}  // namespace plicp2

size_t plicp_jump_bytes_per_job(int n) { return sizeof(short) * 4 * (size_t)n; }

cudaError_t plicp_jump_launch(const PlicpArgs &a, short *jt, int32_t *owner, cudaStream_t st, int *launched) {
    const int n = a.L.n;
    plicp2::owner_reset_kernel<<<(a.B + 255) / 256, 256, 0, st>>>(a.ref_idx, a.B, a.S, owner);
    plicp2::owner_claim_kernel<<<(a.B + 255) / 256, 256, 0, st>>>(a.ref_idx, a.B, a.S, owner);
    if (launched) *launched += 3;
    const int nb8 = (n + 7) >> 3, nb64 = (nb8 + 7) >> 3;
    const size_t smem = sizeof(float) * (size_t)(((n + 31) & ~31) + 3 * nb8 + 3 * nb64);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(plicp2::jump_tables_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    plicp2::jump_tables_kernel<<<a.B, plicp2::JT_THREADS, smem, st>>>(a, jt);
    return cudaGetLastError();
}
