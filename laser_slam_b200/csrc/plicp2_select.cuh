// laser_slam_b200/csrc/plicp2_select.cuh -- block-wide helpers: integer sums, exact order statistics, ld_corr_hash.
#pragma once
#include "plicp2_common.cuh"

namespace PL_NAMESPACE {

// Block-wide integer sum, result to every thread, ONE barrier.  `slot` selects one of 4 scratch rows; a row must
// not be reused until another barrier has passed (callers alternate rows).
template <int T>
__device__ __forceinline__ int block_sum(int v, int *ired, int slot) {
    constexpr int NW = T / 32;
    int w = __reduce_add_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0) ired[slot * 16 + (pl_tid<T>() >> 5)] = w;
    pl_sync<T>();
    int tot = 0;
#pragma unroll
    for (int i = 0; i < NW; i++) tot += ired[slot * 16 + i];
    return tot;
}

// kill_outliers_trim sorts the distances and reads two order statistics; only the two VALUES are
// needed, so this is an 8-bit MSD radix select on the bit patterns of the non-negative doubles in keyS[] for both
// ranks at once.  Per pass: one histogram (both ranks packed in one int: low/high 16 bits), one barrier, then every
// warp redundantly locates the buckets (no second barrier).  A bucket holding a single key ends that rank early
// (the key's owner index is remembered next to the histogram); three rotating histogram sets make the zeroing of
// the next set race-free.
//
// The first two passes cost no pass of their own.  Distances of one iteration share their top byte (sign + 7
// exponent bits: 0x3f for 2^-15 <= d < 1, 0x3e below), so the loop that WRITES the keys (plicp2_kernel.cuh) already
// histograms their second byte into set 1, for the keys whose top byte equals a guess (the top byte of the previous
// iteration's result), and counts the keys below the guess.  When both ranks fall among the guessed keys -- all but
// never otherwise -- the select continues at pass 2; else it starts over at pass 0.

// bucket of rank rem[r] in a 256-bin histogram (both ranks packed 16:16), computed by every lane of a warp:
// digit[r], the rank inside that bucket nrem[r], the bucket's population cnt[r].  total[r] = keys in the histogram.
__device__ __forceinline__ void locate2(const int *hist, int lane, const int rem[2], int digit[2], int nrem[2],
                                        int cnt[2], int total[2]) {
    int loc[8];
#pragma unroll
    for (int q = 0; q < 8; q++) loc[q] = hist[lane * 8 + q];
    int sum = 0;  // packed: the low halves cannot carry into the high ones (fewer than 65536 keys)
#pragma unroll
    for (int q = 0; q < 8; q++) sum += loc[q];
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    const int tot = __shfl_sync(0xffffffffu, incl, 31);
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const int sum_r = (sum >> (16 * r)) & 0xffff, incl_r = (incl >> (16 * r)) & 0xffff;
        const int excl = incl_r - sum_r;
        const bool mine = rem[r] >= excl && rem[r] < incl_r;
        int bsel = 0, acc = excl, cb = 0;
        bool found = false;
#pragma unroll
        for (int q = 0; q < 8; q++) {
            const int c = (loc[q] >> (16 * r)) & 0xffff;
            const bool hit = !found && rem[r] < acc + c;
            bsel = hit ? q : bsel;
            cb = hit ? c : cb;
            found |= hit;
            acc = found ? acc : acc + c;
        }
        const unsigned who = __ballot_sync(0xffffffffu, mine);
        const int src = who ? __ffs(who) - 1 : 0;  // exactly one lane when the rank is inside the population
        digit[r] = __shfl_sync(0xffffffffu, lane * 8 + bsel, src);
        nrem[r] = __shfl_sync(0xffffffffu, rem[r] - acc, src);
        cnt[r] = __shfl_sync(0xffffffffu, cb, src);
        total[r] = (tot >> (16 * r)) & 0xffff;
    }
}

// top_guess / nbelow: the guessed top byte and the number of keys whose top byte is below it (see above); the keys'
// second-byte histogram is in set 1, the owner of a key per bucket in cand set 1 (first half).
template <int T, class Clock>
__device__ __forceinline__ void select2(const Lay &l, int C, unsigned valid, int rank0, int rank1, int top_guess,
                                        const int *nbelow_p, double &v0, double &v1, Clock &pc) {
    const int t = pl_tid<T>(), lane = t & 31;
    const unsigned long long *keyS = PL_SH(const unsigned long long, l.d2j);
    int *hist_all = PL_SH(int, l.hist);
    short *cand_all = PL_SH(short, l.candi);
    unsigned long long prefix[2] = {0, 0};
    int rem[2] = {rank0, rank1};
    bool done[2] = {false, false};
    double val[2] = {0, 0};
    int first_pass = 0;
    pl_sync<T>();  // the keys, their histogram (set 1) and *nbelow_p are in place
    pc.mark(18);
    {
        const int nbelow = *nbelow_p;
        // the fused passes 0+1: set 1 holds the second byte of the keys with the guessed top byte
        const int r2[2] = {rank0 - nbelow, rank1 - nbelow};
        int digit[2], nrem[2], cnt[2], total[2];
        locate2(hist_all + 256, lane, r2, digit, nrem, cnt, total);
        if (r2[0] >= 0 && r2[0] < total[0] && r2[1] >= 0 && r2[1] < total[1]) {
            first_pass = 2;
#pragma unroll
            for (int r = 0; r < 2; r++) {
                prefix[r] = ((unsigned long long)top_guess << 8) | (unsigned long long)digit[r];
                rem[r] = nrem[r];
                if (cnt[r] == 1) {
                    done[r] = true;
                    val[r] = __longlong_as_double((long long)keyS[cand_all[512 + digit[r]]]);
                }
            }
        } else {
            pl_sync<T>();  // slow warps may still be reading set 1, which pass 0 is about to zero
        }
    }
    pc.mark(19);
#pragma unroll 1
    for (int pass = first_pass; pass < 8 && !(done[0] && done[1]); pass++) {
        const int shift = 56 - 8 * pass;
        int *hist = hist_all + 256 * (pass % 3);
        short *cand = cand_all + 512 * (pass % 3);
        int *hnext = hist_all + 256 * ((pass + 1) % 3);
        for (int b = t; b < 256; b += T) hnext[b] = 0;
#pragma unroll 1
        for (int k = 0; k < C; k++) {
            if (!(valid >> k & 1u)) continue;
            const int i = k * T + t;
            const unsigned long long u = keyS[i];
            const unsigned long long hi = pass == 0 ? 0ull : (u >> (shift + 8));
            const int digit = (int)((u >> shift) & 255ull);
            const bool m0 = !done[0] && hi == prefix[0], m1 = !done[1] && hi == prefix[1];
            if (m0 || m1) atomicAdd(&hist[digit], (m0 ? 1 : 0) | (m1 ? 0x10000 : 0));
            if (m0) cand[digit] = (short)i;
            if (m1) cand[256 + digit] = (short)i;
        }
        pl_sync<T>();
        pc.mark(18);
        int digit[2], nrem[2], cnt[2], total[2];
        locate2(hist, lane, rem, digit, nrem, cnt, total);
#pragma unroll
        for (int r = 0; r < 2; r++) {
            if (done[r]) continue;
            prefix[r] = (prefix[r] << 8) | (unsigned long long)digit[r];
            rem[r] = nrem[r];
            if (cnt[r] == 1 || pass == 7) {
                done[r] = true;
                val[r] = pass == 7 ? __longlong_as_double((long long)prefix[r])
                                   : __longlong_as_double((long long)keyS[cand[256 * r + digit[r]]]);
            }
        }
        pc.mark(19);
    }
    v0 = val[0];
    v1 = val[1];
}

// ld_corr_hash (CSM laser_data.c) over the stream v_i = j1 + 1000*j2 (or -1):
//     even i: h ^= (h<<7) ^ v ^ (h>>3)        odd i: h ^= ~((h<<11) ^ v ^ (h>>5))
// Both steps are affine over GF(2), so the hash of the whole stream is the XOR over chunks of
// (linear map of the steps that follow the chunk) applied to (the chunk hashed from 0).  Thread t hashes the chunk
// [t*ch, (t+1)*ch) (ch even) and multiplies by its 32x32 bit matrix hashmat[t] (built on the host by zhsm_set_laser).
__device__ __forceinline__ unsigned corr_hash_chunk(const Lay &l, int n, int t, int ch, const uint32_t *hashmat) {
    const short *cj1 = PL_SH(const short, l.cj1), *cj2 = PL_SH(const short, l.cj2);
    const int lo = t * ch;
    if (lo >= n) return 0u;
    const int hi = lo + ch < n ? lo + ch : n;
    unsigned h = 0;
    for (int i = lo; i < hi; i++) {
        const int j1 = cj1[i];
        const unsigned v = j1 < 0 ? 0xffffffffu : (unsigned)(j1 + 1000 * (int)cj2[i]);
        h ^= ((i & 1) == 0) ? ((h << 7) ^ v ^ (h >> 3)) : (~((h << 11) ^ v ^ (h >> 5)));
    }
    const uint4 *m = reinterpret_cast<const uint4 *>(hashmat + 32 * t);
    unsigned out = 0;
#pragma unroll
    for (int q = 0; q < 8; q++) {
        const uint4 c = __ldg(m + q);
        out ^= c.x & (0u - ((h >> (4 * q + 0)) & 1u));
        out ^= c.y & (0u - ((h >> (4 * q + 1)) & 1u));
        out ^= c.z & (0u - ((h >> (4 * q + 2)) & 1u));
        out ^= c.w & (0u - ((h >> (4 * q + 3)) & 1u));
    }
    return out;
}

}  // namespace PL_NAMESPACE
