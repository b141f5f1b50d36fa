// laser_slam_b200/csrc/icp_sm100.cu -- zhicp on sm_100a: the point-to-point ICP the reference reaches through
// CICPWarpper::ICPMatch (src/zhicp/ZHIcp_Warpper.cpp:125-176) -> mrpt::slam::CICP::Align (icpClassic), one CTA per
// job (cloud pair x initial guess).  SURVEY 8a-9.  MRPT 0.9.6 itself is absent from the reference tree: the algorithm
// is the published one (CICP::ICP_Method_Classic, CPointsMap::computeMatchingWith2D,
// scanmatching::leastSquareErrorRigidTransformation) with the wrapper's option values.
//
// Layout: both clouds of a job live in shared memory as SoA floats.  Per iteration
//   1. all threads transform the current cloud by the pose (double math, float result, as MRPT stores it), reduce its
//      bounding box, and keep the decimated query points;
//   2. nearest map point per query: `parts` threads split the map per query and combine (distance, index)
//      lexicographically with shuffles -- the first closest point wins like a sequential scan with `<`;
//   3. the distance/angle gate, then the "only keep the closest" filter: a shared-memory atomicMin per map point over
//      the float bits of the squared error;
//   4. warp 0 compacts the survivors in query order and accumulates the eight running float sums of the closed-form
//      rigid fit sequentially, one lane per sum (float addition order is part of the result: the loop's stop rule
//      compares pose steps with 1e-6);
//   5. thread 0 solves for the pose and advances the threshold schedule.
#include <float.h>
#include <math.h>

#include "zhsm_internal.h"

namespace icp {

constexpr int T = 256;

struct Sh {
    float2 *m1;                    // [cap] map cloud (x, y) interleaved: one 64-bit load per candidate in the search
    float *m2x, *m2y;              // [cap] current cloud
    float *qx, *qy, *qd;           // [cap] transformed query point, squared error of its match (-1: gated out)
    int *qj;                       // [cap] matched map index
    unsigned *best;                // [cap] per map point: min squared error (float bits) among its matches
    int *list;                     // [cap] kept queries in order
    double *D;                     // [3*cap] covariance Jacobian columns
    double *ctl;                   // [16]
    float *red;                    // [4 * T/32] bounding-box partials
    int *ictl;                     // [8]
};

__host__ __device__ inline size_t sh_bytes(int cap) {
    const size_t a = (size_t)((cap + 3) & ~3);
    return a * 4 * 10 + a * 24 + 16 * 8 + 4 * (T / 32) * 4 + 8 * 4 + 32;  // m1 (8) + 5 float + qj, best, list (4 each) = 40
}

__device__ inline Sh carve(unsigned char *p, int cap) {
    const size_t a = (size_t)((cap + 3) & ~3);
    Sh s;
    s.D = (double *)p; p += a * 24;
    s.ctl = (double *)p; p += 16 * 8;
    s.m1 = (float2 *)p; p += a * 8;
    float **fl[] = {&s.m2x, &s.m2y, &s.qx, &s.qy, &s.qd};
    for (int k = 0; k < 5; k++) { *fl[k] = (float *)p; p += a * 4; }
    s.qj = (int *)p; p += a * 4;
    s.best = (unsigned *)p; p += a * 4;
    s.list = (int *)p; p += a * 4;
    s.red = (float *)p; p += 4 * (T / 32) * 4;
    s.ictl = (int *)p;
    return s;
}

struct Match {
    int nk;       // correspondences kept
    float ratio;  // correspondencesRatio
};

// CPointsMap::computeMatchingWith2D.  Whole-CTA call; the result is valid on every thread and s.list / s.qj / s.qd
// describe the kept correspondences (query q = list[k], current-cloud index offset + q*dec, map index qj[q]).
__device__ __forceinline__ Match compute_matching(const Sh &s, int N1, int N2, double px, double py, double pphi,
                                               float maxDist, float maxAng, int keep_closest, int dec, int offset,
                                               const float gbox[4]) {
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    Match m = {0, 0.f};
    if (!N1 || !N2) return m;
    const double sin_phi = sin(pphi), cos_phi = cos(pphi);
    float lxmin = FLT_MAX, lxmax = -FLT_MAX, lymin = FLT_MAX, lymax = -FLT_MAX;
    for (int i = t; i < N2; i += T) {
        const float ox = s.m2x[i], oy = s.m2y[i];
        const float x = (float)(px + cos_phi * ox - sin_phi * oy);
        const float y = (float)(py + sin_phi * ox + cos_phi * oy);
        lxmin = fminf(lxmin, x); lxmax = fmaxf(lxmax, x);
        lymin = fminf(lymin, y); lymax = fmaxf(lymax, y);
        const int r = i - offset;
        if (r >= 0 && r % dec == 0) { s.qx[r / dec] = x; s.qy[r / dec] = y; }
    }
    for (int o = 16; o; o >>= 1) {
        lxmin = fminf(lxmin, __shfl_xor_sync(0xffffffffu, lxmin, o));
        lxmax = fmaxf(lxmax, __shfl_xor_sync(0xffffffffu, lxmax, o));
        lymin = fminf(lymin, __shfl_xor_sync(0xffffffffu, lymin, o));
        lymax = fmaxf(lymax, __shfl_xor_sync(0xffffffffu, lymax, o));
    }
    if (lane == 0) { s.red[4 * w] = lxmin; s.red[4 * w + 1] = lxmax; s.red[4 * w + 2] = lymin; s.red[4 * w + 3] = lymax; }
    __syncthreads();
    for (int k = 0; k < T / 32; k++) {
        lxmin = fminf(lxmin, s.red[4 * k]); lxmax = fmaxf(lxmax, s.red[4 * k + 1]);
        lymin = fminf(lymin, s.red[4 * k + 2]); lymax = fmaxf(lymax, s.red[4 * k + 3]);
    }
    if (lxmin > gbox[1] || lxmax < gbox[0] || lymin > gbox[3] || lymax < gbox[2]) { __syncthreads(); return m; }
    const int nq = offset < N2 ? (N2 - offset + dec - 1) / dec : 0;
    // nearest neighbour + gate.  A thread searches TWO queries at once over its share of the map (each map point is
    // loaded once and used twice); `parts` threads share a query pair and combine (distance, index) lexicographically.
    const int nqp = (nq + 1) / 2;
    int parts = 1;
    while (parts < 8 && nqp * parts * 2 <= T) parts *= 2;
    const int qper = T / parts;  // query pairs per round
    for (int q0 = 0; q0 < nqp; q0 += qper) {
        const int qp = q0 + t / parts, part = t % parts;
        const int qa = 2 * qp, qb = 2 * qp + 1;
        float bda = FLT_MAX, bdb = FLT_MAX;
        int bja = 0x7fffffff, bjb = 0x7fffffff;
        if (qa < nq) {
            const float xa = s.qx[qa], ya = s.qy[qa];
            const float xb = qb < nq ? s.qx[qb] : xa, yb = qb < nq ? s.qy[qb] : ya;
            const int per = (N1 + parts - 1) / parts;
            const int j0 = part * per, j1 = min(N1, j0 + per);
#pragma unroll 4
            for (int j = j0; j < j1; j++) {
                const float2 mp = s.m1[j];
                const float dxa = mp.x - xa, dya = mp.y - ya, dxb = mp.x - xb, dyb = mp.y - yb;
                const float da = dxa * dxa + dya * dya, db = dxb * dxb + dyb * dyb;
                if (da < bda) { bda = da; bja = j; }
                if (db < bdb) { bdb = db; bjb = j; }
            }
        }
        for (int o = 1; o < parts; o <<= 1) {
            const float oda = __shfl_xor_sync(0xffffffffu, bda, o), odb = __shfl_xor_sync(0xffffffffu, bdb, o);
            const int oja = __shfl_xor_sync(0xffffffffu, bja, o), ojb = __shfl_xor_sync(0xffffffffu, bjb, o);
            if (oda < bda || (oda == bda && oja < bja)) { bda = oda; bja = oja; }
            if (odb < bdb || (odb == bdb && ojb < bjb)) { bdb = odb; bjb = ojb; }
        }
        if (part == 0) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int q = h ? qb : qa;
                if (q >= nq) continue;
                const float bd = h ? bdb : bda;
                int bj = h ? bjb : bja;
                if (bj == 0x7fffffff) bj = 0;
                const float x = s.qx[q], y = s.qy[q];
                const double ex = px - x, ey = py - y;
                const double lim = maxAng * sqrt(ex * ex + ey * ey) + maxDist;
                const bool pass = bd < lim * lim;
                s.qd[q] = pass ? bd : -1.f;
                s.qj[q] = bj;
                if (pass) s.best[bj] = 0x7f7fffffu;
            }
        }
    }
    __syncthreads();
    if (keep_closest) {
        for (int q = t; q < nq; q += T)
            if (s.qd[q] >= 0.f) atomicMin(&s.best[s.qj[q]], __float_as_uint(s.qd[q]));
        __syncthreads();
    }
    if (w == 0) {
        int nk = 0, nOther = 0;
        for (int q0 = 0; q0 < nq; q0 += 32) {
            const int q = q0 + lane;
            bool pass = false, kept = false;
            if (q < nq) {
                const float d = s.qd[q];
                pass = d >= 0.f;
                kept = pass && (!keep_closest || s.best[s.qj[q]] == __float_as_uint(d));
            }
            const unsigned bp = __ballot_sync(0xffffffffu, pass), bk = __ballot_sync(0xffffffffu, kept);
            if (kept) s.list[nk + __popc(bk & ((1u << lane) - 1))] = q;
            nk += __popc(bk);
            nOther += __popc(bp);
        }
        if (lane == 0) { s.ictl[0] = nk; s.ictl[1] = nOther; }
    }
    __syncthreads();
    m.nk = s.ictl[0];
    m.ratio = s.ictl[1] / (float)(N2 / dec);
    return m;
}

__device__ __forceinline__ double wrap_to_pi(double a) {
    a += M_PI;
    const bool neg = a < 0;
    a = fmod(a, 2.0 * M_PI);
    if (neg) a += 2.0 * M_PI;
    return a - M_PI;
}

__device__ __forceinline__ double kernelf(int use_kernel, double x2, double rho2) { return use_kernel ? x2 / (x2 + rho2) : x2; }

__device__ __forceinline__ double parab_deriv(double w1, double q1, double w2, double q2, double w3, double q3, double x0) {
    const double A = ((q3 - q2) / ((w3 - w2) * (w3 - w1))) - ((q1 - q2) / ((w1 - w2) * (w3 - w1)));
    const double B = ((q1 - q2) + (A * ((w2 * w2) - (w1 * w1)))) / (w1 - w2);
    return (2 * A * x0) + B;
}

#define SQD(x) ((x) * (x))

__global__ void __launch_bounds__(T, 3) icp_kernel(IcpArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const Sh s = carve(smem_raw, a.cap);
    const int job = blockIdx.x, t = threadIdx.x, lane = t & 31, w = t >> 5;
    const zhsm_icp_options &o = a.opt;
    const int rc = a.ref_idx[job], cc = a.cur_idx[job];
    if ((unsigned)rc >= (unsigned)a.C || (unsigned)cc >= (unsigned)a.C) {
        // device-resident job lists are not checked by the host: a bad index fails its own job, never the batch
        if (t == 0) {
            zhsm_result r = {};
            r.status = ZHSM_JOB_INVALID_SCAN;
            a.results[job] = r;
        }
        return;
    }
    const int N1 = a.cnt[rc], N2 = a.cnt[cc];
    const float *g1 = a.pts + 2 * (size_t)a.off[rc], *g2 = a.pts + 2 * (size_t)a.off[cc];
    float gxmin = FLT_MAX, gxmax = -FLT_MAX, gymin = FLT_MAX, gymax = -FLT_MAX;
    for (int j = t; j < N1; j += T) {
        const float x = g1[j], y = g1[N1 + j];
        s.m1[j] = make_float2(x, y);
        gxmin = fminf(gxmin, x); gxmax = fmaxf(gxmax, x); gymin = fminf(gymin, y); gymax = fmaxf(gymax, y);
    }
    for (int i = t; i < N2; i += T) { s.m2x[i] = g2[i]; s.m2y[i] = g2[N2 + i]; }
    for (int ofs = 16; ofs; ofs >>= 1) {
        gxmin = fminf(gxmin, __shfl_xor_sync(0xffffffffu, gxmin, ofs));
        gxmax = fmaxf(gxmax, __shfl_xor_sync(0xffffffffu, gxmax, ofs));
        gymin = fminf(gymin, __shfl_xor_sync(0xffffffffu, gymin, ofs));
        gymax = fmaxf(gymax, __shfl_xor_sync(0xffffffffu, gymax, ofs));
    }
    if (lane == 0) { s.red[4 * w] = gxmin; s.red[4 * w + 1] = gxmax; s.red[4 * w + 2] = gymin; s.red[4 * w + 3] = gymax; }
    __syncthreads();
    float gbox[4] = {FLT_MAX, -FLT_MAX, FLT_MAX, -FLT_MAX};
    for (int k = 0; k < T / 32; k++) {
        gbox[0] = fminf(gbox[0], s.red[4 * k]); gbox[1] = fmaxf(gbox[1], s.red[4 * k + 1]);
        gbox[2] = fminf(gbox[2], s.red[4 * k + 2]); gbox[3] = fmaxf(gbox[3], s.red[4 * k + 3]);
    }
    __syncthreads();

    double mx = a.init ? a.init[3 * (size_t)job] : 0, my = a.init ? a.init[3 * (size_t)job + 1] : 0,
           mphi = a.init ? a.init[3 * (size_t)job + 2] : 0;
    double lx = 0, ly = 0, lphi = 0;
    float umbral_dist = o.thresholdDist, umbral_ang = o.thresholdAng, ratio = 0;
    const int dec = o.corresponding_points_decimation < 1 ? 1 : o.corresponding_points_decimation;
    int nIter = 0, nCorr = 0, offset = 0, last_offset = 0;
    bool keep = false;
    if (N2 > 0) {
        do {
            last_offset = offset;
            const Match m = compute_matching(s, N1, N2, mx, my, mphi, umbral_dist, umbral_ang,
                                             o.onlyClosestCorrespondences, dec, offset, gbox);
            nCorr = m.nk;
            ratio = m.ratio;
            if (!nCorr) keep = false;
            else {
                // leastSquareErrorRigidTransformation: eight running float sums (SumXa SumYa SumXb SumYb Sxx Sxy Syx Syy)
                // in correspondence order.  All threads gather the points and evaluate the products (rounded to float
                // before they are added, as the reference's `Sxx += xa * xb` does); lanes 0..7 of warp 0 each add one
                // array in order.
                {
                    const size_t ca = (size_t)((a.cap + 3) & ~3);
                    float *tb = (float *)s.D;  // 6 arrays fit in the covariance scratch; qx / qy are free after matching
                    for (int k = t; k < nCorr; k += T) {
                        const int q = s.list[k], j = s.qj[q], i = offset + q * dec;
                        const float xa = s.m1[j].x, ya = s.m1[j].y, xb = s.m2x[i], yb = s.m2y[i];
                        tb[k] = xa; tb[ca + k] = ya; tb[2 * ca + k] = xb; tb[3 * ca + k] = yb;
                        tb[4 * ca + k] = xa * xb; tb[5 * ca + k] = xa * yb;
                        s.qx[k] = ya * xb; s.qy[k] = ya * yb;
                    }
                    __syncthreads();
                }
                if (w == 0) {
                    float acc = 0;
                    if (lane < 8) {
                        const size_t ca = (size_t)((a.cap + 3) & ~3);
                        const float *arr = lane < 6 ? (const float *)s.D + lane * ca : (lane == 6 ? s.qx : s.qy);
                        float r0 = 0;
                        int k = 0;
                        for (; k + 4 <= nCorr; k += 4) {
                            const float v0 = arr[k], v1 = arr[k + 1], v2 = arr[k + 2], v3 = arr[k + 3];
                            r0 += v0; r0 += v1; r0 += v2; r0 += v3;
                        }
                        for (; k < nCorr; k++) r0 += arr[k];
                        acc = r0;
                    }
                    const float SumXa = __shfl_sync(0xffffffffu, acc, 0), SumYa = __shfl_sync(0xffffffffu, acc, 1);
                    const float SumXb = __shfl_sync(0xffffffffu, acc, 2), SumYb = __shfl_sync(0xffffffffu, acc, 3);
                    const float Sxx = __shfl_sync(0xffffffffu, acc, 4), Sxy = __shfl_sync(0xffffffffu, acc, 5);
                    const float Syx = __shfl_sync(0xffffffffu, acc, 6), Syy = __shfl_sync(0xffffffffu, acc, 7);
                    if (lane == 0 && nCorr >= 2) {
                        const int N = nCorr;
                        const float N_inv = 1.0f / N;
                        const float mean_x_a = SumXa * N_inv, mean_y_a = SumYa * N_inv;
                        const float mean_x_b = SumXb * N_inv, mean_y_b = SumYb * N_inv;
                        const float Ax = N * (Sxx + Syy) - SumXa * SumXb - SumYa * SumYb;
                        const float Ay = SumXa * SumYb + N * (Syx - Sxy) - SumXb * SumYa;
                        const double phi = (Ax != 0 || Ay != 0) ? atan2((double)Ay, (double)Ax) : 0;
                        const double ccos = cos(phi), csin = sin(phi);
                        s.ctl[0] = mean_x_a - mean_x_b * ccos + mean_y_b * csin;
                        s.ctl[1] = mean_y_a - mean_x_b * csin - mean_y_b * ccos;
                        s.ctl[2] = phi;
                    }
                }
                __syncthreads();
                if (nCorr >= 2) { mx = s.ctl[0]; my = s.ctl[1]; mphi = s.ctl[2]; }
                __syncthreads();
                keep = true;
                if (!(fabs(lx - mx) > o.minAbsStep_trans || fabs(ly - my) > o.minAbsStep_trans ||
                      fabs(wrap_to_pi(lphi - mphi)) > o.minAbsStep_rot)) {
                    umbral_dist *= o.ALFA;
                    umbral_ang *= o.ALFA;
                    if (umbral_dist < o.smallestThresholdDist) keep = false;
                    if (o.rotate_decimation_offset == 2) offset = (offset + 1) % dec;
                }
                lx = mx; ly = my; lphi = mphi;
            }
            nIter++;
            if (o.rotate_decimation_offset == 1) offset = (offset + 1) % dec;
            if (nIter >= o.maxIterations && umbral_dist > o.smallestThresholdDist) umbral_dist *= o.ALFA;
        } while ((keep && nIter < o.maxIterations) || (nIter >= o.maxIterations && umbral_dist > o.smallestThresholdDist));
    }

    double cov[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, mse = 0;
    float quality = 0, goodness = 1;
    if (nCorr) {
        if (t == 0) {  // not part of the wrapper's interface: mean squared distance of the final correspondences
            double sum = 0;
            for (int k = 0; k < nCorr; k++) sum += s.qd[s.list[k]];
            mse = sum / nCorr;
        }
        // covariance (D D^T + 1e-6 I)^-1; D = three-point numerical derivative of the (kernelised) squared residuals
        const double ccos = cos(mphi), csin = sin(mphi);
        const double Axy = o.Axy_aprox_derivatives, rho2 = (double)o.kernel_rho * o.kernel_rho;
        for (int k = t; k < nCorr; k += T) {
            const int q = s.list[k], j = s.qj[q], i = last_offset + q * dec;
            const float tx = s.m1[j].x, ty = s.m1[j].y, ox = s.m2x[i], oy = s.m2y[i];
            const float oxt = (float)(mx + ccos * ox - csin * oy);
            const float oyt = (float)(my + csin * ox + ccos * oy);
            double w1, w2, w3, q1, q2, q3;
            w1 = oxt - Axy; w2 = oxt; w3 = oxt + Axy;
            q1 = kernelf(o.use_kernel, SQD(tx - w1) + SQD(ty - oyt), rho2);
            q2 = kernelf(o.use_kernel, SQD(tx - w2) + SQD(ty - oyt), rho2);
            q3 = kernelf(o.use_kernel, SQD(tx - w3) + SQD(ty - oyt), rho2);
            s.D[3 * k] = parab_deriv(w1, q1, w2, q2, w3, q3, oxt);
            w1 = oyt - Axy; w2 = oyt; w3 = oyt + Axy;
            q1 = kernelf(o.use_kernel, SQD(tx - oxt) + SQD(ty - w1), rho2);
            q2 = kernelf(o.use_kernel, SQD(tx - oxt) + SQD(ty - w2), rho2);
            q3 = kernelf(o.use_kernel, SQD(tx - oxt) + SQD(ty - w3), rho2);
            s.D[3 * k + 1] = parab_deriv(w1, q1, w2, q2, w3, q3, oyt);
            const double ww[3] = {mphi - Axy, mphi, mphi + Axy};
            double qq[3];
            for (int c = 0; c < 3; c++) {
                const double cx = mx + cos(ww[c]) * ox - sin(ww[c]) * oy;
                const double cy = my + sin(ww[c]) * ox + cos(ww[c]) * oy;
                qq[c] = kernelf(o.use_kernel, SQD(tx - cx) + SQD(ty - cy), rho2);
            }
            s.D[3 * k + 2] = parab_deriv(ww[0], qq[0], ww[1], qq[1], ww[2], qq[2], mphi);
        }
        __syncthreads();
        if (t < 9) {  // one thread per element of D D^T, added in correspondence order
            const int r = t / 3, c = t % 3;
            double acc = 0;
            for (int k = 0; k < nCorr; k++) acc += s.D[3 * k + r] * s.D[3 * k + c];
            if (r == c) acc += 1e-6;
            s.ctl[t] = acc;
        }
        __syncthreads();
        double A[9];
        for (int k = 0; k < 9; k++) A[k] = s.ctl[k];
        __syncthreads();
        const double det = A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) + A[2] * (A[3] * A[7] - A[4] * A[6]);
        if (det == 0 || isnan(det)) { for (int k = 0; k < 9; k++) cov[k] = nan(""); }
        else {
            cov[0] = (A[4] * A[8] - A[5] * A[7]) / det; cov[1] = (A[2] * A[7] - A[1] * A[8]) / det; cov[2] = (A[1] * A[5] - A[2] * A[4]) / det;
            cov[4] = (A[0] * A[8] - A[2] * A[6]) / det; cov[5] = (A[2] * A[3] - A[0] * A[5]) / det; cov[8] = (A[0] * A[4] - A[1] * A[3]) / det;
        }
        goodness = ratio;
        // quality: growth of the mean squared matching error under +-0.9*threshold shifts of the pose
        const float A2 = umbral_dist * 0.9f;
        double E[5];
        for (int sft = 0; sft < 5; sft++) {
            const double sx = sft == 1 ? -(double)A2 : sft == 2 ? (double)A2 : 0.0;
            const double sy = sft == 3 ? -(double)A2 : sft == 4 ? (double)A2 : 0.0;
            const Match m = compute_matching(s, N1, N2, mx + sx, my + sy, mphi, umbral_dist, umbral_ang,
                                             o.onlyClosestCorrespondences, dec, 0, gbox);
            if (t == 0) {
                double sum = 0;
                for (int k = 0; k < m.nk; k++) sum += s.qd[s.list[k]];
                s.ctl[10] = m.nk ? sum / m.nk : 0;
            }
            __syncthreads();
            E[sft] = s.ctl[10];
            __syncthreads();
        }
        double mm = E[1] - E[0];
        for (int sft = 2; sft < 5; sft++) if (E[sft] - E[0] > mm) mm = E[sft] - E[0];
        quality = (float)(-mm / (E[0] + 1e-1));
    }
    if (t == 0) {
        zhsm_result r = {};
        r.x[0] = mx; r.x[1] = my; r.x[2] = mphi;
        r.cov[0] = cov[0]; r.cov[1] = cov[1]; r.cov[2] = cov[2]; r.cov[3] = cov[4]; r.cov[4] = cov[5]; r.cov[5] = cov[8];
        r.valid = nCorr > 0;
        r.status = nCorr > 0 ? ZHSM_JOB_OK : ZHSM_JOB_TOO_FEW_POINTS;
        r.iterations = nIter;
        r.nvalid = nCorr;
        r.goodness = goodness;
        r.quality = quality;
        r.error = mse;
        a.results[job] = r;
    }
}

// Multi-start reduction: for pair p the best of its G jobs results[p*G .. p*G+G).
//   criterion 0 (zhicp):  highest goodness, then lowest error, then lowest start index
//   criterion 1 (PL-ICP): lowest error per valid correspondence (sm_result.error / nvalid), then lowest start index
// -- the order a sequential caller trying the starts one after the other would prefer them.  One thread per pair.
__global__ void best_of_starts_kernel(const zhsm_result *results, int P, int G, int criterion, int32_t *best,
                                      zhsm_result *best_results) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    int bi = -1;
    double bk0 = 0, bk1 = 0;
    for (int g = 0; g < G; g++) {
        const zhsm_result &r = results[(size_t)p * G + g];
        if (!r.valid) continue;
        const double k0 = criterion == 0 ? -(double)r.goodness : r.error / (double)(r.nvalid > 0 ? r.nvalid : 1);
        const double k1 = criterion == 0 ? r.error : 0.0;
        if (bi < 0 || k0 < bk0 || (k0 == bk0 && k1 < bk1)) { bi = g; bk0 = k0; bk1 = k1; }
    }
    if (best) best[p] = bi;
    if (best_results) {
        if (bi >= 0) best_results[p] = results[(size_t)p * G + bi];
        else { zhsm_result z = {}; z.status = ZHSM_JOB_TOO_FEW_POINTS; best_results[p] = z; }
    }
}

}  // namespace icp

size_t icp_smem_bytes(int cap) { return icp::sh_bytes(cap); }

cudaError_t icp_launch(const IcpArgs &a, cudaStream_t st) {
    const size_t smem = icp::sh_bytes(a.cap);
    cudaError_t e = cudaFuncSetAttribute(icp::icp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    icp::icp_kernel<<<a.B, icp::T, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t best_of_starts_launch(const zhsm_result *results, int P, int G, int criterion, int32_t *best,
                                  zhsm_result *best_results, cudaStream_t st) {
    icp::best_of_starts_kernel<<<(P + 127) / 128, 128, 0, st>>>(results, P, G, criterion, best, best_results);
    return cudaGetLastError();
}
