// laser_slam_b200/csrc/plicp_kernel2.cuh -- PL-ICP kernel entry: load, icp_loop x (1+6), result record.
#pragma once

namespace plicp {

// ld_corr_hash (laser_data.c) over the stream s.str[0..n)
__device__ inline unsigned corr_hash_seq(const int *str, int n) {
    unsigned hash = 0;
    for (unsigned i = 0; i < (unsigned)n; i++) {
        unsigned v = (unsigned)str[i];
        hash ^= ((i & 1) == 0) ? ((hash << 7) ^ v ^ (hash >> 3)) : (~((hash << 11) ^ v ^ (hash >> 5)));
    }
    return hash & 0x7FFFFFFFu;
}

template <int C>
__global__ void __launch_bounds__(256) plicp_kernel(PlicpArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int n = a.L.n;
    const int t = threadIdx.x;
    const int nt = blockDim.x;
    const int job = blockIdx.x;
    const Smem s = carve(smem_raw, n, nt);
    Ctl *ctl = s.ctl;
    const LaserDev L = a.L;
    const zhsm_plicp_params &P = a.P;

    // ---------------- load + fromPMScan2LDP + ld_valid_fields + ld_invalid_if_outside + ld_compute_cartesian
    const float *ref_row = a.scans + (size_t)a.ref_idx[job] * n;
    const float *sens_row = a.scans + (size_t)a.sens_idx[job] * n;
    int bad_field = 0, nvf_ref = 0, nvf_sens = 0;
    for (int i = t; i < n; i += nt) {
        double2 cs = a.cs[i];
        double rr = reading_of(ref_row[i], L, P, &bad_field);
        nvf_ref += (rr != -1.0);
        if (rr < 0) rr = -1.0;
        s.refr[i] = rr;
        s.refp[i] = rr >= 0 ? make_double2(cs.x * rr, cs.y * rr) : make_double2(nan(""), nan(""));
        double rs = reading_of(sens_row[i], L, P, &bad_field);
        nvf_sens += (rs != -1.0);
        s.sensp[i] = rs >= 0 ? make_double2(cs.x * rs, cs.y * rs) : make_double2(nan(""), nan(""));
    }
    if (t == 0) {
        for (int k = 0; k < 3; k++) ctl->x_old[k] = a.first_guess ? a.first_guess[3 * (size_t)job + k] : 0.0;
    }
    int tot_bad = block_sum_i(bad_field, s);
    int tot_ref = block_sum_i(nvf_ref, s);
    int tot_sens = block_sum_i(nvf_sens, s);
    zhsm_result *res = a.results + job;
    int16_t *oj1 = a.corr_j1 ? a.corr_j1 + (size_t)job * n : nullptr;
    int16_t *oj2 = a.corr_j2 ? a.corr_j2 + (size_t)job * n : nullptr;
    const bool guess_nan = isnan(ctl->x_old[0]) || isnan(ctl->x_old[1]) || isnan(ctl->x_old[2]);
    if (tot_bad || tot_ref < n * 0.10 || tot_sens < n * 0.10 || guess_nan) {
        // ld_valid_fields failed (sm_icp returns valid=0 before iterating) or icp_loop refused a NaN first guess
        if (t == 0) {
            zhsm_result r = {};
            r.status = ZHSM_JOB_INVALID_SCAN;
            r.n_loops = guess_nan && !(tot_bad || tot_ref < n * 0.10 || tot_sens < n * 0.10) ? 1 : 0;
            *res = r;
        }
        if (oj1) for (int i = t; i < n; i += nt) { oj1[i] = -1; oj2[i] = -1; }
        return;
    }
    // ld_create_jump_tables(laser_ref) (laser_data.c), one ray per thread step
    for (int i = t; i < n; i += nt) {
        double ri = s.refr[i];
        int j;
        j = i + 1; while (j < n && s.refr[j] >= 0 && s.refr[j] <= ri) j++;
        s.jt[0 * n + i] = (short)(j - i);
        j = i + 1; while (j < n && s.refr[j] >= 0 && s.refr[j] >= ri) j++;
        s.jt[1 * n + i] = (short)(j - i);
        j = i - 1; while (j >= 0 && s.refr[j] >= 0 && s.refr[j] <= ri) j--;
        s.jt[2 * n + i] = (short)(j - i);  // down_bigger
        j = i - 1; while (j >= 0 && s.refr[j] >= 0 && s.refr[j] >= ri) j--;
        s.jt[3 * n + i] = (short)(j - i);  // down_smaller
    }
    if (t == 0) { ctl->cs = cos(ctl->x_old[2]); ctl->sn = sin(ctl->x_old[2]); }
    __syncthreads();

    const double min_theta = theta_of(L, 0), max_theta = theta_of(L, n - 1);
    const double C1 = (double)n / (max_theta - min_theta);
    const double maxd2 = P.max_correspondence_dist * P.max_correspondence_dist;
    const int i0 = t * C;
    const int max_it = P.max_iterations < ZHSM_MAX_ITER ? P.max_iterations : ZHSM_MAX_ITER;
    const double fail_perc = 0.05;

    Corr<C> c;
    int guess = -1;
    // thread-0 bookkeeping of sm_icp (icp.c)
    unsigned hashes[ZHSM_MAX_ITER];
    double best_x[3] = {0, 0, 0}, x_first[3] = {0, 0, 0}, best_error = 0, loop_error = 0;
    int iterations = 0, nvalid0 = 0, loop_valid = 0, n_loops = 0, restarted = 0, valid = 0, status = ZHSM_JOB_OK;

    for (int loop = 0; loop < 7; loop++) {
        int iteration = 0;
        int all_ok = 1;
        for (iteration = 0; iteration < max_it; iteration++) {
            const double x_old[3] = {ctl->x_old[0], ctl->x_old[1], ctl->x_old[2]};
            const double cs = ctl->cs, sn = ctl->sn;
            for (int j = t; j < n; j += nt) s.d2j[j] = (unsigned long long)__double_as_longlong(1000000.0);
            find_correspondences<C>(s, L, n, i0, C1, min_theta, maxd2, x_old, cs, sn, guess, c);
            int num_corr = block_sum_i(__popc(c.valid), s);
            if (num_corr < fail_perc * n) { all_ok = 0; status = ZHSM_JOB_TOO_FEW_POINTS; break; }
            // kill_outliers_double (icp_outliers.c)
            if (P.outliers_remove_doubles) {
#pragma unroll
                for (int k = 0; k < C; k++)
                    if (c.valid >> k & 1u) atomicMin(&s.d2j[c.j1[k]], (unsigned long long)__double_as_longlong(c.d2[k]));
                __syncthreads();
#pragma unroll
                for (int k = 0; k < C; k++)
                    if (c.valid >> k & 1u) {
                        double dj = __longlong_as_double((long long)s.d2j[c.j1[k]]);
                        if (c.d2[k] > (3.0 * 3.0) * dj) c.valid &= ~(1u << k);
                    }
            }
            // kill_outliers_trim
            double dist[C];
#pragma unroll
            for (int k = 0; k < C; k++) {
                dist[k] = 0;
                if (c.valid >> k & 1u) {
                    double2 p = s.sensp[i0 + k];
                    double wx = cs * p.x - sn * p.y + x_old[0];
                    double wy = sn * p.x + cs * p.y + x_old[1];
                    double2 q1 = s.refp[c.j1[k]], q2 = s.refp[c.j2[k]];
                    dist[k] = dist_to_segment(q1.x, q1.y, q2.x, q2.y, wx, wy);
                }
            }
            int kcnt = block_sum_i(__popc(c.valid), s);
            int order = (int)floor(kcnt * P.outliers_maxPerc);
            order = order < kcnt - 1 ? order : kcnt - 1; order = order > 0 ? order : 0;
            int order2 = (int)floor(kcnt * P.outliers_adaptive_order);
            order2 = order2 < kcnt - 1 ? order2 : kcnt - 1; order2 = order2 > 0 ? order2 : 0;
            double lim1, d_o2;
            select2<C>(s, dist, c.valid, order, order2, lim1, d_o2);
            double lim2 = P.outliers_adaptive_mult * d_o2;
            double error_limit = lim1 < lim2 ? lim1 : lim2;
            // kill + per-thread partial sums: total_error, nvalid, GPC sums (compute_next_estimate + gpc_solve)
            double S[NSUM];
#pragma unroll
            for (int q = 0; q < NSUM; q++) S[q] = 0;
            int nkept = 0;
#pragma unroll
            for (int k = 0; k < C; k++) {
                int sv = -1;
                if (c.valid >> k & 1u) {
                    if (dist[k] > error_limit) {
                        c.valid &= ~(1u << k);
                        c.j1[k] = -1; c.j2[k] = -1;
                    } else {
                        nkept++;
                        S[14] += dist[k];
                        sv = c.j1[k] + 1000 * c.j2[k];
                        double2 p = s.sensp[i0 + k];
                        double2 q1 = s.refp[c.j1[k]], q2 = s.refp[c.j2[k]];
                        double px = p.x, py = p.y, qx = q1.x, qy = q1.y;
                        double diff0 = q1.x - q2.x, diff1 = q1.y - q2.y;
                        double one_on_norm = 1 / sqrt(diff0 * diff0 + diff1 * diff1);
                        double ca = +diff1 * one_on_norm, sa = -diff0 * one_on_norm;
                        double C00 = ca * ca, C01 = ca * sa, C11 = sa * sa;
                        S[0] += C00;
                        S[1] += C01;
                        S[2] += +px * C00 + py * C01;
                        S[3] += -py * C00 + px * C01;
                        S[4] += C11;
                        S[5] += +px * C01 + py * C11;
                        S[6] += +px * C11 - py * C01;
                        S[7] += (px * px) * (+C00) + (px * py) * (+C01 + C01) + (py * py) * (+C11);
                        S[8] += (px * px) * (+C01) + (px * py) * (-C00 + C11) + (py * py) * (-C01);
                        S[9] += (px * px) * (+C11) + (px * py) * (-C01 - C01) + (py * py) * (+C00);
                        S[10] += C00 * qx + C01 * qy;
                        S[11] += C01 * qx + C11 * qy;
                        S[12] += qx * (C00 * px + C01 * py) + qy * (C01 * px + C11 * py);
                        S[13] += qx * (C00 * (-py) + C01 * px) + qy * (C01 * (-py) + C11 * px);
                    }
                }
                if (i0 + k < n) s.str[i0 + k] = sv;
            }
#pragma unroll
            for (int q = 0; q < 15; q++) {
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) S[q] += __shfl_down_sync(0xffffffffu, S[q], o);
            }
            if ((t & 31) == 0) {
#pragma unroll
                for (int q = 0; q < 15; q++) s.red[(t >> 5) * NSUM + q] = S[q];
            }
            int num_after = block_sum_i(nkept, s);  // barriers also publish s.red and s.str
            if (t == 0) {
                int brk = 0;
                const int nw = (nt + 31) >> 5;
                double T[15];
                for (int q = 0; q < 15; q++) { double v = 0; for (int w = 0; w < nw; w++) v += s.red[w * NSUM + q]; T[q] = v; }
                loop_error = T[14];
                loop_valid = num_after;
                double x_new[3];
                if (num_after < fail_perc * n) { brk = 2; status = ZHSM_JOB_TOO_FEW_POINTS; }
                else if (!gpc_solve_sums(T, x_new)) { brk = 2; status = ZHSM_JOB_SINGULAR; }
                else {
                    ctl->x_new[0] = x_new[0]; ctl->x_new[1] = x_new[1]; ctl->x_new[2] = x_new[2];
                    double delta[3];
                    pose_diff(x_new, x_old, delta);
                    hashes[iteration] = corr_hash_seq(s.str, n);
                    if (P.use_point_to_line_distance) {
                        for (int q = iteration - 1; q >= 0; q--) if (hashes[q] == hashes[iteration]) { brk = 1; break; }
                    }
                    if (!brk) {
                        double da = sqrt(delta[0] * delta[0] + delta[1] * delta[1]);
                        double db = fabs(delta[2]);
                        if (da < P.epsilon_xy && db < P.epsilon_theta) brk = 1;
                    }
                    if (!brk) {
                        ctl->x_old[0] = x_new[0]; ctl->x_old[1] = x_new[1]; ctl->x_old[2] = x_new[2];
                        ctl->cs = cos(x_new[2]); ctl->sn = sin(x_new[2]);
                    }
                }
                ctl->brk = brk;
            }
            __syncthreads();
            int brk = ctl->brk;
            if (brk == 2) all_ok = 0;
            if (brk) break;
        }
        // ---- after icp_loop: sm_icp bookkeeping on thread 0 (icp.c)
        if (t == 0) {
            int my_iterations = iteration + 1;
            int next = 0;
            n_loops++;
            if (loop == 0) {
                iterations = my_iterations;
                if (!all_ok) { valid = 0; nvalid0 = 0; }
                else {
                    valid = 1; status = ZHSM_JOB_OK;
                    best_error = loop_error; nvalid0 = loop_valid;
                    for (int k = 0; k < 3; k++) { best_x[k] = ctl->x_new[k]; x_first[k] = ctl->x_new[k]; }
                    if (P.restart && (loop_error / loop_valid) > P.restart_threshold_mean_error) { restarted = 1; next = 1; }
                }
            } else {
                if (!all_ok) { next = 0; status = ZHSM_JOB_OK; }  // a failing restart breaks the restart loop
                else {
                    iterations += my_iterations;
                    if (loop_error < best_error) { best_error = loop_error; for (int k = 0; k < 3; k++) best_x[k] = ctl->x_new[k]; }
                    next = loop < 6;
                }
            }
            if (next) {
                const int aidx = loop;  // perturbation index for the next loop (loop+1)-1
                const double dt = P.restart_dt, dth = P.restart_dtheta;
                const double pert[6][3] = {{dt, 0, 0}, {-dt, 0, 0}, {0, dt, 0}, {0, -dt, 0}, {0, 0, dth}, {0, 0, -dth}};
                for (int k = 0; k < 3; k++) ctl->x_old[k] = x_first[k] + pert[aidx][k];
                ctl->cs = cos(ctl->x_old[2]); ctl->sn = sin(ctl->x_old[2]);
            }
            ctl->next = next;
            ctl->isum = valid | (restarted << 1);
        }
        __syncthreads();
        if (!ctl->next) break;
    }
    // ---------------- final correspondences + result record
    const int fin = ctl->isum;
    const bool v_ok = fin & 1, v_restarted = fin & 2;
    if (v_ok && v_restarted) {
        // "recompute correspondences in case of restarts" (icp.c): world coords at best x + tricks, no outlier pass
        __syncthreads();
        if (t == 0) {
            for (int k = 0; k < 3; k++) ctl->x_old[k] = best_x[k];
            ctl->cs = cos(best_x[2]); ctl->sn = sin(best_x[2]);
        }
        __syncthreads();
        const double x_old[3] = {ctl->x_old[0], ctl->x_old[1], ctl->x_old[2]};
        find_correspondences<C>(s, L, n, i0, C1, min_theta, maxd2, x_old, ctl->cs, ctl->sn, guess, c);
    }
#pragma unroll
    for (int k = 0; k < C; k++) {
        int i = i0 + k;
        if (i >= n) continue;
        bool cv = v_ok && (c.valid >> k & 1u);
        s.str[i] = cv ? c.j1[k] + 1000 * c.j2[k] : -1;
        if (oj1) { oj1[i] = cv ? (int16_t)c.j1[k] : (int16_t)-1; oj2[i] = cv ? (int16_t)c.j2[k] : (int16_t)-1; }
    }
    __syncthreads();
    if (t == 0) {
        zhsm_result r = {};
        r.valid = valid;
        r.status = valid ? ZHSM_JOB_OK : status;
        r.iterations = iterations;
        r.n_loops = n_loops;
        if (valid) {
            r.x[0] = best_x[0]; r.x[1] = best_x[1]; r.x[2] = best_x[2];
            r.error = best_error;
            r.nvalid = nvalid0;
            r.corr_hash = corr_hash_seq(s.str, n);
        }
        *res = r;
    }
}

}  // namespace plicp

cudaError_t plicp_launch(const PlicpArgs &a, int sm_count, cudaStream_t st, int *launched) {
    (void)sm_count;
    const int n = a.L.n;
    const int nt = 256;
    int C = (n + nt - 1) / nt;
    if (C & 1) C++;
    if (C < 2) C = 2;
    size_t smem = plicp::smem_bytes(n, nt);
    cudaError_t e;
#define PLICP_CASE(CC)                                                                                          \
    case CC:                                                                                                    \
        e = cudaFuncSetAttribute(plicp::plicp_kernel<CC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return e;                                                                         \
        plicp::plicp_kernel<CC><<<a.B, nt, smem, st>>>(a);                                                      \
        break;
    switch (C) {
        PLICP_CASE(2)
        PLICP_CASE(4)
        PLICP_CASE(6)
        PLICP_CASE(8)
        default:
            return cudaErrorInvalidValue;
    }
#undef PLICP_CASE
    if (launched) *launched += 1;
    return cudaGetLastError();
}
