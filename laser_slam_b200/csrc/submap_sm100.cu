// laser_slam_b200/csrc/submap_sm100.cu -- SURVEY 8f-3: the submap cloud builder that creates zhicp's inputs.
// CMapNode::computePMAP (src/mapGraph/MapNode.cpp:726-757) = translate2GlobalFrame (:387-460) of every scan of the node,
// computeBoundary (:714-725), then CPMap (src/mapGraph/PMap.cpp): updateMap (:47-95) traces every beam through a 5 cm
// grid with the Bresenham walk of line.cpp (grid_line_core :4-104) counting hits and visits, computeProbsOfMap
// (:114-131), getPointCloud (:133-142) = the cells whose hit/visit ratio is exactly 1, in row-major order.
// Rasterisation as plain CUDA: one thread per beam, integer atomics on the two count grids (order-free), then an
// ordered two-level compaction so that the cloud comes out in the reference's cell order (the ICP's float sums depend
// on point order).
//
// Two behaviours of the reference that are reproduced, not corrected: grid_line()'s "reverse if the line does not start
// at `start`" never triggers (grid_line_core clears the vector, so its size() is 0, line.cpp:112-120), hence for beams
// enumerated end->start the extra "last cell" hit lands on the ROBOT's cell; and the end cell is counted as a hit once
// directly and once more as the last cell of forward-enumerated lines.
#include <float.h>
#include <math.h>
#include <string.h>

#include "zh_libm.cuh"
#include "zhsm_internal.h"

namespace submap {

__device__ __forceinline__ unsigned okey(float v) {
    const unsigned u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ inline float oval(unsigned k) {
    unsigned u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

// translate2GlobalFrame + computeBoundary.  bounds: [0] min x, [1] min y as ordered keys (atomicMin), [2] max x, [3] max y
__global__ void points_kernel(const float *scans, const double *traj, int K, int n, const float *co, const float *si,
                              int max_range_cm, float *gx, float *gy, unsigned *bounds) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= K * n) return;
    const int k = g / n, i = g % n;
    const float b = scans[g];
    if (b > (float)max_range_cm) { gx[g] = nanf(""); return; }
    const double rx = traj[3 * k], ry = traj[3 * k + 1];
    const float th = (float)traj[3 * k + 2];
    const float fcos = zhm::cosf_glibc(th), fsin = zhm::sinf_glibc(th);
    const float x = (float)(((double)b / 100.0) * (double)co[i]), y = (float)(((double)b / 100.0) * (double)si[i]);
    const float tx = (float)((double)(fcos * x - fsin * y) + rx), ty = (float)((double)(fsin * x + fcos * y) + ry);
    gx[g] = tx;
    gy[g] = ty;
    atomicMin(&bounds[0], okey(tx)); atomicMin(&bounds[1], okey(ty));
    atomicMax(&bounds[2], okey(tx)); atomicMax(&bounds[3], okey(ty));
}

struct Grid {
    int size_x, size_y, center_x, center_y;
};

// CPMap::mapPosfromPoint (PMap.cpp:96-112) with offset 0, resolution 0.05, little_tail 0.0005f
__device__ __forceinline__ int map_pos(double p, int center) {
    const float little_tail = 0.0005f, offset = 0.f;
    const double v = (p - (double)offset + (double)little_tail) / 0.05;
    return p < (double)offset ? (int)((double)center + v - 1) : (int)((double)center + v);
}

__device__ __forceinline__ void visit(const Grid &G, int *hit, int *sum, int x, int y, bool last) {
    if (x >= 0 && x < G.size_x && y >= 0 && y < G.size_y) {
        if (last) atomicAdd(&hit[(size_t)x * G.size_y + y], 1);
        atomicAdd(&sum[(size_t)x * G.size_y + y], 1);
    }
}

// CPMap::updateMap for one beam per thread
__global__ void raycast_kernel(const float *gx, const float *gy, const double *traj, int K, int n, Grid G, int *hit, int *sum) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= K * n) return;
    const float ex_f = gx[g];
    if (isnan(ex_f)) return;
    const int k = g / n;
    const float rx = (float)traj[3 * k], ry = (float)traj[3 * k + 1];
    const int sx = map_pos((double)rx, G.center_x), sy = map_pos((double)ry, G.center_y);
    const int ex = map_pos((double)ex_f, G.center_x), ey = map_pos((double)gy[g], G.center_y);
    if (ex >= 0 && ex < G.size_x && ey >= 0 && ey < G.size_y) atomicAdd(&hit[(size_t)ex * G.size_y + ey], 1);
    // grid_line_core (line.cpp:4-104); cells are visited in the order the reference stores them, the last one (index
    // cnt-1, "j > cnt-2") also counts as a hit
    const int dx = abs(ex - sx), dy = abs(ey - sy);
    if (dy <= dx) {
        int d = 2 * dy - dx;
        const int incr1 = 2 * dy, incr2 = 2 * (dy - dx);
        int x, y, xend, ydirflag;
        if (sx > ex) { x = ex; y = ey; ydirflag = -1; xend = sx; } else { x = sx; y = sy; ydirflag = 1; xend = ex; }
        const int ystep = ((ey - sy) * ydirflag) > 0 ? 1 : -1;
        visit(G, hit, sum, x, y, x >= xend);
        while (x < xend) {
            x++;
            if (d < 0) d += incr1; else { y += ystep; d += incr2; }
            visit(G, hit, sum, x, y, x >= xend);
        }
    } else {
        int d = 2 * dx - dy;
        const int incr1 = 2 * dx, incr2 = 2 * (dx - dy);
        int x, y, yend, xdirflag;
        if (sy > ey) { y = ey; x = ex; yend = sy; xdirflag = -1; } else { y = sy; x = sx; yend = ey; xdirflag = 1; }
        const int xstep = ((ex - sx) * xdirflag) > 0 ? 1 : -1;
        visit(G, hit, sum, x, y, y >= yend);
        while (y < yend) {
            y++;
            if (d < 0) d += incr1; else { x += xstep; d += incr2; }
            visit(G, hit, sum, x, y, y >= yend);
        }
    }
}

// occupied = probability exactly 1 (computeProbsOfMap + getPointCloud): visits > 0 and hits >= visits
__device__ __forceinline__ bool occupied(const int *hit, const int *sum, size_t c) { return sum[c] > 0 && hit[c] >= sum[c]; }

__global__ void rowcount_kernel(const int *hit, const int *sum, Grid G, int *rowcnt) {
    const int x = blockIdx.x;
    int c = 0;
    for (int y = threadIdx.x; y < G.size_y; y += blockDim.x) c += occupied(hit, sum, (size_t)x * G.size_y + y);
    c = __reduce_add_sync(0xffffffffu, c);
    __shared__ int w[8];
    if ((threadIdx.x & 31) == 0) w[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) { int t = 0; for (int k = 0; k < (int)(blockDim.x >> 5); k++) t += w[k]; rowcnt[x] = t; }
}

__global__ void rowscan_kernel(int *rowcnt, int size_x, int *total) {  // exclusive scan, one thread: size_x is a few thousand
    if (threadIdx.x || blockIdx.x) return;
    int run = 0;
    for (int x = 0; x < size_x; x++) { const int c = rowcnt[x]; rowcnt[x] = run; run += c; }
    *total = run;
}

__global__ void emit_kernel(const int *hit, const int *sum, Grid G, const int *rowoff, float *out_x, float *out_y, int cap) {
    const int x = blockIdx.x, lane = threadIdx.x;  // one warp per row: ordered by ballots
    int base = rowoff[x];
    for (int y0 = 0; y0 < G.size_y; y0 += 32) {
        const int y = y0 + lane;
        const bool occ = y < G.size_y && occupied(hit, sum, (size_t)x * G.size_y + y);
        const unsigned m = __ballot_sync(0xffffffffu, occ);
        if (occ) {
            const int p = base + __popc(m & ((1u << lane) - 1));
            if (p < cap) {
                out_x[p] = (float)((double)(x - G.center_x) * 0.05);
                out_y[p] = (float)((double)(y - G.center_y) * 0.05);
            }
        }
        base += __popc(m);
    }
}

}  // namespace submap

extern "C" zhsm_status zhsm_submap_cloud(zhsm_handle *h, const float *scans, const double *traj, int K, float *out_x,
                                         float *out_y, int cap, int *count, int *dims) {
    if (!h) return ZHSM_ERR_ARG;
    if (!scans || !traj || !out_x || !out_y || !count || K <= 0 || cap < 0) { h->err = "zhsm_submap_cloud: bad argument"; return ZHSM_ERR_ARG; }
    if (!h->have_laser) { h->err = "zhsm_set_laser not called"; return ZHSM_ERR_ARG; }
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    StreamScope scope(h, st);
    const int n = h->laser.n;
    const size_t tot = (size_t)K * n;
    void *p;
    zhsm_status s;
    if ((s = zhsm_stage(h, 5, tot * 4, &p)) != ZHSM_OK) return s;
    float *d_scans = (float *)p;
    if ((s = zhsm_stage(h, 6, sizeof(double) * 3 * K, &p)) != ZHSM_OK) return s;
    double *d_traj = (double *)p;
    if ((s = zhsm_stage(h, 7, tot * 4, &p)) != ZHSM_OK) return s;
    float *d_gx = (float *)p;
    if ((s = zhsm_stage(h, 8, tot * 4, &p)) != ZHSM_OK) return s;
    float *d_gy = (float *)p;
    if ((s = zhsm_stage(h, 9, 64, &p)) != ZHSM_OK) return s;
    unsigned *d_bounds = (unsigned *)p;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(d_scans, scans, tot * 4, cudaMemcpyHostToDevice, st));
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(d_traj, traj, sizeof(double) * 3 * K, cudaMemcpyHostToDevice, st));
    const unsigned init[4] = {0xffffffffu, 0xffffffffu, 0u, 0u};
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(d_bounds, init, sizeof(init), cudaMemcpyHostToDevice, st));
    submap::points_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(d_scans, d_traj, K, n, h->d_co, h->d_si,
                                                                        h->laser.max_range_cm, d_gx, d_gy, d_bounds);
    ZHSM_CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    unsigned bk[4];
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(bk, d_bounds, sizeof(bk), cudaMemcpyDeviceToHost, st));
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    *count = 0;
    if (dims) dims[0] = dims[1] = 0;
    if (bk[0] == 0xffffffffu) return ZHSM_OK;  // no beam in range: the reference would build an empty map
    // computeBoundary starts from +-MAX_VAL = 10000 (MapNode.cpp:716), then computePMAP's size arithmetic (:735-736)
    float min_x = fminf(10000.f, submap::oval(bk[0])), min_y = fminf(10000.f, submap::oval(bk[1]));
    float max_x = fmaxf(-10000.f, submap::oval(bk[2])), max_y = fmaxf(-10000.f, submap::oval(bk[3]));
    submap::Grid G;
    G.size_x = (int)((((double)(max_x - min_x) + 0.05) / 0.05) * 2);
    G.size_y = (int)((((double)(max_y - min_y) + 0.05) / 0.05) * 2);
    G.center_x = G.size_x / 2;
    G.center_y = G.size_y / 2;
    if (dims) { dims[0] = G.size_x; dims[1] = G.size_y; }
    if (G.size_x <= 0 || G.size_y <= 0) return ZHSM_OK;
    const size_t cells = (size_t)G.size_x * G.size_y;
    if (cells > ((size_t)1 << 28)) { h->err = "zhsm_submap_cloud: map larger than 2^28 cells"; return ZHSM_ERR_UNSUPPORTED; }
    if ((s = zhsm_stage(h, 10, cells * 4, &p)) != ZHSM_OK) return s;
    int *d_hit = (int *)p;
    if ((s = zhsm_stage(h, 11, cells * 4, &p)) != ZHSM_OK) return s;
    int *d_sum = (int *)p;
    if ((s = zhsm_stage(h, 12, sizeof(int) * ((size_t)G.size_x + 1), &p)) != ZHSM_OK) return s;
    int *d_row = (int *)p;
    const size_t ocap = cap > 0 ? (size_t)cap : 1;
    if ((s = zhsm_stage(h, 13, ocap * 4, &p)) != ZHSM_OK) return s;
    float *d_ox = (float *)p;
    if ((s = zhsm_stage(h, 14, ocap * 4, &p)) != ZHSM_OK) return s;
    float *d_oy = (float *)p;
    ZHSM_CUDA_TRY(h, cudaMemsetAsync(d_hit, 0, cells * 4, st));
    ZHSM_CUDA_TRY(h, cudaMemsetAsync(d_sum, 0, cells * 4, st));
    submap::raycast_kernel<<<(unsigned)((tot + 127) / 128), 128, 0, st>>>(d_gx, d_gy, d_traj, K, n, G, d_hit, d_sum);
    submap::rowcount_kernel<<<G.size_x, 128, 0, st>>>(d_hit, d_sum, G, d_row);
    submap::rowscan_kernel<<<1, 32, 0, st>>>(d_row, G.size_x, d_row + G.size_x);
    submap::emit_kernel<<<G.size_x, 32, 0, st>>>(d_hit, d_sum, G, d_row, d_ox, d_oy, cap);
    ZHSM_CUDA_TRY(h, cudaGetLastError());
    h->launches += 4;
    int total = 0;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(&total, d_row + G.size_x, sizeof(int), cudaMemcpyDeviceToHost, st));
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    *count = total;
    const int ncopy = total < cap ? total : cap;
    if (ncopy > 0) {
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(out_x, d_ox, sizeof(float) * ncopy, cudaMemcpyDeviceToHost, st));
        ZHSM_CUDA_TRY(h, cudaMemcpyAsync(out_y, d_oy, sizeof(float) * ncopy, cudaMemcpyDeviceToHost, st));
        ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    }
    return ZHSM_OK;
}
