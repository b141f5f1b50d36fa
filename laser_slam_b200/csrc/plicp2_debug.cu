// laser_slam_b200/csrc/plicp2_debug.cu -- test hooks that run device-only building blocks of the PL-ICP kernel in
// isolation (tests/ compare them with numpy): the warp-cooperative greatest-real-root solver of gpc_solve.
#include "plicp2_solve.cuh"
#include "zhsm_internal.h"

__global__ void quartic_root_kernel(const double *coef, int B, double *root, int *ok) {
    const int w = blockIdx.x;
    if (w >= B) return;
    double a[5];
    for (int i = 0; i < 5; i++) a[i] = coef[5 * (size_t)w + i];
    double r = 0;
    const bool good = plicp2::greatest_real_root4_warp(a, &r);
    if (threadIdx.x == 0) { root[w] = r; ok[w] = good ? 1 : 0; }
}

// coef: [B][5] host doubles a0..a4; root/ok: [B] host outputs
extern "C" zhsm_status zhsm_debug_quartic_roots(zhsm_handle *h, const double *coef, int B, double *root, int *ok) {
    if (!h || !coef || !root || !ok || B <= 0) return ZHSM_ERR_ARG;
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    double *d_c = nullptr, *d_r = nullptr;
    int *d_ok = nullptr;
    ZHSM_CUDA_TRY(h, cudaMalloc(&d_c, sizeof(double) * 5 * B));
    ZHSM_CUDA_TRY(h, cudaMalloc(&d_r, sizeof(double) * B));
    ZHSM_CUDA_TRY(h, cudaMalloc(&d_ok, sizeof(int) * B));
    ZHSM_CUDA_TRY(h, cudaMemcpy(d_c, coef, sizeof(double) * 5 * B, cudaMemcpyHostToDevice));
    quartic_root_kernel<<<B, 32, 0, h->stream>>>(d_c, B, d_r, d_ok);
    h->launches++;
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    ZHSM_CUDA_TRY(h, cudaMemcpy(root, d_r, sizeof(double) * B, cudaMemcpyDeviceToHost));
    ZHSM_CUDA_TRY(h, cudaMemcpy(ok, d_ok, sizeof(int) * B, cudaMemcpyDeviceToHost));
    cudaFree(d_c); cudaFree(d_r); cudaFree(d_ok);
    return ZHSM_OK;
}
