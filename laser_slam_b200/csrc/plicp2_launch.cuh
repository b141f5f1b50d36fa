// laser_slam_b200/csrc/plicp2_launch.cuh -- launch helper shared by the per-CTA-size instance files.
#pragma once
#define PLICP2_LAUNCH(TT)                                                                                        \
    {                                                                                                            \
        size_t smem = plicp2::smem_bytes(a.L.n, TT);                                                             \
        smem += (size_t)a.smem_pad; /* tuning (ZHSM_PLICP_SMEM_PAD, read in zhsm_create): fewer CTAs per SM */    \
        cudaError_t e = cudaFuncSetAttribute(plicp2::plicp2_kernel<TT>,                                          \
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);            \
        if (e != cudaSuccess) return e;                                                                          \
        plicp2::plicp2_kernel<TT><<<a.B, TT, smem, st>>>(a);                                                     \
        return cudaGetLastError();                                                                               \
    }
