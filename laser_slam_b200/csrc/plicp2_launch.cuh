// laser_slam_b200/csrc/plicp2_launch.cuh -- launch helper shared by the per-CTA-size instance files.
#pragma once
#define PLICP2_CASE(TT, CC)                                                                                      \
    case CC: {                                                                                                   \
        size_t smem = plicp2::smem_bytes(a.L.n, TT);                                                             \
        cudaError_t e = cudaFuncSetAttribute(plicp2::plicp2_kernel<TT, CC>,                                      \
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);            \
        if (e != cudaSuccess) return e;                                                                          \
        plicp2::plicp2_kernel<TT, CC><<<a.B, TT, smem, st>>>(a);                                                 \
        break;                                                                                                   \
    }
