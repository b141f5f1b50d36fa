// laser_slam_b200/csrc/plicp2_launch.cuh -- launch helper shared by the instance files: one CTA of TT threads per job, or
// (PLICP_NS > 1, the multi-start instance) one CTA of PLICP_NS * TT threads per PLICP_NS jobs that share their scans.
#pragma once
#define PLICP2_LAUNCH(TT)                                                                                        \
    {                                                                                                            \
        size_t smem = PL_NAMESPACE::smem_bytes(a.L.n, TT);                                                       \
        smem += (size_t)a.smem_pad; /* tuning (ZHSM_PLICP_SMEM_PAD, read in zhsm_create): fewer CTAs per SM */    \
        if (smem > 227 * 1024 || a.B % PLICP_NS != 0) return cudaErrorInvalidValue;                              \
        cudaError_t e = cudaFuncSetAttribute(PL_NAMESPACE::plicp2_kernel<TT>,                                    \
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);            \
        if (e != cudaSuccess) return e;                                                                          \
        PL_NAMESPACE::plicp2_kernel<TT><<<a.B / PLICP_NS, TT * PLICP_NS, smem, st>>>(a);                         \
        return cudaGetLastError();                                                                               \
    }
