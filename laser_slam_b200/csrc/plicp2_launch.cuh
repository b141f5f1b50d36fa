// laser_slam_b200/csrc/plicp2_launch.cuh -- launch helper shared by the instance files: one CTA of TT threads per job, or
// (PLICP_NS > 1, the multi-start instance) one CTA of PLICP_NS * TT threads per PLICP_NS jobs that share their scans.
#pragma once
#include <cstdio>
#define PLICP2_LAUNCH(TT)                                                                                        \
    {                                                                                                            \
        size_t smem = PL_NAMESPACE::smem_bytes(a.L.n, TT);                                                       \
        smem += (size_t)(a.smem_pad & ~15); /* tuning (ZHSM_PLICP_SMEM_PAD, read in zhsm_create): fewer CTAs per SM */    \
        if (smem > 227 * 1024 || a.B % PLICP_NS != 0) return cudaErrorInvalidValue;                              \
        cudaError_t e = cudaFuncSetAttribute(PL_NAMESPACE::plicp2_kernel<TT>,                                    \
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);            \
        if (e != cudaSuccess) return e;                                                                          \
        if (a.smem_pad & 1) { /* tuning: an odd pad asks for the resident CTAs per SM on stderr */               \
            int nb = 0;                                                                                          \
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, PL_NAMESPACE::plicp2_kernel<TT>, TT * PLICP_NS, smem); \
            fprintf(stderr, "plicp2_kernel<%d>: %zu B shared, %d CTAs per SM\n", TT, smem, nb);                   \
        }                                                                                                        \
        const PL_NAMESPACE::Lay lay = PLICP_NS == 1 ? PL_NAMESPACE::make_layout(a.L.n, TT)                       \
                                                    : PL_NAMESPACE::make_layout_slots(a.L.n, TT, PLICP_NS, 0);   \
        PL_NAMESPACE::plicp2_kernel<TT><<<a.B / PLICP_NS, TT * PLICP_NS, smem, st>>>(a, lay);                    \
        return cudaGetLastError();                                                                               \
    }
