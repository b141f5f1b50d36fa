/* oracle/graph_oracle.h -- TEST INFRASTRUCTURE ONLY.  See graph_oracle.c. */
#ifndef ORC_GRAPH_ORACLE_H
#define ORC_GRAPH_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#define ORC_GATE_OVERLAPPED 1 /* CMapGraph::isOverlappedArea */
#define ORC_GATE_LOOPY 2      /* CMapGraph::isLoopyArea */

/* flags[m*N + j] for cur = cur_idx[m], ref = j.  cov_xy = per node (m_root_Cov(0,0), m_root_Cov(1,1)) or NULL for the
 * constant-covariance mode (m_bUseConstantCov, the reference's default). */
void orc_gate_candidates(const double *pose, const double *ld, const double *ru, const double *cov_xy, int N,
                         const int32_t *cur_idx, int M, float beta, uint8_t *flags);
/* CMapGraph::isAvaliableArea for M (node, trans) pairs */
void orc_available_area(const double *trans, const double *cov_xy, int M, uint8_t *ok);
/* deleteMisMatch + calculateWeightedPose + calculateFinalPose's choice, for S matched sets stored back to back:
 * set s = entries [set_off[s], set_off[s+1]).  keep[e] in {0,1}; final_pose[s] = weighted pose of the kept entries;
 * final_idx[s] = entry (absolute index) closest to it, -1 for an empty set. */
void orc_consensus(const double *poses, const float *good, const int32_t *set_off, int S, uint8_t *keep,
                   double *final_pose, int32_t *final_idx);
#ifdef __cplusplus
}
#endif
#endif
