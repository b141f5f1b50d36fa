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
/* matchNodePairICP's hand-over (src/mapGraph/MapNode.cpp:638-651): symmetric 3x3 from cov6 = xx xy xt yy yt tt
 * ((0,0), (1,1) times scale), information = covariance.inverse() by Eigen's 3x3 cofactor formula; const_diag != NULL:
 * the m_bUseConstantCov override (src/mapGraph/MapGraph.cpp:250-261,1427-1429).  info = row-major 3x3 per record. */
void orc_cov_to_information(const double *cov6, int M, double scale, const double *const_diag, double *info);
#ifdef __cplusplus
}
#endif
#endif
