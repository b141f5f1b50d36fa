/* oracle/icp_oracle.h -- TEST INFRASTRUCTURE ONLY.  See icp_oracle.c.  PARITY UNPINNED (MRPT 0.9.6 absent). */
#ifndef ORC_ICP_ORACLE_H
#define ORC_ICP_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int maxIterations;           /* 100   ZHIcp_Warpper.cpp:46 */
    float thresholdAng;          /* 10deg ZHIcp_Warpper.cpp:47 */
    float thresholdDist;         /* 0.75  ZHIcp_Warpper.cpp:48 */
    float ALFA;                  /* 0.5   ZHIcp_Warpper.cpp:49 */
    float smallestThresholdDist; /* 0.05  ZHIcp_Warpper.cpp:50 */
    int onlyClosestCorrespondences; /* MRPT default true */
    float minAbsStep_trans, minAbsStep_rot; /* MRPT default 1e-6 */
    /* MRPT's TConfigParams::corresponding_points_decimation (default 5): only every k-th point of the current cloud
     * is matched; goodness = matched / (N2 / k) (documented as "set to 1 for the MRPT < 0.9.5 behaviour").
     * rotate_decimation_offset: 0 = always start at point 0; 1 = advance the start every iteration; 2 (default) =
     * advance it each time the thresholds are scaled by ALFA, which is how MRPT's CICP uses offset_other_map_points.
     * Both are parameters because the pinned release is absent. */
    int corresponding_points_decimation;
    int rotate_decimation_offset;
    float kernel_rho;            /* 0.07 */
    int use_kernel;              /* true */
    float Axy_aprox_derivatives; /* 0.05 */
} orc_icp_options;

typedef struct {
    double pose[3];
    double cov[9];
    int nIterations;
    int nCorrespondences;
    float goodness;
    float quality;
} orc_icp_result;

void orc_icp_default_options(orc_icp_options *o);

/* m1 = reference cloud (the map), m2 = current cloud; result = pose of m2's frame in m1's frame. */
void orc_mrpt_icp(const float *m1x, const float *m1y, int N1, const float *m2x, const float *m2y, int N2,
                  const double init[3], const orc_icp_options *opt_or_null, orc_icp_result *res);

/* clouds packed as x[0..N) y[0..N) at float offset 2*off[c]; job k matches clouds (ref_idx[k], cur_idx[k]). */
void orc_mrpt_icp_batch(const float *pts, const int *off, const int *cnt, const int *ref_idx, const int *cur_idx,
                        const double *init, int njobs, const orc_icp_options *opt_or_null, orc_icp_result *out,
                        int nthreads);

#ifdef __cplusplus
}
#endif
#endif
