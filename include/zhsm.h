/*
 * include/zhsm.h -- C ABI of the B200 batched 2D scan-matching engine (libzhsm.so).
 *
 * Drop-in boundary for the reference's three matchers.  Every entry point is
 * plain C: pointers + sizes, no C++/torch types.  Status codes replace the reference's `throw 1`
 * (src/zhpsm/ZHPolar_Match.cpp:1095,1106,1239,1831) and `output->valid==0`
 * (src/zhcsm/ZHCanonical_Matcher.cpp:275-283); nothing aborts a batch.
 *
 * Scans live in a "scan pool": [S][n] float32 ranges in cm, row-major, exactly PMScan::r
 * (src/zhpsm/PolarParameter.h:105-184).  A batch is a list of JOBS (ref scan index, sens scan index,
 * optional first guess) so consecutive-scan odometry, many-to-one loop-closure candidates and
 * multi-start share one entry point and each scan is read from HBM once per job.
 *
 * There is NO CPU fallback: every compute entry returns ZHSM_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef ZHSM_H
#define ZHSM_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct zhsm_handle zhsm_handle;

typedef enum {
    ZHSM_OK = 0,
    ZHSM_ERR_ARG = 1,
    ZHSM_ERR_CUDA = 2,
    ZHSM_ERR_NOMEM = 3,
    ZHSM_ERR_UNSUPPORTED = 4
} zhsm_status;

/* per-job status in zhsm_result.status */
enum {
    ZHSM_JOB_OK = 0,
    ZHSM_JOB_TOO_FEW_POINTS = 1, /* pm_icp/pm_translation_estimation "not enough points"; icp_loop <5% correspondences */
    ZHSM_JOB_SINGULAR = 2,       /* determinant too small / gpc_solve failed */
    ZHSM_JOB_SEARCH_FAILED = 3,  /* pm_orientation_search failed */
    ZHSM_JOB_INVALID_SCAN = 4    /* ld_valid_fields rejected a scan (sm_icp returns valid=0 before iterating) */
};

/* POD image of Base_PARAM (src/zhpsm/PolarParameter.h:42-83) */
typedef struct zhsm_laser {
    int32_t n;                /* pm_l_points */
    float fi_min;             /* pm_fi_min [rad] */
    float dfi;                /* pm_dfi [rad] */
    int32_t min_range_cm;     /* pm_min_range (PM_MIN_RANGE = 10) */
    int32_t max_range_cm;     /* pm_max_range */
    int32_t min_valid_points; /* pm_min_valid_points */
    int32_t scan_window;      /* pm_scan_window */
    int32_t flags;            /* ZHSM_LASER_* */
} zhsm_laser;
/* pm_orientation_search's branch for pm_laser_name == "HOKUYO_UTM_30LX" (ZHPolar_Match.cpp:1182-1191): a beam only
 * counts towards a shift's error when |new_r - ref_r| < PM_MAX_ERROR (100 cm).  The reference keys it on the sensor's
 * name; here it is a flag of the geometry so that any Hokuyo-class laser can ask for it. */
#define ZHSM_LASER_UTM_GATE 1

/* The sm_params fields CCanonicalMatcher::sm_init sets (src/zhcsm/ZHCanonical_Matcher.cpp:83-157).
 * zhsm_plicp_default_params() fills the reference's values. */
typedef struct zhsm_plicp_params {
    double max_angular_correction_deg;
    double max_linear_correction;
    double epsilon_xy;
    double epsilon_theta;
    double max_correspondence_dist;
    double sigma;
    double restart_threshold_mean_error;
    double restart_dt;
    double restart_dtheta;
    double outliers_maxPerc;
    double outliers_adaptive_order;
    double outliers_adaptive_mult;
    double min_reading; /* m */
    double max_reading; /* m */
    int32_t max_iterations;
    int32_t use_corr_tricks;          /* only 1 (the reference's setting) is implemented */
    int32_t restart;
    int32_t use_point_to_line_distance; /* only 1 */
    int32_t outliers_remove_doubles;
    int32_t do_compute_covariance;
    int32_t reserved[2];
} zhsm_plicp_params;

/* Fixed 128-byte result record; also the NCCL all-gather record. */
typedef struct zhsm_result {
    double x[3];        /* PL-ICP / ICP: m,m,rad.  PSM family: cm,cm,rad (PMScan units) */
    double cov[6];      /* upper triangle xx xy xt yy yt tt */
    double error;       /* sm_result.error / returned avg error */
    int32_t valid;      /* sm_result.valid */
    int32_t iterations; /* sm_result.iterations (sum over restarts) */
    int32_t nvalid;     /* sm_result.nvalid */
    int32_t status;     /* ZHSM_JOB_* */
    float goodness;
    float quality;
    uint32_t corr_hash; /* ld_corr_hash of the final correspondences */
    int32_t n_loops;    /* icp_loop executions (1 + restarts) */
    int32_t reserved[4];
} zhsm_result;

/* flags for batch calls */
#define ZHSM_SCANS_ON_DEVICE 1u   /* scan pool pointer is device memory */
#define ZHSM_JOBS_ON_DEVICE 2u    /* ref_idx/sens_idx/first_guess pointers are device memory */
#define ZHSM_RESULTS_ON_DEVICE 4u /* results (and corr outputs) are device memory */
#define ZHSM_ALL_ON_DEVICE 7u
/* zhsm_plicp_batch only: median-filter the whole scan pool first (the device copy when the pool is a host
 * buffer), i.e. the CCanonicalMatcher::FMatch(PMScan*,PMScan*,bFirst=true) flow (:229-249) in one call.
 * With ZHSM_WRITEBACK_SCANS a host pool receives the filtered ranges back, like FMatch mutates lsa->r. */
#define ZHSM_FILTER_SCANS 8u
#define ZHSM_WRITEBACK_SCANS 16u
/* zhsm_plicp_batch only, multi-start batches (loop-closure candidates with G first guesses per scan pair, the order
 * zhsm_best_of_starts reads: pair-major): the caller promises that B is a multiple of 4 and that jobs 4k .. 4k+3 share
 * both scan indices.  Long scans (> 1472 beams) are then matched four jobs per CTA over ONE staged copy of the two
 * scans (twice the jobs per SM); shorter scans ignore the flag.  The same records either way: every integer field and
 * every correspondence equal, poses equal to rounding (1e-12).  A job that breaks the promise is failed with
 * ZHSM_JOB_INVALID_SCAN, never matched against the wrong scans.  MEASURED SLOWER than the default path on B200
 * (68 k against 90 k matches/s at 2048 beams): leave it off unless your own measurement says otherwise. */
#define ZHSM_MULTISTART 32u

zhsm_status zhsm_create(int device, zhsm_handle **out);
void zhsm_destroy(zhsm_handle *h);
const char *zhsm_last_error(zhsm_handle *h);

/* Sets the sensor geometry; builds the device-side tables (double cos/sin of the beam angles computed on the
 * host with libm so Cartesian points are bit-identical to ld_compute_cartesian, and the GF(2) matrices that
 * let ld_corr_hash be evaluated in parallel). */
zhsm_status zhsm_set_laser(zhsm_handle *h, const zhsm_laser *laser);

void zhsm_plicp_default_params(zhsm_plicp_params *p, int pm_min_range_cm, int pm_max_range_cm);

/* CCanonicalMatcher::median_filter (ZHCanonical_Matcher.cpp:159-189) == CPolarMatch::pm_median_filter
 * (ZHPolar_Match.cpp:1610-1639), in place over S scans. */
zhsm_status zhsm_median_filter_batch(zhsm_handle *h, float *scans, int S, unsigned flags, void *cuda_stream);

/* Batched CCanonicalMatcher::FMatch(LDP,LDP) (ZHCanonical_Matcher.cpp:251-285): fromPMScan2LDP on both scans
 * (:193-227) + sm_icp (:274).  Scans must already be median filtered (the wrapper filters each scan once when
 * it arrives, :233-237).  first_guess may be NULL (= 0,0,0 as :270-272) or [B][3] doubles (m,m,rad).
 * corr_j1/corr_j2 may be NULL or [B][n] int16 (-1 = no correspondence).
 * sm_icp's restart loop (params->restart: six perturbed icp_loops when the mean error is above
 * restart_threshold_mean_error) runs either inside the job, one loop after the other, or as a second batch of one-loop
 * jobs -- always for batches of 512 to 4096 jobs, for larger ones when the handle's previous batch restarted on 2 % of
 * its jobs or more, never for smaller ones.  Same records either way, up to the diagnostics in reserved[].  The second form needs the number of
 * restarting jobs on the host: the call then BLOCKS until the first pass is done, also with the *_ON_DEVICE flags
 * (what it queues after that is asynchronous as before; device-resident job arrays are not read after the call
 * returns, the scan pool and the result buffers are, as ever, until the stream has run). */
zhsm_status zhsm_plicp_batch(zhsm_handle *h, const float *scans, int S, const int32_t *ref_idx,
                             const int32_t *sens_idx, const double *first_guess, int B,
                             const zhsm_plicp_params *params, zhsm_result *results, int16_t *corr_j1,
                             int16_t *corr_j2, unsigned flags, void *cuda_stream);

/* zhsm_plicp_batch for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.
 * Same kernels and same records as zhsm_plicp_batch, but it allocates nothing: every buffer is the CALLER'S DEVICE
 * memory (d_scans [S][n] float cm, already median filtered; d_ref_idx / d_sens_idx [B]; d_first_guess [B][3] or NULL;
 * d_results [B]; d_corr_j1 / d_corr_j2 [B][n] int16 or both NULL), and the kernels' scratch (jump tables, 8 bytes per
 * ray and job, + a per-scan owner table) lives in one static device array of fixed size that every GPU loading the
 * library holds until exit (8 MiB + 32 KiB).  Limits: B * beams <= ZHSM_NOALLOC_MAX_RAYS (970 jobs at 1080 beams, 512
 * at 2048, 5793 at 181) and S <= ZHSM_NOALLOC_MAX_SCANS; beyond them the call fails with ZHSM_ERR_ARG and says so in
 * zhsm_last_error.  zhsm_create / zhsm_set_laser (set-up, they do allocate) must have run; the first call on a device
 * also resolves the array's address once (cudaGetSymbolAddress).  Asynchronous on cuda_stream (NULL = the handle's):
 * the records are ready after zhsm_sync.  All calls on a device share the array, which keeps the last call's
 * contents (nothing in it needs clearing: each launch re-initialises what it reads); calls that OVERLAP on the GPU --
 * two streams, or two handles on one device -- race on it: keep them on one stream or order them with events. */
#define ZHSM_NOALLOC_MAX_RAYS (1u << 20)
#define ZHSM_NOALLOC_MAX_SCANS 8192
zhsm_status zhsm_plicp_batch_noalloc(zhsm_handle *h, const float *d_scans, int S, const int32_t *d_ref_idx,
                                     const int32_t *d_sens_idx, const double *d_first_guess, int B,
                                     const zhsm_plicp_params *params, zhsm_result *d_results, int16_t *d_corr_j1,
                                     int16_t *d_corr_j2, void *cuda_stream);

/* ---- zhpsm: polar scan matching and pm_icp (src/zhpsm/ZHPolar_Match.cpp) ---------------------------------------
 * Scans are preprocessed once (zhsm_preprocess_batch) and then matched any number of times.  bad holds the
 * reference's PM_RANGE / PM_OCCLUDED / PM_EMPTY bit flags (PolarParameter.h:5-9) as uint8, seg the segment numbers of
 * pm_segment_scan as int16.  Poses are PMScan units: cm, cm, rad.  The ref scan of a job sits at the origin, which is
 * how every caller uses the matchers (CPolarMatch::FMatch zeroes both poses, ZHPolar_Match.cpp:97-99). */
enum {
    ZHSM_ALGO_PSM = 0,     /* CPolarMatch::pm_psm (:890-1003) */
    ZHSM_ALGO_PM_ICP = 1,  /* CPolarMatch::pm_icp (:1653-2021) */
    ZHSM_ALGO_PSM_ICP = 2, /* pm_psm, then pm_icp seeded with its result (runlog :561-563) */
    ZHSM_ALGO_FMATCH = 3   /* CPolarMatch::FMatch (:81-121) on preprocessed scans: zero poses, pm_icp, pm_error_index2 */
};

/* pm_preprocessScan (:861-866) over S scans: pm_median_filter in place, pm_find_far_points (bad |= PM_RANGE),
 * pm_segment_scan.  bad is in/out (pass zeros for fresh scans).  ZHSM_SCANS_ON_DEVICE covers all three arrays. */
zhsm_status zhsm_preprocess_batch(zhsm_handle *h, float *scans, uint8_t *bad, int16_t *seg, int S, unsigned flags,
                                  void *cuda_stream);

/* Batched matcher.  init_pose: [B][3] float act poses or NULL (zeros).  Per job: results[k].x = pose (cm, cm, rad),
 * .error = the matcher's return value, .iterations, .valid / .status (the reference's `throw 1` becomes
 * ZHSM_JOB_TOO_FEW_POINTS / SINGULAR / SEARCH_FAILED and leaves the pose at init_pose, like the reference leaves
 * lsa untouched).  ZHSM_ALGO_FMATCH also fills .cov[0], .cov[3] with the pair FMatch returns (avg_err_x, avg_err_y in
 * m^2; -1,-1 on failure) and .nvalid with pm_error_index2's associated points.
 * flags: ZHSM_SCANS_ON_DEVICE (scans, bad, seg), ZHSM_JOBS_ON_DEVICE (ref_idx, act_idx, init_pose),
 * ZHSM_RESULTS_ON_DEVICE. */
zhsm_status zhsm_psm_batch(zhsm_handle *h, const float *scans, const uint8_t *bad, const int16_t *seg, int S,
                           const int32_t *ref_idx, const int32_t *act_idx, const float *init_pose, int B, int algo,
                           zhsm_result *results, unsigned flags, void *cuda_stream);

/* pm_scan_project (:1356-1479) of scan idx[k] at pose[k]: new_r [B][n] float, new_bad [B][n] uint8.  Host buffers. */
zhsm_status zhsm_scan_project_batch(zhsm_handle *h, const float *scans, const uint8_t *bad, const int16_t *seg, int S,
                                    const int32_t *idx, const float *pose, int B, float *new_r, uint8_t *new_bad);

/* pm_is_corridor (:2058-2186) over S preprocessed scans: out[k] = 1 corridor, 0 room, -3 / -2 where the reference
 * throws 3 / 2 (fewer than two usable line directions).  out is a host array unless ZHSM_RESULTS_ON_DEVICE. */
zhsm_status zhsm_is_corridor_batch(zhsm_handle *h, const float *scans, const uint8_t *bad, const int16_t *seg, int S,
                                   int8_t *out, unsigned flags, void *cuda_stream);

/* ---- zhicp: point-to-point ICP (src/zhicp/ZHIcp_Warpper.cpp -> mrpt::slam::CICP, icpClassic) ---------------------
 * Options of mrpt::slam::CICP::TConfigParams the path reads.  zhsm_icp_default_options() fills the values
 * CICPWarpper::initICP sets (ZHIcp_Warpper.cpp:43-53) and MRPT's defaults for the rest. */
typedef struct zhsm_icp_options {
    int32_t maxIterations;              /* 100 (:46) */
    float thresholdAng;                 /* DEG2RAD(10) (:47) */
    float thresholdDist;                /* 0.75 (:48) */
    float ALFA;                         /* 0.5 (:49) */
    float smallestThresholdDist;        /* 0.05 (:50) */
    int32_t onlyClosestCorrespondences; /* 1 */
    float minAbsStep_trans;             /* 1e-6 */
    float minAbsStep_rot;               /* 1e-6 */
    int32_t corresponding_points_decimation; /* 5 */
    int32_t rotate_decimation_offset;   /* 0 never, 1 every iteration, 2 whenever the thresholds shrink (default) */
    float kernel_rho;                   /* 0.07 */
    int32_t use_kernel;                 /* 1 */
    float Axy_aprox_derivatives;        /* 0.05 */
} zhsm_icp_options;

void zhsm_icp_default_options(zhsm_icp_options *o);

/* Batched CICPWarpper::ICPMatch (ZHIcp_Warpper.cpp:125-176).  Point clouds live in a cloud pool: cloud c has cnt[c]
 * points stored at pts[2*off[c]] as x[0..cnt) followed by y[0..cnt) (metres, float32: the rx/ry, ax/ay arrays the
 * wrapper receives).  off/cnt ([C], HOST arrays always) are the pool's table of contents; pts is [2*npts] and may be
 * device memory (ZHSM_SCANS_ON_DEVICE).  Job k aligns cloud cur_idx[k] to the map cloud ref_idx[k] from the initial
 * guess init[k] ([B][3] doubles m,m,rad, NULL = zeros); a multi-start batch lists a pair G times with G guesses.
 * Per job: results[k].x = pose (getResult, :188-206), .cov = upper triangle of the 3x3 covariance, .iterations /
 * .goodness / .quality = getMatchQuality (:208-213), .nvalid = correspondences of the last iteration, .error = their
 * mean squared distance (an extra; MRPT does not report it); .valid = 0 when no correspondence was found.  The value ICPMatch returns is .goodness. */
zhsm_status zhsm_icp_batch(zhsm_handle *h, const float *pts, long long npts, const int32_t *off, const int32_t *cnt,
                           int C, const int32_t *ref_idx, const int32_t *cur_idx, const double *init, int B,
                           const zhsm_icp_options *opt, zhsm_result *results, unsigned flags, void *cuda_stream);

/* Multi-start reduction: results holds P pairs x G starts (pair-major).  best[p] = the preferred start among the
 * valid ones (-1 if none), best_results[p] = its record.  criterion ZHSM_BEST_GOODNESS (zhicp: highest goodness, then
 * lowest .error, then lowest start index -- the callers' acceptance is a goodness threshold, MapGraph.cpp:42,1385) or
 * ZHSM_BEST_MEAN_ERROR (PL-ICP: lowest error / nvalid, then lowest start index).  Either output may be NULL.
 * ZHSM_RESULTS_ON_DEVICE covers results, best and best_results. */
#define ZHSM_BEST_GOODNESS 0
#define ZHSM_BEST_MEAN_ERROR 1
zhsm_status zhsm_best_of_starts(zhsm_handle *h, const zhsm_result *results, int P, int G, int criterion,
                                int32_t *best, zhsm_result *best_results, unsigned flags, void *cuda_stream);

/* ---- loop-closure gating and consensus (src/mapGraph/MapGraph.cpp) --------------------------------
 * The arithmetic either side of the ICP batches in CMapGraph::addMapNode's loop search (:1342-1442).  All arrays are
 * HOST arrays (the graph lives on the host).  Node table: pose[N][3] = m_rootPose (x, y, theta), ld[N][2] / ru[N][2]
 * = m_ld_corner / m_ru_corner in the node frame, cov_xy[N][2] = (m_root_Cov(0,0), m_root_Cov(1,1)) or NULL for the
 * constant-covariance mode (m_bUseConstantCov, :229).
 * flags[m*N + j], cur = cur_idx[m], other node j: bit ZHSM_GATE_OVERLAPPED = isOverlappedArea(cur, j) (:957-994),
 * bit ZHSM_GATE_LOOPY = isLoopyArea(cur, j, beta) (:996-1032; the caller passes beta = 225, :1350).  A pair becomes an
 * ICP job when both bits are set and j is outside the adjacency window (:1345-1346). */
#define ZHSM_GATE_OVERLAPPED 1
#define ZHSM_GATE_LOOPY 2
zhsm_status zhsm_gate_candidates(zhsm_handle *h, const double *pose, const double *ld, const double *ru,
                                 const double *cov_xy, int N, const int32_t *cur_idx, int M, float beta, uint8_t *flags);
/* isAvaliableArea (:918-944) for M matched transforms trans[M][3] (cov_xy[M][2] of the current node or NULL) */
zhsm_status zhsm_available_area(zhsm_handle *h, const double *trans, const double *cov_xy, int M, uint8_t *ok);
/* S matched sets stored back to back (set s = entries [set_off[s], set_off[s+1]), poses[e][3] = the node pose each
 * match implies, good[e] = its goodness): deleteMisMatch (:169-209) -> keep[e]; calculateWeightedPose (:146-167) over
 * the kept entries -> final_pose[s][3]; the entry calculateFinalPose picks (:110-123) -> final_idx[s] (absolute entry
 * index, -1 for an empty set). */
zhsm_status zhsm_consensus(zhsm_handle *h, const double *poses, const float *good, const int32_t *set_off, int S,
                           uint8_t *keep, double *final_pose, int32_t *final_idx);

/* ---- hand-over formats -----------------------------------------------------------------------
 * covariance -> information, what the pose graph consumes from a matcher result: CMapNode::matchNodePairICP
 * (src/mapGraph/MapNode.cpp:638-651) fills a symmetric 3x3 from getResult's six numbers cov6[k] = xx xy xt yy yt tt
 * ((0,0) and (1,1) times `scale`, 1 in the reference) and sets information = covariance.inverse(); the loop-closure
 * caller replaces it by the constant diag(const_diag) when m_bUseConstantCov (src/mapGraph/MapGraph.cpp:250-261,
 * 1427-1429: m_cov_loopnode = diag(10,10,10)).  const_diag = NULL: the inverse; else the override.  info[k] is the
 * row-major 3x3, ok[k] (may be NULL) = 0 where the covariance is singular (Eigen would return inf/NaN; so does info).
 * Host arrays.  Also usable on zhsm_result.cov of PL-ICP (same six-number layout). */
zhsm_status zhsm_cov_to_information(zhsm_handle *h, const double *cov6, int M, double scale, const double *const_diag,
                                    double *info, uint8_t *ok);

/* The client -> server frame record (src/tcp_slam/clientFrontend.cpp:300-305, 405-407 sendFrameInfo[Cov]; written by
 * src/tcp_slam/clientLocal.cpp:189-199, parsed by src/tcp_slam/serverSocket.cpp:60-115): pose of the frame, the
 * synchronisation number, optionally the six covariance numbers, and the np ranges of the scan [cm] -- a scan-pool row.
 * zhsm_frame_pack writes one QDataStream block (big-endian: quint16 length, x y th as reals, qint32 syn, [6 doubles],
 * quint16 np, np reals; a real is an 8-byte double, QDataStream's default since Qt 4.6, or 4 bytes with
 * ZHSM_FRAME_QT_SINGLE) and returns its size (-1: does not fit cap / np > 65535 / block > 65535 bytes).
 * zhsm_frame_unpack returns the bytes consumed, 0 when the buffer does not yet hold a whole block (the server waits
 * for more, :69-74), -1 on a malformed block or np > cap_r.  Pure host code, no handle. */
typedef struct zhsm_frame {
    float x, y, th;  /* m, m, rad (curr_pose) */
    int32_t syn;     /* syn_num */
    int32_t np;      /* ranges that follow */
    int32_t reserved;
    double cov[6];   /* xx xy xt yy yt tt (USE_COV builds) */
} zhsm_frame;
#define ZHSM_FRAME_WITH_COV 1u
#define ZHSM_FRAME_QT_SINGLE 2u
long long zhsm_frame_pack(const zhsm_frame *f, const float *r, unsigned flags, uint8_t *buf, long long cap);
long long zhsm_frame_unpack(const uint8_t *buf, long long len, unsigned flags, zhsm_frame *f, float *r, int cap_r);

/* ---- submap cloud builder (src/mapGraph/MapNode.cpp:726-757 computePMAP + src/mapGraph/PMap.cpp) -------
 * Builds the point cloud zhicp aligns: K scans (scans[K][n] float cm, the handle's laser geometry) taken at traj[K][3]
 * (x, y, theta in the node frame, doubles like m_trajectory) are traced through a 5 cm hit/visit grid and the cells
 * whose hit ratio is exactly 1 are returned as (x, y) in metres, in the reference's row-major cell order.
 * *count = number of cloud points (may exceed cap; only cap are written); dims (optional) = grid size.  Host arrays. */
zhsm_status zhsm_submap_cloud(zhsm_handle *h, const float *scans, const double *traj, int K, float *out_x, float *out_y,
                              int cap, int *count, int *dims);

/* ---- several GPUs of one box from ONE process ---------------------------------------------------------
 * The reference has no multi-GPU code; north_star: batches "sharded across the 8 GPUs of one box, with NCCL over NVLink
 * used only to all-gather result poses and covariances for the pose graph", host side C++.  A group = one handle per
 * device + an NCCL communicator over them (ncclCommInitAll; libnccl.so.2 is loaded with dlopen here, libzhsm.so itself
 * does not depend on it).  zhsm_group_plicp_batch is zhsm_plicp_batch with HOST arrays: the B jobs are cut into
 * contiguous blocks of ceil(B / n) (GPU k owns jobs [k * block, (k+1) * block)), each GPU receives only the slice of the
 * scan pool its block references, all GPUs match at once, and ONE ncclAllGather of the 128-byte records leaves every
 * record on every GPU (zhsm_group_gathered: device memory of GPU k, [n * block] records, the first B in job order,
 * valid until the next group call) and in results[B] on the host.  flags: 0 or ZHSM_FILTER_SCANS.  One caller thread per
 * group.  The torchrun path (bench.py, one process per GPU) uses the same partition and the same gather. */
typedef struct zhsm_group zhsm_group;
zhsm_status zhsm_group_create(const int *devices, int n, zhsm_group **out);
void zhsm_group_destroy(zhsm_group *g);
int zhsm_group_size(zhsm_group *g);
zhsm_handle *zhsm_group_handle(zhsm_group *g, int k);
const char *zhsm_group_last_error(zhsm_group *g);
zhsm_status zhsm_group_set_laser(zhsm_group *g, const zhsm_laser *laser);
zhsm_status zhsm_group_plicp_batch(zhsm_group *g, const float *scans, int S, const int32_t *ref_idx,
                                   const int32_t *sens_idx, const double *first_guess, int B,
                                   const zhsm_plicp_params *params, zhsm_result *results, unsigned flags);
const zhsm_result *zhsm_group_gathered(zhsm_group *g, int k, int *block, int *B);

/* blocks until everything queued on the handle's stream (or the given one) is done */
zhsm_status zhsm_sync(zhsm_handle *h, void *cuda_stream);

/* number of kernels this handle has launched since creation (bench.py's gpu_launches) */
uint64_t zhsm_launch_count(zhsm_handle *h);
/* elapsed GPU milliseconds of the most recent batch call's main kernel, measured with CUDA events on
 * the launching stream (valid after zhsm_sync) */
float zhsm_last_kernel_ms(zhsm_handle *h);

/* ---- workload tooling (bench.py / tests; not part of the reference's interface) ----------------------------
 * Synthetic scan pairs of SURVEY 8d (C4/C5) generated on the device: scans[2p] = ref, scans[2p+1] = sens taken
 * from a pose displaced by truth[p] = (dx,dy,dtheta) uniform in +-dmax.  Philox counter = first_pair + p. */
zhsm_status zhsm_synth_pairs(zhsm_handle *h, float *d_scans, double *d_truth, long long first_pair, int npairs,
                             float dmax_xy_m, float dmax_th_rad, float noise_m, float dropout,
                             unsigned long long seed, void *cuda_stream);
/* measured CUDA-core FMA peak in TFLOP/s: which = 32 (FFMA) or 64 (DFMA) */
double zhsm_fp_peak_tflops(zhsm_handle *h, int which);

/* test hook, host only: ld_corr_hash of a correspondence stream (j1 + 1000*j2, or -1) evaluated sequentially
 * (seq) and with the kernel's chunk/GF(2)-matrix decomposition (par).  Returns 0 on success. */
int zhsm_debug_corr_hash(const int32_t *stream, int n, uint32_t *seq, uint32_t *par);
/* test hook, device: greatest real root of B quartics a0 + a1 x + .. + a4 x^4 (host arrays coef[B][5], root[B],
 * ok[B]) by the warp-cooperative solver the PL-ICP kernel uses for gpc_solve's lambda. */
/* test hook, device: out[k] = f(x[k], y[k]) with f = 0 sinf, 1 cosf, 2 atan2f(x = y-arg, y = x-arg) evaluated by the
 * device routines that mirror the host libm bit for bit (csrc/zh_libm.cuh).  Host arrays. */
zhsm_status zhsm_debug_libm(zhsm_handle *h, int which, const float *x, const float *y, float *out, int N);
zhsm_status zhsm_debug_quartic_roots(zhsm_handle *h, const double *coef, int B, double *root, int *ok);

#ifdef __cplusplus
}
#endif
#endif /* ZHSM_H */
