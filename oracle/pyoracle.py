"""ctypes bindings for the TEST-ONLY oracles under oracle/.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product package (laser_slam_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "liboracle.so")
_REF = os.path.join(_HERE, "_ref", "libzhpsm_ref.so")


def build(quiet=True):
    """Compile oracle/_build/liboracle.so and (when /root/reference exists) oracle/_ref/libzhpsm_ref.so."""
    subprocess.run(["make", "-C", _HERE, "all"], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


class SmParams(C.Structure):
    _fields_ = [
        ("laser_ref", C.c_void_p), ("laser_sens", C.c_void_p),
        ("first_guess", C.c_double * 3),
        ("max_angular_correction_deg", C.c_double), ("max_linear_correction", C.c_double),
        ("max_iterations", C.c_int), ("epsilon_xy", C.c_double), ("epsilon_theta", C.c_double),
        ("max_correspondence_dist", C.c_double), ("sigma", C.c_double),
        ("use_corr_tricks", C.c_int), ("restart", C.c_int),
        ("restart_threshold_mean_error", C.c_double), ("restart_dt", C.c_double),
        ("restart_dtheta", C.c_double), ("clustering_threshold", C.c_double),
        ("orientation_neighbourhood", C.c_int), ("use_point_to_line_distance", C.c_int),
        ("do_alpha_test", C.c_int), ("do_alpha_test_thresholdDeg", C.c_double),
        ("outliers_maxPerc", C.c_double), ("outliers_adaptive_order", C.c_double),
        ("outliers_adaptive_mult", C.c_double), ("do_visibility_test", C.c_int),
        ("outliers_remove_doubles", C.c_int), ("do_compute_covariance", C.c_int),
        ("debug_verify_tricks", C.c_int), ("use_ml_weights", C.c_int), ("use_sigma_weights", C.c_int),
        ("min_reading", C.c_double), ("max_reading", C.c_double),
    ]


class SmResult(C.Structure):
    _fields_ = [
        ("valid", C.c_int), ("x", C.c_double * 3), ("iterations", C.c_int), ("nvalid", C.c_int),
        ("error", C.c_double), ("cov_x_m", C.c_double * 9), ("corr_hash", C.c_uint), ("n_loops", C.c_int),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        _lib = C.CDLL(_LIB)
        _lib.orc_sm_default_params.argtypes = [C.POINTER(SmParams), C.c_int, C.c_int]
        _lib.orc_median_filter.argtypes = [C.c_void_p, C.c_int]
        _lib.orc_zhcsm_fmatch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int,
                                          C.POINTER(SmParams), C.c_void_p, C.c_void_p, C.POINTER(SmResult),
                                          C.c_void_p, C.c_void_p]
        _lib.orc_zhcsm_fmatch.restype = C.c_int
        _lib.orc_zhcsm_fmatch_batch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_float,
                                                C.c_float, C.c_int, C.c_int, C.POINTER(SmParams), C.c_void_p,
                                                C.c_void_p, C.c_int]
        _lib.orc_poly_greatest_real_root.argtypes = [C.c_void_p, C.c_void_p]
        _lib.orc_count_div100_mismatches.argtypes = [C.c_uint, C.c_uint]
        _lib.orc_count_div100_mismatches.restype = C.c_longlong
        _lib.orc_poly_greatest_real_root.restype = C.c_int
    return _lib


RESULT_DTYPE = np.dtype([("valid", "i4"), ("_p0", "i4"), ("x", "f8", 3), ("iterations", "i4"), ("nvalid", "i4"),
                         ("error", "f8"), ("cov", "f8", 9), ("corr_hash", "u4"), ("n_loops", "i4")])
assert RESULT_DTYPE.itemsize == C.sizeof(SmResult), (RESULT_DTYPE.itemsize, C.sizeof(SmResult))


def default_params(min_range_cm=10, max_range_cm=5000, **over):
    p = SmParams()
    lib().orc_sm_default_params(C.byref(p), int(min_range_cm), int(max_range_cm))
    for k, v in over.items():
        setattr(p, k, v)
    return p


def median_filter(r):
    """In-place 5-tap median of a float32 vector (returns a filtered copy)."""
    out = np.ascontiguousarray(r, dtype=np.float32).copy()
    lib().orc_median_filter(out.ctypes.data, out.size)
    return out


def median_filter_batch(scans, nthreads=1):
    """In place over a float32 [S][n] array."""
    assert scans.dtype == np.float32 and scans.flags.c_contiguous
    lib().orc_median_filter_batch.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib().orc_median_filter_batch(scans.ctypes.data, scans.shape[0], scans.shape[1], int(nthreads))
    return scans


def zhcsm_fmatch(ref_r_cm, sens_r_cm, fi_min, dfi, min_range_cm, max_range_cm, params=None, first_guess=None,
                 want_corr=False):
    """CCanonicalMatcher::FMatch on already median-filtered float32 cm ranges."""
    ref = np.ascontiguousarray(ref_r_cm, dtype=np.float32)
    sens = np.ascontiguousarray(sens_r_cm, dtype=np.float32)
    n = ref.size
    pose = np.zeros(3, np.float32)
    res = SmResult()
    j1 = np.full(n, -1, np.int32)
    j2 = np.full(n, -1, np.int32)
    fg = None if first_guess is None else np.ascontiguousarray(first_guess, dtype=np.float64)
    ret = lib().orc_zhcsm_fmatch(ref.ctypes.data, sens.ctypes.data, n, np.float32(fi_min), np.float32(dfi),
                                 int(min_range_cm), int(max_range_cm),
                                 C.byref(params) if params is not None else None,
                                 None if fg is None else fg.ctypes.data, pose.ctypes.data, C.byref(res),
                                 j1.ctypes.data if want_corr else None, j2.ctypes.data if want_corr else None)
    out = dict(ret=ret, pose_cm=pose, valid=res.valid, x=np.array(res.x[:]), iterations=res.iterations,
               nvalid=res.nvalid, error=res.error, cov=np.array(res.cov_x_m[:]).reshape(3, 3),
               corr_hash=res.corr_hash, n_loops=res.n_loops)
    if want_corr:
        out["j1"], out["j2"] = j1, j2
    return out


def count_div100_mismatches(lo, hi):
    """floats in [lo, hi) whose exact-FMA quotient f/100 (the CUDA kernel's) differs from the division"""
    lo_b, hi_b = (int(np.float32(v).view(np.uint32)) for v in (lo, hi))
    return int(lib().orc_count_div100_mismatches(lo_b, hi_b))


def zhcsm_fmatch_batch(scans_cm, ref_idx, sens_idx, fi_min, dfi, min_range_cm, max_range_cm, params=None,
                       first_guess=None, nthreads=1):
    scans = np.ascontiguousarray(scans_cm, dtype=np.float32)
    ri = np.ascontiguousarray(ref_idx, dtype=np.int32)
    si = np.ascontiguousarray(sens_idx, dtype=np.int32)
    out = np.zeros(ri.size, RESULT_DTYPE)
    fg = None if first_guess is None else np.ascontiguousarray(first_guess, dtype=np.float64)
    lib().orc_zhcsm_fmatch_batch(scans.ctypes.data, scans.shape[1], ri.ctypes.data, si.ctypes.data, ri.size,
                                 np.float32(fi_min), np.float32(dfi), int(min_range_cm), int(max_range_cm),
                                 C.byref(params) if params is not None else None,
                                 None if fg is None else fg.ctypes.data, out.ctypes.data, int(nthreads))
    return out


# ---------------------------------------------------------------------------------------------
# oracle/_ref/libzhpsm_ref.so: the reference's own zhpsm sources compiled by oracle/Makefile
# ---------------------------------------------------------------------------------------------
_ref = None


def have_ref():
    return os.path.exists(_REF)


def ref():
    global _ref
    if _ref is None:
        if not os.path.exists(_REF):
            raise RuntimeError("oracle/_ref/libzhpsm_ref.so is absent (build it where /root/reference exists: "
                               "make -C oracle ref)")
        _ref = C.CDLL(_REF)
        _ref.zhpsm_ref_create.restype = C.c_void_p
        _ref.zhpsm_ref_create.argtypes = [C.c_char_p, C.c_float, C.c_int]
        _ref.zhpsm_ref_destroy.argtypes = [C.c_void_p]
        _ref.zhpsm_ref_set_laser_name.argtypes = [C.c_void_p, C.c_char_p]
        _ref.zhpsm_ref_params.argtypes = [C.c_void_p] + [C.c_void_p] * 7
        _ref.zhpsm_ref_tables.argtypes = [C.c_void_p] * 4
        _ref.zhpsm_ref_preprocess.argtypes = [C.c_void_p] * 4
        _ref.zhpsm_ref_median.argtypes = [C.c_void_p] * 2
        _ref.zhpsm_ref_scan_project.argtypes = [C.c_void_p] * 7
        _ref.zhpsm_ref_match.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 9
        _ref.zhpsm_ref_match.restype = C.c_int
        _ref.zhpsm_ref_error_index2.argtypes = [C.c_void_p] * 11
        _ref.zhpsm_ref_fmatch.argtypes = [C.c_void_p] * 7 + [C.c_int, C.c_void_p, C.c_void_p]
        _ref.zhpsm_ref_is_corridor.argtypes = [C.c_void_p] * 4
        _ref.zhpsm_ref_is_corridor.restype = C.c_int
        _ref.zhpsm_ref_match_batch.argtypes = [C.c_char_p, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                               C.c_void_p, C.c_int]
    return _ref


class RefPolarMatch:
    """The reference's CPolarMatch (src/zhpsm/ZHPolar_Match.h:16-88), compiled unmodified."""

    def __init__(self, laser, fi_min=float("nan"), max_range_cm=0):
        self.laser = laser.encode()
        self.fi_min_override = float(fi_min)
        self.max_range_override = int(max_range_cm)
        self.h = ref().zhpsm_ref_create(self.laser, C.c_float(fi_min), int(max_range_cm))
        if not self.h:
            raise ValueError("unknown laser " + laser)
        n, mn, mx, mv, w = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
        fi, dfi = C.c_float(), C.c_float()
        ref().zhpsm_ref_params(self.h, C.byref(n), C.byref(fi), C.byref(dfi), C.byref(mn), C.byref(mx),
                               C.byref(mv), C.byref(w))
        self.n, self.fi_min, self.dfi = n.value, np.float32(fi.value), np.float32(dfi.value)
        self.min_range, self.max_range, self.min_valid, self.window = mn.value, mx.value, mv.value, w.value

    def __del__(self):
        try:
            if self.h:
                ref().zhpsm_ref_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def set_laser_name(self, name):
        """e.g. "HOKUYO_UTM_30LX": turns on the delta < PM_MAX_ERROR gate of pm_orientation_search (:1182-1191)."""
        ref().zhpsm_ref_set_laser_name(self.h, name.encode())

    def tables(self):
        fi, si, co = (np.zeros(self.n, np.float32) for _ in range(3))
        ref().zhpsm_ref_tables(self.h, fi.ctypes.data, si.ctypes.data, co.ctypes.data)
        return fi, si, co

    def preprocess(self, r, bad=None):
        r = np.ascontiguousarray(r, np.float32).copy()
        bad = np.zeros(self.n, np.int32) if bad is None else np.ascontiguousarray(bad, np.int32).copy()
        seg = np.zeros(self.n, np.int32)
        ref().zhpsm_ref_preprocess(self.h, r.ctypes.data, bad.ctypes.data, seg.ctypes.data)
        return r, bad, seg

    def median(self, r):
        r = np.ascontiguousarray(r, np.float32).copy()
        ref().zhpsm_ref_median(self.h, r.ctypes.data)
        return r

    def is_corridor(self, r, bad, seg):
        """pm_is_corridor on a preprocessed scan: 1 / 0, or -3 / -2 when the reference throws 3 / 2."""
        r, bad, seg = _f32(r), _i32(bad), _i32(seg)
        return int(ref().zhpsm_ref_is_corridor(self.h, r.ctypes.data, bad.ctypes.data, seg.ctypes.data))

    def scan_project(self, r, bad, seg, pose):
        r, bad, seg = _f32(r), _i32(bad), _i32(seg)
        pose = np.asarray(pose, np.float32)
        new_r = np.zeros(self.n, np.float32)
        new_bad = np.zeros(self.n, np.int32)
        ref().zhpsm_ref_scan_project(self.h, r.ctypes.data, bad.ctypes.data, seg.ctypes.data, pose.ctypes.data,
                                     new_r.ctypes.data, new_bad.ctypes.data)
        return new_r, new_bad

    def match(self, algo, ref_scan, act_scan, act_pose=(0, 0, 0), ref_pose=(0, 0, 0)):
        """algo 0 pm_psm, 1 pm_icp, 2 pm_psm->pm_icp; scans = (r,bad,seg) preprocessed."""
        rr, rb, rs = _f32(ref_scan[0]), _i32(ref_scan[1]), _i32(ref_scan[2])
        ar, ab, as_ = _f32(act_scan[0]), _i32(act_scan[1]), _i32(act_scan[2])
        rp = np.asarray(ref_pose, np.float32)
        pose = np.asarray(act_pose, np.float32).copy()
        err = C.c_float()
        st = ref().zhpsm_ref_match(self.h, int(algo), rr.ctypes.data, rb.ctypes.data, rs.ctypes.data, rp.ctypes.data,
                                   ar.ctypes.data, ab.ctypes.data, as_.ctypes.data, pose.ctypes.data,
                                   C.cast(C.byref(err), C.c_void_p))
        return st, pose, err.value

    def error_index2(self, ref_scan, act_scan, act_pose, ref_pose=(0, 0, 0)):
        rr, rb, rs = _f32(ref_scan[0]), _i32(ref_scan[1]), _i32(ref_scan[2])
        ar, ab, as_ = _f32(act_scan[0]), _i32(act_scan[1]), _i32(act_scan[2])
        rp, ap = np.asarray(ref_pose, np.float32), np.asarray(act_pose, np.float32)
        out = np.zeros(2, np.float32)
        assoc = C.c_int()
        ref().zhpsm_ref_error_index2(self.h, rr.ctypes.data, rb.ctypes.data, rs.ctypes.data, rp.ctypes.data,
                                     ar.ctypes.data, ab.ctypes.data, as_.ctypes.data, ap.ctypes.data,
                                     out.ctypes.data, C.cast(C.byref(assoc), C.c_void_p))
        return out, assoc.value

    def match_batch(self, algo, r, bad, seg, ref_idx, act_idx, nthreads=1):
        r, bad, seg = _f32(r), _i32(bad), _i32(seg)
        ri, ai = _i32(ref_idx), _i32(act_idx)
        poses = np.zeros((ri.size, 3), np.float32)
        status = np.zeros(ri.size, np.int32)
        ref().zhpsm_ref_match_batch(self.laser, C.c_float(self.fi_min_override), self.max_range_override, int(algo),
                                    r.ctypes.data, bad.ctypes.data, seg.ctypes.data, self.n, ri.ctypes.data,
                                    ai.ctypes.data, ri.size, poses.ctypes.data, status.ctypes.data, int(nthreads))
        return poses, status


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


def _i32(a):
    return np.ascontiguousarray(a, np.int32)


# ---------------------------------------------------------------------------------------------
# zhicp: MRPT 0.9.6 CICP (classic) restatement, oracle/icp_oracle.c  -- PARITY UNPINNED
# ---------------------------------------------------------------------------------------------
class IcpOptions(C.Structure):
    _fields_ = [("maxIterations", C.c_int), ("thresholdAng", C.c_float), ("thresholdDist", C.c_float),
                ("ALFA", C.c_float), ("smallestThresholdDist", C.c_float), ("onlyClosestCorrespondences", C.c_int),
                ("minAbsStep_trans", C.c_float), ("minAbsStep_rot", C.c_float),
                ("corresponding_points_decimation", C.c_int), ("rotate_decimation_offset", C.c_int),
                ("kernel_rho", C.c_float), ("use_kernel", C.c_int), ("Axy_aprox_derivatives", C.c_float)]


ICP_RESULT_DTYPE = np.dtype([("pose", "f8", 3), ("cov", "f8", 9), ("nIterations", "i4"), ("nCorrespondences", "i4"),
                             ("goodness", "f4"), ("quality", "f4")])


def icp_default_options(**over):
    o = IcpOptions()
    lib().orc_icp_default_options.argtypes = [C.POINTER(IcpOptions)]
    lib().orc_icp_default_options(C.byref(o))
    for k, v in over.items():
        setattr(o, k, v)
    return o


def pack_clouds(clouds):
    """list of (x[N], y[N]) -> (pts float32 [2*sum N] as x-block then y-block per cloud, off int32, cnt int32)."""
    cnt = np.array([len(c[0]) for c in clouds], np.int32)
    off = np.zeros(len(clouds), np.int32)
    off[1:] = np.cumsum(cnt)[:-1]
    pts = np.zeros(2 * int(cnt.sum()), np.float32)
    for k, (x, y) in enumerate(clouds):
        o, n = 2 * int(off[k]), int(cnt[k])
        pts[o:o + n] = x
        pts[o + n:o + 2 * n] = y
    return pts, off, cnt


def mrpt_icp_batch(pts, off, cnt, ref_idx, cur_idx, init, options=None, nthreads=1):
    """CICPWarpper::ICPMatch for every job k: clouds ref_idx[k] (map) and cur_idx[k], initial guess init[k]."""
    pts = np.ascontiguousarray(pts, np.float32)
    off = np.ascontiguousarray(off, np.int32)
    cnt = np.ascontiguousarray(cnt, np.int32)
    ri = np.ascontiguousarray(ref_idx, np.int32)
    ci = np.ascontiguousarray(cur_idx, np.int32)
    init = np.ascontiguousarray(init, np.float64).reshape(-1, 3)
    assert init.shape[0] == ri.size == ci.size
    out = np.zeros(ri.size, ICP_RESULT_DTYPE)
    assert ICP_RESULT_DTYPE.itemsize == 112
    L = lib()
    L.orc_mrpt_icp_batch.argtypes = [C.c_void_p] * 6 + [C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    L.orc_mrpt_icp_batch(pts.ctypes.data, off.ctypes.data, cnt.ctypes.data, ri.ctypes.data, ci.ctypes.data,
                         init.ctypes.data, ri.size, C.byref(options) if options is not None else None,
                         out.ctypes.data, int(nthreads))
    return out


# ---------------------------------------------------------------------------------------------
# loop-closure gating and consensus (src/mapGraph/MapGraph.cpp), oracle/graph_oracle.c
# ---------------------------------------------------------------------------------------------
def gate_candidates(pose, ld, ru, cur_idx, cov_xy=None, beta=225.0):
    pose = np.ascontiguousarray(pose, np.float64); ld = np.ascontiguousarray(ld, np.float64)
    ru = np.ascontiguousarray(ru, np.float64); cur_idx = np.ascontiguousarray(cur_idx, np.int32)
    cov = None if cov_xy is None else np.ascontiguousarray(cov_xy, np.float64)
    N, M = pose.shape[0], cur_idx.size
    flags = np.zeros((M, N), np.uint8)
    L = lib()
    L.orc_gate_candidates.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_void_p]
    L.orc_gate_candidates(pose.ctypes.data, ld.ctypes.data, ru.ctypes.data, None if cov is None else cov.ctypes.data, N,
                          cur_idx.ctypes.data, M, C.c_float(beta), flags.ctypes.data)
    return flags


def available_area(trans, cov_xy=None):
    trans = np.ascontiguousarray(trans, np.float64).reshape(-1, 3)
    cov = None if cov_xy is None else np.ascontiguousarray(cov_xy, np.float64)
    ok = np.zeros(trans.shape[0], np.uint8)
    L = lib()
    L.orc_available_area.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.orc_available_area(trans.ctypes.data, None if cov is None else cov.ctypes.data, trans.shape[0], ok.ctypes.data)
    return ok


def consensus(poses, good, set_off):
    poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 3)
    good = np.ascontiguousarray(good, np.float32); set_off = np.ascontiguousarray(set_off, np.int32)
    S = set_off.size - 1
    keep = np.zeros(poses.shape[0], np.uint8); fp = np.zeros((S, 3)); fi = np.zeros(S, np.int32)
    L = lib()
    L.orc_consensus.argtypes = [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 3
    L.orc_consensus(poses.ctypes.data, good.ctypes.data, set_off.ctypes.data, S, keep.ctypes.data, fp.ctypes.data,
                    fi.ctypes.data)
    return keep, fp, fi


def synth_pairs(npairs, n, fi_min, dfi, max_range_cm, first_pair=0, dmax_xy_m=0.10, dmax_th_rad=0.0872664626,
                noise_m=0.01, dropout=0.01, seed=20261104, nthreads=1):
    """The HOST twin of the device workload generator (oracle/synth_oracle.c): SURVEY 8d's synthetic scan pairs without
    loading the product library.  -> (scans float32 [2*npairs][n] cm, truth float64 [npairs][3])."""
    scans = np.zeros((2 * npairs, n), np.float32)
    truth = np.zeros((npairs, 3), np.float64)
    L = lib()
    L.orc_synth_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int,
                                  C.c_float, C.c_float, C.c_float, C.c_float, C.c_ulonglong, C.c_int]
    L.orc_synth_pairs(scans.ctypes.data, truth.ctypes.data, int(first_pair), int(npairs), int(n), float(fi_min), float(dfi),
                      int(max_range_cm), float(dmax_xy_m), float(dmax_th_rad), float(noise_m), float(dropout), int(seed),
                      int(nthreads))
    return scans, truth


def cov_to_information(cov6, scale=1.0, const_diag=None):
    """matchNodePairICP's information = covariance.inverse() (Eigen 3x3 cofactor formula), or the constant override."""
    cov6 = np.ascontiguousarray(cov6, np.float64).reshape(-1, 6)
    info = np.zeros((cov6.shape[0], 3, 3))
    cd = None if const_diag is None else np.ascontiguousarray(const_diag, np.float64)
    L = lib()
    L.orc_cov_to_information.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.c_void_p]
    L.orc_cov_to_information(cov6.ctypes.data, cov6.shape[0], float(scale), None if cd is None else cd.ctypes.data,
                             info.ctypes.data)
    return info


def ref_oplus_point(pose, pt):
    """OrientedPoint2D::oplus(Point2D) executed by the compiled reference (src/zhpsm/point.cpp:63-69)."""
    R = ref()
    R.zhpsm_ref_oplus_point.argtypes = [C.c_void_p] * 3
    pose = np.ascontiguousarray(pose, np.float64); pt = np.ascontiguousarray(pt, np.float64); out = np.zeros(2)
    R.zhpsm_ref_oplus_point(pose.ctypes.data, pt.ctypes.data, out.ctypes.data)
    return out


# ---------------------------------------------------------------------------------------------
# submap cloud builder: the reference's own CPMap (src/mapGraph/PMap.cpp, line.cpp) compiled into oracle/_ref
# ---------------------------------------------------------------------------------------------
_PMAP = os.path.join(_HERE, "_ref", "libpmap_ref.so")
_pmap = None


def have_pmap_ref():
    return os.path.exists(_PMAP)


def ref_submap_cloud(scans_cm, traj, min_angle, dfi, max_range_cm, cap=1 << 20):
    """CMapNode::computePMAP (src/mapGraph/MapNode.cpp:726-757) on K scans [K][n] (cm) taken at traj [K][3]:
    returns (x, y) float32 of the occupied cells, in the reference's row-major cell order, and the grid size."""
    global _pmap
    if _pmap is None:
        if not os.path.exists(_PMAP):
            raise RuntimeError("oracle/_ref/libpmap_ref.so is absent (make -C oracle ref, where /root/reference exists)")
        _pmap = C.CDLL(_PMAP)
        _pmap.pmap_ref_submap_cloud.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int,
                                                C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        _pmap.pmap_ref_submap_cloud.restype = C.c_int
    scans = np.ascontiguousarray(scans_cm, np.float32)
    traj = np.ascontiguousarray(traj, np.float64)
    x = np.zeros(cap, np.float32); y = np.zeros(cap, np.float32); dims = np.zeros(2, np.int32)
    cnt = _pmap.pmap_ref_submap_cloud(scans.ctypes.data, traj.ctypes.data, scans.shape[0], scans.shape[1],
                                      C.c_float(min_angle), C.c_float(dfi), int(max_range_cm), x.ctypes.data,
                                      y.ctypes.data, cap, dims.ctypes.data)
    assert cnt <= cap
    return x[:cnt].copy(), y[:cnt].copy(), dims
