"""ctypes binding of include/zhsm.h.  Host buffers are numpy arrays; device buffers are anything with
`data_ptr()` (torch CUDA tensors) -- only raw pointers cross the boundary."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

PM_PI = 3.141592654  # src/zhpsm/PolarParameter.h:38 (the reference's truncated pi)
PM_D2R = PM_PI / 180.0

# Base_PARAM tables (src/zhpsm/PolarParameter.h:70-83): points, fov_deg, max_range_cm, min_valid, window, fi_min_deg
LASER_TABLE = {
    "LMS511": (361, 180, 5000, 80, 40, 0),
    "LMS211": (181, 180, 5000, 40, 20, -90),
    "LMS151": (541, 270, 5000, 100, 50, -45),
}


class ZhsmError(RuntimeError):
    pass


class Laser(C.Structure):
    _fields_ = [("n", C.c_int32), ("fi_min", C.c_float), ("dfi", C.c_float), ("min_range_cm", C.c_int32),
                ("max_range_cm", C.c_int32), ("min_valid_points", C.c_int32), ("scan_window", C.c_int32),
                ("flags", C.c_int32)]


def laser_from_table(name, fi_min=None, max_range_cm=None):
    """Base_PARAM ctor arithmetic (PolarParameter.h:57-66): pm_dfi = (fov*D2R)/(points-1.0) narrowed to float."""
    pts, fov, mx, mv, win, fmin_deg = LASER_TABLE[name]
    lz = Laser()
    lz.n = pts
    lz.fi_min = np.float32(fmin_deg * PM_D2R) if fi_min is None else np.float32(fi_min)
    lz.dfi = np.float32((fov * PM_D2R) / (pts - 1.0))
    lz.min_range_cm = 10
    lz.max_range_cm = mx if max_range_cm is None else int(max_range_cm)
    lz.min_valid_points = mv
    lz.scan_window = win
    return lz


class PlicpParams(C.Structure):
    _fields_ = [(k, C.c_double) for k in (
        "max_angular_correction_deg", "max_linear_correction", "epsilon_xy", "epsilon_theta",
        "max_correspondence_dist", "sigma", "restart_threshold_mean_error", "restart_dt", "restart_dtheta",
        "outliers_maxPerc", "outliers_adaptive_order", "outliers_adaptive_mult", "min_reading", "max_reading")] + [
        (k, C.c_int32) for k in ("max_iterations", "use_corr_tricks", "restart", "use_point_to_line_distance",
                                 "outliers_remove_doubles", "do_compute_covariance")] + [("reserved", C.c_int32 * 2)]


class IcpOptions(C.Structure):
    """zhsm_icp_options (include/zhsm.h): mrpt::slam::CICP::TConfigParams as CICPWarpper::initICP sets them."""
    _fields_ = [("maxIterations", C.c_int32), ("thresholdAng", C.c_float), ("thresholdDist", C.c_float),
                ("ALFA", C.c_float), ("smallestThresholdDist", C.c_float), ("onlyClosestCorrespondences", C.c_int32),
                ("minAbsStep_trans", C.c_float), ("minAbsStep_rot", C.c_float),
                ("corresponding_points_decimation", C.c_int32), ("rotate_decimation_offset", C.c_int32),
                ("kernel_rho", C.c_float), ("use_kernel", C.c_int32), ("Axy_aprox_derivatives", C.c_float)]


class Frame(C.Structure):
    """zhsm_frame: the client -> server frame record (src/tcp_slam/clientFrontend.cpp:300-305, 405-407)."""
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("th", C.c_float), ("syn", C.c_int32), ("np", C.c_int32),
                ("reserved", C.c_int32), ("cov", C.c_double * 6)]


FRAME_WITH_COV, FRAME_QT_SINGLE = 1, 2


def frame_pack(x, y, th, syn, r, cov=None, single=False):
    """One QDataStream block of the reference's TCP frame record (bytes)."""
    L = load_library()
    r = np.ascontiguousarray(r, np.float32)
    f = Frame(float(x), float(y), float(th), int(syn), int(r.size), 0)
    flags = (FRAME_WITH_COV if cov is not None else 0) | (FRAME_QT_SINGLE if single else 0)
    if cov is not None:
        for k in range(6):
            f.cov[k] = float(cov[k])
    buf = np.zeros(2 + 128 + 8 * r.size, np.uint8)
    n = L.zhsm_frame_pack(C.byref(f), r.ctypes.data, flags, buf.ctypes.data, buf.size)
    if n < 0:
        raise ZhsmError("zhsm_frame_pack: frame does not fit a quint16-length block")
    return buf[:n].tobytes()


def frame_unpack(data, with_cov=False, single=False, cap=65535):
    """-> (consumed bytes, dict) ; consumed 0 = incomplete block, raises on a malformed one."""
    L = load_library()
    buf = np.frombuffer(data, np.uint8)
    r = np.zeros(cap, np.float32)
    f = Frame()
    flags = (FRAME_WITH_COV if with_cov else 0) | (FRAME_QT_SINGLE if single else 0)
    n = L.zhsm_frame_unpack(buf.ctypes.data if buf.size else None, buf.size, flags, C.byref(f), r.ctypes.data, cap)
    if n < 0:
        raise ZhsmError("zhsm_frame_unpack: malformed frame block")
    if n == 0:
        return 0, None
    return int(n), dict(x=f.x, y=f.y, th=f.th, syn=f.syn, r=r[:f.np].copy(), cov=np.array(list(f.cov)))


RESULT_DTYPE = np.dtype([("x", "f8", 3), ("cov", "f8", 6), ("error", "f8"), ("valid", "i4"), ("iterations", "i4"),
                         ("nvalid", "i4"), ("status", "i4"), ("goodness", "f4"), ("quality", "f4"),
                         ("corr_hash", "u4"), ("n_loops", "i4"), ("reserved", "i4", 4)])
assert RESULT_DTYPE.itemsize == 128

SCANS_ON_DEVICE, JOBS_ON_DEVICE, RESULTS_ON_DEVICE, ALL_ON_DEVICE = 1, 2, 4, 7
FILTER_SCANS, WRITEBACK_SCANS, MULTISTART = 8, 16, 32
ALGO_PSM, ALGO_PM_ICP, ALGO_PSM_ICP, ALGO_FMATCH = 0, 1, 2, 3
BEST_GOODNESS, BEST_MEAN_ERROR = 0, 1
LASER_UTM_GATE = 1  # zhsm_laser.flags: pm_orientation_search's HOKUYO_UTM_30LX branch

_lib = None


def lib_path():
    v = os.environ.get("ZHSM_LIB_VARIANT")  # tuning builds only (laser_slam_b200.build.build(variant=...))
    return os.path.join(_HERE, "libzhsm_%s.so" % v if v else "libzhsm.so")


def load_library():
    """Loads libzhsm.so; raises loudly when it is missing (there is no fallback path)."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not os.path.exists(p):
        raise ZhsmError("libzhsm.so is not built: run `python -m laser_slam_b200.build` (nvcc, sm_100a)")
    L = C.CDLL(p)
    vp, i32, u32 = C.c_void_p, C.c_int, C.c_uint
    L.zhsm_create.argtypes = [i32, C.POINTER(vp)]
    L.zhsm_destroy.argtypes = [vp]
    L.zhsm_destroy.restype = None
    L.zhsm_last_error.argtypes = [vp]
    L.zhsm_last_error.restype = C.c_char_p
    L.zhsm_set_laser.argtypes = [vp, C.POINTER(Laser)]
    L.zhsm_plicp_default_params.argtypes = [C.POINTER(PlicpParams), i32, i32]
    L.zhsm_plicp_default_params.restype = None
    L.zhsm_median_filter_batch.argtypes = [vp, vp, i32, u32, vp]
    L.zhsm_plicp_batch.argtypes = [vp, vp, i32, vp, vp, vp, i32, C.POINTER(PlicpParams), vp, vp, vp, u32, vp]
    L.zhsm_plicp_batch_noalloc.argtypes = [vp, vp, i32, vp, vp, vp, i32, C.POINTER(PlicpParams), vp, vp, vp, vp]
    L.zhsm_cov_to_information.argtypes = [vp, vp, i32, C.c_double, vp, vp, vp]
    L.zhsm_frame_pack.argtypes = [C.POINTER(Frame), vp, u32, vp, C.c_longlong]
    L.zhsm_frame_pack.restype = C.c_longlong
    L.zhsm_frame_unpack.argtypes = [vp, C.c_longlong, u32, C.POINTER(Frame), vp, i32]
    L.zhsm_frame_unpack.restype = C.c_longlong
    L.zhsm_group_create.argtypes = [vp, i32, C.POINTER(vp)]
    L.zhsm_group_destroy.argtypes = [vp]
    L.zhsm_group_destroy.restype = None
    L.zhsm_group_size.argtypes = [vp]
    L.zhsm_group_handle.argtypes = [vp, i32]
    L.zhsm_group_handle.restype = vp
    L.zhsm_group_last_error.argtypes = [vp]
    L.zhsm_group_last_error.restype = C.c_char_p
    L.zhsm_group_set_laser.argtypes = [vp, C.POINTER(Laser)]
    L.zhsm_group_plicp_batch.argtypes = [vp, vp, i32, vp, vp, vp, i32, C.POINTER(PlicpParams), vp, u32]
    L.zhsm_group_gathered.argtypes = [vp, i32, C.POINTER(i32), C.POINTER(i32)]
    L.zhsm_group_gathered.restype = vp
    L.zhsm_sync.argtypes = [vp, vp]
    L.zhsm_launch_count.argtypes = [vp]
    L.zhsm_launch_count.restype = C.c_uint64
    L.zhsm_last_kernel_ms.argtypes = [vp]
    L.zhsm_last_kernel_ms.restype = C.c_float
    L.zhsm_synth_pairs.argtypes = [vp, vp, vp, C.c_longlong, i32, C.c_float, C.c_float, C.c_float, C.c_float,
                                   C.c_ulonglong, vp]
    L.zhsm_fp_peak_tflops.argtypes = [vp, i32]
    L.zhsm_fp_peak_tflops.restype = C.c_double
    L.zhsm_preprocess_batch.argtypes = [vp, vp, vp, vp, i32, u32, vp]
    L.zhsm_psm_batch.argtypes = [vp, vp, vp, vp, i32, vp, vp, vp, i32, i32, vp, u32, vp]
    L.zhsm_scan_project_batch.argtypes = [vp, vp, vp, vp, i32, vp, vp, i32, vp, vp]
    L.zhsm_debug_libm.argtypes = [vp, i32, vp, vp, vp, i32]
    L.zhsm_is_corridor_batch.argtypes = [vp, vp, vp, vp, i32, vp, u32, vp]
    L.zhsm_gate_candidates.argtypes = [vp, vp, vp, vp, vp, i32, vp, i32, C.c_float, vp]
    L.zhsm_available_area.argtypes = [vp, vp, vp, i32, vp]
    L.zhsm_consensus.argtypes = [vp, vp, vp, vp, i32, vp, vp, vp]
    L.zhsm_submap_cloud.argtypes = [vp, vp, vp, i32, vp, vp, i32, vp, vp]
    L.zhsm_icp_default_options.argtypes = [C.POINTER(IcpOptions)]
    L.zhsm_icp_default_options.restype = None
    L.zhsm_icp_batch.argtypes = [vp, vp, C.c_longlong, vp, vp, i32, vp, vp, vp, i32, C.POINTER(IcpOptions), vp, u32, vp]
    L.zhsm_best_of_starts.argtypes = [vp, vp, i32, i32, i32, vp, vp, u32, vp]
    _lib = L
    return L


def _ptr(a):
    if a is None:
        return None
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    return a.ctypes.data


def _is_dev(a):
    return hasattr(a, "data_ptr")


_TORCH_DTYPES = {"float32": "torch.float32", "float64": "torch.float64", "int32": "torch.int32", "int16": "torch.int16",
                 "uint8": "torch.uint8", "int8": "torch.int8"}


def _check_dev(a, what, dtype, device, cols=None, rows=None):
    """A device tensor gets the checks the numpy path makes: only its raw pointer crosses the boundary, so a sliced,
    mistyped or misplaced tensor would otherwise be read as garbage (or fault the whole batch)."""
    if a is None:
        return
    if not a.is_cuda:
        raise ZhsmError(f"{what}: expected a CUDA tensor")
    if a.device.index != device:
        raise ZhsmError(f"{what}: tensor is on cuda:{a.device.index}, the engine's handle on cuda:{device}")
    if str(a.dtype) != _TORCH_DTYPES[dtype]:
        raise ZhsmError(f"{what}: dtype {a.dtype}, expected {dtype}")
    if not a.is_contiguous():
        raise ZhsmError(f"{what}: tensor must be contiguous")
    if cols is not None and (a.dim() != 2 or a.shape[1] != cols):
        raise ZhsmError(f"{what}: shape {tuple(a.shape)}, expected [*][{cols}]")
    if rows is not None and a.shape[0] != rows:
        raise ZhsmError(f"{what}: {a.shape[0]} rows, expected {rows}")


def _same_side(what, *arrs):
    sides = {_is_dev(a) for a in arrs if a is not None}
    if len(sides) > 1:
        raise ZhsmError(f"{what}: these arrays must all be host (numpy) or all device (torch) arrays")


class Engine:
    """One handle = one GPU + one stream.  Mirrors the three matcher classes' batched form."""

    def __init__(self, device=0):
        self.L = load_library()
        h = C.c_void_p()
        st = self.L.zhsm_create(int(device), C.byref(h))
        if st != 0:
            raise ZhsmError(f"zhsm_create(device={device}) failed with status {st}: no usable sm_100 GPU "
                            "(this engine has no CPU fallback)")
        self.h = h
        self.device = int(device)
        self.laser = None

    def close(self):
        if getattr(self, "h", None):
            self.L.zhsm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st, what):
        if st != 0:
            raise ZhsmError(f"{what}: status {st}: {self.L.zhsm_last_error(self.h).decode()}")

    def set_laser(self, laser):
        self.laser = laser
        self._check(self.L.zhsm_set_laser(self.h, C.byref(laser)), "zhsm_set_laser")

    def default_plicp_params(self):
        p = PlicpParams()
        self.L.zhsm_plicp_default_params(C.byref(p), self.laser.min_range_cm, self.laser.max_range_cm)
        return p

    def median_filter(self, scans, stream=None):
        """In place.  numpy float32 [S][n] (host) or a CUDA tensor."""
        dev = _is_dev(scans)
        if not dev:
            assert scans.dtype == np.float32 and scans.flags.c_contiguous and scans.shape[1] == self.laser.n
        else:
            _check_dev(scans, "median_filter: scans", "float32", self.device, cols=self.laser.n)
        S = scans.shape[0]
        self._check(self.L.zhsm_median_filter_batch(self.h, _ptr(scans), S, SCANS_ON_DEVICE if dev else 0, stream),
                    "zhsm_median_filter_batch")
        return scans

    def plicp(self, scans, ref_idx, sens_idx, first_guess=None, params=None, results=None, want_corr=False,
              stream=None, filter_scans=False, writeback=False, multistart=False):
        """Batched CCanonicalMatcher::FMatch on median-filtered scans.  Host (numpy) or device (torch) buffers;
        all job arrays must live on the same side."""
        n = self.laser.n
        dev_scans = _is_dev(scans)
        dev_jobs = _is_dev(ref_idx)
        B = int(ref_idx.shape[0])
        S = int(scans.shape[0])
        flags = (SCANS_ON_DEVICE if dev_scans else 0) | (JOBS_ON_DEVICE if dev_jobs else 0)
        flags |= (FILTER_SCANS if filter_scans else 0) | (WRITEBACK_SCANS if writeback else 0)
        flags |= MULTISTART if multistart else 0   # jobs in runs of four that share both scans (see zhsm.h)
        _same_side("plicp: ref_idx / sens_idx / first_guess", ref_idx, sens_idx, first_guess)
        if not dev_scans:
            assert scans.dtype == np.float32 and scans.flags.c_contiguous and scans.shape[1] == n
        else:
            _check_dev(scans, "plicp: scans", "float32", self.device, cols=n)
        if dev_jobs:
            _check_dev(ref_idx, "plicp: ref_idx", "int32", self.device)
            _check_dev(sens_idx, "plicp: sens_idx", "int32", self.device, rows=B)
            _check_dev(first_guess, "plicp: first_guess", "float64", self.device, cols=3, rows=B)
        if results is not None and _is_dev(results):
            _check_dev(results, "plicp: results", "uint8", self.device, cols=128, rows=B)
        if not dev_jobs:
            ref_idx = np.ascontiguousarray(ref_idx, np.int32)
            sens_idx = np.ascontiguousarray(sens_idx, np.int32)
            if first_guess is not None:
                first_guess = np.ascontiguousarray(first_guess, np.float64)
        j1 = j2 = None
        if results is None:
            results = np.zeros(B, RESULT_DTYPE)
        if _is_dev(results):
            flags |= RESULTS_ON_DEVICE
            if want_corr:
                import torch
                j1 = torch.empty((B, n), dtype=torch.int16, device=results.device)
                j2 = torch.empty((B, n), dtype=torch.int16, device=results.device)
        elif want_corr:
            j1 = np.zeros((B, n), np.int16)
            j2 = np.zeros((B, n), np.int16)
        st = self.L.zhsm_plicp_batch(self.h, _ptr(scans), S, _ptr(ref_idx), _ptr(sens_idx), _ptr(first_guess), B,
                                     C.byref(params) if params is not None else None, _ptr(results), _ptr(j1),
                                     _ptr(j2), flags, stream)
        self._check(st, "zhsm_plicp_batch")
        if want_corr:
            return results, j1, j2
        return results

    def plicp_noalloc(self, scans, ref_idx, sens_idx, results, first_guess=None, params=None, corr_j1=None,
                      corr_j2=None, stream=None):
        """zhsm_plicp_batch_noalloc: every buffer is a CUDA tensor of the caller's; nothing is allocated (the kernels'
        scratch is a static device array: B * beams <= 1 048 576, S <= 8192).  Asynchronous: sync() before reading."""
        for a in (scans, ref_idx, sens_idx, results):
            if not _is_dev(a):
                raise ZhsmError("plicp_noalloc: scans, job lists and results must be device tensors")
        B, n = int(ref_idx.shape[0]), self.laser.n
        _check_dev(scans, "plicp_noalloc: scans", "float32", self.device, cols=n)
        _check_dev(ref_idx, "plicp_noalloc: ref_idx", "int32", self.device)
        _check_dev(sens_idx, "plicp_noalloc: sens_idx", "int32", self.device, rows=B)
        _check_dev(first_guess, "plicp_noalloc: first_guess", "float64", self.device, cols=3, rows=B)
        _check_dev(results, "plicp_noalloc: results", "uint8", self.device, cols=128, rows=B)
        _check_dev(corr_j1, "plicp_noalloc: corr_j1", "int16", self.device, cols=n, rows=B)
        _check_dev(corr_j2, "plicp_noalloc: corr_j2", "int16", self.device, cols=n, rows=B)
        st = self.L.zhsm_plicp_batch_noalloc(self.h, _ptr(scans), int(scans.shape[0]), _ptr(ref_idx), _ptr(sens_idx),
                                             _ptr(first_guess), int(ref_idx.shape[0]),
                                             C.byref(params) if params is not None else None, _ptr(results),
                                             _ptr(corr_j1), _ptr(corr_j2), stream)
        self._check(st, "zhsm_plicp_batch_noalloc")
        return results

    # ---- zhpsm family (CPolarMatch) -------------------------------------------------------------------------
    def preprocess(self, scans, bad=None, stream=None):
        """pm_preprocessScan over a pool: median filter in place, far points, segmentation.
        Host: numpy float32 [S][n] -> (scans, bad uint8, seg int16).  Device: torch tensors (bad must be given)."""
        S, n = scans.shape
        if _is_dev(scans):
            import torch
            if bad is None:
                bad = torch.zeros((S, n), dtype=torch.uint8, device=scans.device)
            seg = torch.zeros((S, n), dtype=torch.int16, device=scans.device)
            flags = SCANS_ON_DEVICE
        else:
            assert scans.dtype == np.float32 and scans.flags.c_contiguous
            bad = np.zeros((S, n), np.uint8) if bad is None else np.ascontiguousarray(bad, np.uint8)
            seg = np.zeros((S, n), np.int16)
            flags = 0
        self._check(self.L.zhsm_preprocess_batch(self.h, _ptr(scans), _ptr(bad), _ptr(seg), S, flags, stream),
                    "zhsm_preprocess_batch")
        return scans, bad, seg

    def psm(self, scans, bad, seg, ref_idx, act_idx, algo=ALGO_PSM, init_pose=None, results=None, stream=None):
        """Batched pm_psm / pm_icp / pm_psm->pm_icp / FMatch on preprocessed scans (all arrays on one side)."""
        dev = _is_dev(scans)
        dev_jobs = _is_dev(ref_idx)
        B, S = int(ref_idx.shape[0]), int(scans.shape[0])
        flags = (SCANS_ON_DEVICE if dev else 0) | (JOBS_ON_DEVICE if dev_jobs else 0)
        _same_side("psm: scans / bad / seg", scans, bad, seg)
        _same_side("psm: ref_idx / act_idx / init_pose", ref_idx, act_idx, init_pose)
        n = self.laser.n
        if dev:
            _check_dev(scans, "psm: scans", "float32", self.device, cols=n)
            _check_dev(bad, "psm: bad", "uint8", self.device, cols=n, rows=S)
            _check_dev(seg, "psm: seg", "int16", self.device, cols=n, rows=S)
        if dev_jobs:
            _check_dev(ref_idx, "psm: ref_idx", "int32", self.device)
            _check_dev(act_idx, "psm: act_idx", "int32", self.device, rows=B)
            _check_dev(init_pose, "psm: init_pose", "float32", self.device, cols=3, rows=B)
        if results is not None and _is_dev(results):
            _check_dev(results, "psm: results", "uint8", self.device, cols=128, rows=B)
        if not dev_jobs:
            ref_idx = np.ascontiguousarray(ref_idx, np.int32)
            act_idx = np.ascontiguousarray(act_idx, np.int32)
            if init_pose is not None:
                init_pose = np.ascontiguousarray(init_pose, np.float32)
        if results is None:
            results = np.zeros(B, RESULT_DTYPE)
        if _is_dev(results):
            flags |= RESULTS_ON_DEVICE
        self._check(self.L.zhsm_psm_batch(self.h, _ptr(scans), _ptr(bad), _ptr(seg), S, _ptr(ref_idx), _ptr(act_idx),
                                          _ptr(init_pose), B, int(algo), _ptr(results), flags, stream), "zhsm_psm_batch")
        return results

    def is_corridor(self, scans, bad, seg, stream=None):
        """pm_is_corridor per preprocessed scan: int8 [S] = 1 corridor, 0 room, -3 / -2 where the reference throws."""
        S = int(scans.shape[0])
        if _is_dev(scans):
            import torch
            out = torch.empty(S, dtype=torch.int8, device=scans.device)
            flags = SCANS_ON_DEVICE | RESULTS_ON_DEVICE
        else:
            out = np.zeros(S, np.int8)
            flags = 0
        self._check(self.L.zhsm_is_corridor_batch(self.h, _ptr(scans), _ptr(bad), _ptr(seg), S, _ptr(out), flags, stream),
                    "zhsm_is_corridor_batch")
        return out

    def scan_project(self, scans, bad, seg, idx, pose):
        """pm_scan_project of scans[idx[k]] at pose[k] (host arrays) -> (new_r float32 [B][n], new_bad uint8 [B][n])."""
        idx = np.ascontiguousarray(idx, np.int32)
        pose = np.ascontiguousarray(pose, np.float32)
        B, n = idx.shape[0], scans.shape[1]
        new_r = np.zeros((B, n), np.float32)
        new_bad = np.zeros((B, n), np.uint8)
        self._check(self.L.zhsm_scan_project_batch(self.h, _ptr(scans), _ptr(bad), _ptr(seg), scans.shape[0], _ptr(idx),
                                                   _ptr(pose), B, _ptr(new_r), _ptr(new_bad)), "zhsm_scan_project_batch")
        return new_r, new_bad

    def default_icp_options(self, **over):
        o = IcpOptions()
        self.L.zhsm_icp_default_options(C.byref(o))
        for k, v in over.items():
            setattr(o, k, v)
        return o

    def icp(self, pts, off, cnt, ref_idx, cur_idx, init=None, options=None, results=None, stream=None):
        """Batched CICPWarpper::ICPMatch over a cloud pool (see zhsm_icp_batch).  off/cnt are host int32 arrays;
        pts / the job arrays / results may each be numpy (host) or torch CUDA tensors."""
        off = np.ascontiguousarray(off, np.int32)
        cnt = np.ascontiguousarray(cnt, np.int32)
        flags = (SCANS_ON_DEVICE if _is_dev(pts) else 0) | (JOBS_ON_DEVICE if _is_dev(ref_idx) else 0)
        _same_side("icp: ref_idx / cur_idx / init", ref_idx, cur_idx, init)
        if _is_dev(pts):
            _check_dev(pts, "icp: pts", "float32", self.device)
        if _is_dev(ref_idx):
            _check_dev(ref_idx, "icp: ref_idx", "int32", self.device)
            _check_dev(cur_idx, "icp: cur_idx", "int32", self.device, rows=int(ref_idx.shape[0]))
            _check_dev(init, "icp: init", "float64", self.device, cols=3, rows=int(ref_idx.shape[0]))
        if results is not None and _is_dev(results):
            _check_dev(results, "icp: results", "uint8", self.device, cols=128, rows=int(ref_idx.shape[0]))
        if not _is_dev(pts):
            pts = np.ascontiguousarray(pts, np.float32)
        if not _is_dev(ref_idx):
            ref_idx = np.ascontiguousarray(ref_idx, np.int32)
            cur_idx = np.ascontiguousarray(cur_idx, np.int32)
            if init is not None:
                init = np.ascontiguousarray(init, np.float64)
        B = int(ref_idx.shape[0])
        npts = int(pts.shape[0]) // 2
        if results is None:
            results = np.zeros(B, RESULT_DTYPE)
        if _is_dev(results):
            flags |= RESULTS_ON_DEVICE
        self._check(self.L.zhsm_icp_batch(self.h, _ptr(pts), npts, _ptr(off), _ptr(cnt), int(off.size), _ptr(ref_idx),
                                          _ptr(cur_idx), _ptr(init), B, C.byref(options) if options is not None else None,
                                          _ptr(results), flags, stream), "zhsm_icp_batch")
        return results

    def best_of_starts(self, results, P, G, criterion=BEST_GOODNESS, stream=None):
        """Per pair the preferred start (see zhsm_best_of_starts): returns (best index [P], best record [P])."""
        if _is_dev(results):
            import torch
            best = torch.empty(P, dtype=torch.int32, device=results.device)
            rec = torch.empty((P, 128), dtype=torch.uint8, device=results.device)
            flags = RESULTS_ON_DEVICE
        else:
            best = np.zeros(P, np.int32)
            rec = np.zeros(P, RESULT_DTYPE)
            flags = 0
        self._check(self.L.zhsm_best_of_starts(self.h, _ptr(results), int(P), int(G), int(criterion), _ptr(best), _ptr(rec), flags,
                                               stream), "zhsm_best_of_starts")
        return best, rec

    def gate_candidates(self, pose, ld, ru, cur_idx, cov_xy=None, beta=225.0):
        """isOverlappedArea / isLoopyArea flags [M][N] (bit 1 / bit 2) of nodes cur_idx against every node."""
        pose = np.ascontiguousarray(pose, np.float64); ld = np.ascontiguousarray(ld, np.float64)
        ru = np.ascontiguousarray(ru, np.float64); cur_idx = np.ascontiguousarray(cur_idx, np.int32)
        cov = None if cov_xy is None else np.ascontiguousarray(cov_xy, np.float64)
        N, M = pose.shape[0], cur_idx.size
        flags = np.zeros((M, N), np.uint8)
        self._check(self.L.zhsm_gate_candidates(self.h, _ptr(pose), _ptr(ld), _ptr(ru), _ptr(cov), N, _ptr(cur_idx), M,
                                                float(beta), _ptr(flags)), "zhsm_gate_candidates")
        return flags

    def available_area(self, trans, cov_xy=None):
        trans = np.ascontiguousarray(trans, np.float64).reshape(-1, 3)
        cov = None if cov_xy is None else np.ascontiguousarray(cov_xy, np.float64)
        ok = np.zeros(trans.shape[0], np.uint8)
        self._check(self.L.zhsm_available_area(self.h, _ptr(trans), _ptr(cov), trans.shape[0], _ptr(ok)), "zhsm_available_area")
        return ok

    def consensus(self, poses, good, set_off):
        """deleteMisMatch + calculateWeightedPose + calculateFinalPose's choice per matched set."""
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 3)
        good = np.ascontiguousarray(good, np.float32); set_off = np.ascontiguousarray(set_off, np.int32)
        S = set_off.size - 1
        keep = np.zeros(poses.shape[0], np.uint8); fp = np.zeros((S, 3)); fi = np.zeros(S, np.int32)
        self._check(self.L.zhsm_consensus(self.h, _ptr(poses), _ptr(good), _ptr(set_off), S, _ptr(keep), _ptr(fp), _ptr(fi)),
                    "zhsm_consensus")
        return keep, fp, fi

    def cov_to_information(self, cov6, scale=1.0, const_diag=None):
        """matchNodePairICP's information = covariance.inverse() (MapNode.cpp:638-651), or the constant override
        (MapGraph.cpp:1427-1429) when const_diag is given: -> (info [M][3][3], ok [M])."""
        cov6 = np.ascontiguousarray(cov6, np.float64).reshape(-1, 6)
        M = cov6.shape[0]
        info = np.zeros((M, 3, 3)); ok = np.zeros(M, np.uint8)
        cd = None if const_diag is None else np.ascontiguousarray(const_diag, np.float64)
        self._check(self.L.zhsm_cov_to_information(self.h, _ptr(cov6), M, float(scale), _ptr(cd), _ptr(info), _ptr(ok)),
                    "zhsm_cov_to_information")
        return info, ok

    def submap_cloud(self, scans_cm, traj, cap=1 << 20):
        """CMapNode::computePMAP: occupied 5 cm cells of K scans taken at traj [K][3]; returns (x, y, grid dims)."""
        scans = np.ascontiguousarray(scans_cm, np.float32); traj = np.ascontiguousarray(traj, np.float64)
        x = np.zeros(cap, np.float32); y = np.zeros(cap, np.float32)
        cnt = C.c_int(0); dims = np.zeros(2, np.int32)
        self._check(self.L.zhsm_submap_cloud(self.h, _ptr(scans), _ptr(traj), scans.shape[0], _ptr(x), _ptr(y), cap,
                                             C.byref(cnt), _ptr(dims)), "zhsm_submap_cloud")
        if cnt.value > cap:
            raise ZhsmError(f"submap cloud has {cnt.value} points, capacity {cap}")
        return x[:cnt.value].copy(), y[:cnt.value].copy(), dims

    def debug_libm(self, which, x, y=None):
        x = np.ascontiguousarray(x, np.float32)
        y = np.zeros_like(x) if y is None else np.ascontiguousarray(y, np.float32)
        out = np.zeros_like(x)
        self._check(self.L.zhsm_debug_libm(self.h, int(which), _ptr(x), _ptr(y), _ptr(out), x.size), "zhsm_debug_libm")
        return out

    def synth_pairs(self, d_scans, d_truth, first_pair, npairs, dmax_xy_m=0.10, dmax_th_rad=0.0872664626,
                    noise_m=0.01, dropout=0.01, seed=20261104, stream=None):
        """SURVEY 8d C4 generator, on the device: d_scans [2*npairs][n] float32, d_truth [npairs][3] float64."""
        self._check(self.L.zhsm_synth_pairs(self.h, _ptr(d_scans), _ptr(d_truth), int(first_pair), int(npairs),
                                            dmax_xy_m, dmax_th_rad, noise_m, dropout, int(seed), stream),
                    "zhsm_synth_pairs")

    def fp_peak_tflops(self, which):
        return float(self.L.zhsm_fp_peak_tflops(self.h, int(which)))

    def sync(self, stream=None):
        self._check(self.L.zhsm_sync(self.h, stream), "zhsm_sync")

    def launch_count(self):
        return int(self.L.zhsm_launch_count(self.h))

    def last_kernel_ms(self):
        return float(self.L.zhsm_last_kernel_ms(self.h))


class Group:
    """zhsm_group_*: n GPUs of one box driven from this process (one handle per device + an NCCL communicator).
    plicp() is Engine.plicp with host arrays, sharded by contiguous job blocks, the records all-gathered."""

    def __init__(self, devices):
        self.L = load_library()
        devs = np.ascontiguousarray(devices, np.int32)
        g = C.c_void_p()
        st = self.L.zhsm_group_create(devs.ctypes.data, int(devs.size), C.byref(g))
        if st != 0:
            raise ZhsmError(f"zhsm_group_create({list(devices)}) failed with status {st}")
        self.g, self.n, self.laser = g, int(devs.size), None

    def close(self):
        if getattr(self, "g", None):
            self.L.zhsm_group_destroy(self.g)
            self.g = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st, what):
        if st != 0:
            raise ZhsmError(f"{what}: status {st}: {self.L.zhsm_group_last_error(self.g).decode()}")

    def set_laser(self, laser):
        self.laser = laser
        self._check(self.L.zhsm_group_set_laser(self.g, C.byref(laser)), "zhsm_group_set_laser")

    def plicp(self, scans, ref_idx, sens_idx, first_guess=None, params=None, filter_scans=False):
        assert scans.dtype == np.float32 and scans.flags.c_contiguous and scans.shape[1] == self.laser.n
        ri, si = np.ascontiguousarray(ref_idx, np.int32), np.ascontiguousarray(sens_idx, np.int32)
        fg = None if first_guess is None else np.ascontiguousarray(first_guess, np.float64)
        res = np.zeros(ri.size, RESULT_DTYPE)
        self._check(self.L.zhsm_group_plicp_batch(self.g, _ptr(scans), int(scans.shape[0]), _ptr(ri), _ptr(si), _ptr(fg), int(ri.size),
                                                  C.byref(params) if params is not None else None, _ptr(res),
                                                  FILTER_SCANS if filter_scans else 0), "zhsm_group_plicp_batch")
        return res

    def gathered(self, k):
        """(device pointer of GPU k's copy of all records, block, B) of the last batch."""
        block, B = C.c_int(), C.c_int()
        p = self.L.zhsm_group_gathered(self.g, int(k), C.byref(block), C.byref(B))
        return p, block.value, B.value
