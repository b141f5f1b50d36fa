"""The C++ CARMEN reader (laser_slam_b200/host/zh_carmen.cpp) against the reference's own
CPolarMatch::readCarmon (compiled into oracle/_ref) on a log file written from the golden fixtures, and against the
real logs where /root/reference exists.  No GPU needed: text parsing is host work."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import pyoracle as po
from tests.support import load_log

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "laser_slam_b200", "libzhsm_host.so")


class Laser(C.Structure):
    _fields_ = [("n", C.c_int32), ("fi_min", C.c_float), ("dfi", C.c_float), ("min_range_cm", C.c_int32),
                ("max_range_cm", C.c_int32), ("min_valid_points", C.c_int32), ("scan_window", C.c_int32),
                ("reserved", C.c_int32)]


@pytest.fixture(scope="module")
def host():
    if not os.path.exists(HOST):
        from laser_slam_b200 import build
        build.build_host()
    L = C.CDLL(HOST)
    L.zhsm_read_carmen.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(Laser)]
    L.zhsm_read_carmen.restype = C.c_int
    return L


def _read_ours(host, path, laser, tag, n, cap=4000):
    scans = np.zeros((cap, n), np.float32); poses = np.zeros((cap, 3)); geom = Laser()
    S = host.zhsm_read_carmen(path.encode(), laser.encode(), tag.encode(), scans.ctypes.data, poses.ctypes.data, cap,
                              C.byref(geom))
    return S, scans[:max(S, 0)], poses[:max(S, 0)], geom


def _read_ref(path, laser, tag, n, cap=4000):
    R = po.ref()
    R.zhpsm_ref_read_carmen.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    R.zhpsm_ref_read_carmen.restype = C.c_int
    scans = np.zeros((cap, n), np.float32); poses = np.zeros((cap, 3), np.float32)
    fi = C.c_float(); mx = C.c_int()
    S = R.zhpsm_ref_read_carmen(laser.encode(), path.encode(), tag.encode(), scans.ctypes.data, poses.ctypes.data, cap,
                                C.byref(fi), C.byref(mx))
    return S, scans[:max(S, 0)], poses[:max(S, 0)], fi.value, mx.value


def _write_log(path, g, K, tag="ROBOTLASER1", short_every=7):
    """A CARMEN file in the layout of the reference's data/*.log: header fields, readings in metres, remission count,
    laser pose, then trailing fields the reader ignores.  Some readings are made shorter than PM_MIN_RANGE."""
    r = g["ranges_cm"][:K].astype(np.float64) / 100.0
    with open(path, "w") as f:
        f.write("# comment line\nPARAM robot_x 1\nODOM 0 0 0 0 0 0 0 host 0\n")
        for k in range(K):
            vals = r[k].copy()
            vals[k % short_every] = 0.05
            lp = g["laser_pose"][k]
            f.write(f"{tag} 0 {float(g['start']):.6f} 3.141593 0.017453 {float(g['max_range_m']):.1f} 0.010000 0 {vals.size} "
                    + " ".join(f"{v:.2f}" for v in vals) + f" 0 {lp[0]:.6f} {lp[1]:.6f} {lp[2]:.6f} 0 0 0 0 0 0 0 0 0 host 0.0\n")
            if k % 5 == 0:
                f.write("FLASER 3 1.0 2.0 3.0 0 0 0 0 0 0 host 0\n")
            if k == 3:
                f.write(f"{tag} 0 0 0 0 0 0 0 0\n")   # a record without beams is skipped (:222)


@pytest.mark.parametrize("log", ["intel-lab", "fr079"])
def test_reader_matches_compiled_reference_on_written_log(tmp_path, host, log):
    if not po.have_ref():
        pytest.skip("oracle/_ref not built")
    g = load_log(log)
    n = int(g["n"])
    path = str(tmp_path / "log.txt")
    _write_log(path, g, 120)
    S, scans, poses, geom = _read_ours(host, path, g["laser_name"], "ROBOTLASER1", n)
    Sr, rscans, rposes, fi, mx = _read_ref(path, g["laser_name"], "ROBOTLASER1", n)
    assert S == Sr == 120
    assert np.array_equal(scans, rscans)
    assert np.array_equal(poses.astype(np.float32), rposes)
    assert geom.fi_min == fi and geom.max_range_cm == mx and geom.n == n
    assert (scans == mx + 1).sum() >= 120                       # the short readings and padded beams


def test_reader_failures(tmp_path, host):
    g = load_log("fr079")
    n = int(g["n"])
    assert _read_ours(host, str(tmp_path / "missing.log"), "LMS511", "ROBOTLASER1", n)[0] == -1
    path = str(tmp_path / "log.txt")
    _write_log(path, g, 10)
    assert _read_ours(host, path, "LMS511", "NOSUCHTAG", n)[0] == -1       # no record matched
    assert _read_ours(host, path, "LMS211", "ROBOTLASER1", 181)[0] == -1   # 360 beams do not fit a 181-beam laser
    assert _read_ours(host, path, "LMS999", "ROBOTLASER1", n)[0] == -1


@pytest.mark.parametrize("log", ["intel-lab", "fr079", "mit-cscail"])
def test_reader_on_the_real_logs(host, log):
    path = f"/root/reference/data/{log}.log"
    if not os.path.exists(path):
        pytest.skip("reference data not present on this machine")
    g = load_log(log)
    S, scans, poses, geom = _read_ours(host, path, g["laser_name"], "ROBOTLASER1", int(g["n"]))
    assert S == g["scans"].shape[0]
    assert np.array_equal(scans, g["scans"])
    assert np.allclose(poses, g["laser_pose"], atol=1e-9)
    assert geom.fi_min == g["fi_min"] and geom.max_range_cm == int(g["max_range"])
