"""GPU parity of the zhpsm family (pm_preprocessScan, pm_scan_project, pm_psm, pm_icp, FMatch) against goldens that
were produced by EXECUTING the reference's own sources (oracle/_ref, tools/make_goldens.py): parity is pinned.

Everything on this path is float; the device code mirrors the reference's expression shapes, its summation order and
the host libm's sinf/cosf/atan2f bit for bit, so most results are expected to be IDENTICAL, not merely close.
Tolerance where a difference is allowed: 1e-2 cm (1e-4 m) and 1e-5 rad, the north star's figures."""
import ctypes as C
import zlib

import numpy as np
import pytest

from laser_slam_b200 import engine as E
from tests.support import LOGS, load_log, make_laser

pytestmark = pytest.mark.gpu


def _crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def test_device_libm(engine):
    """csrc/zh_libm.cuh returns the host libm's bits (glibc sinf/cosf/atan2f) on random arguments."""
    m = C.CDLL("libm.so.6")
    for f in (m.sinf, m.cosf):
        f.argtypes = [C.c_float]; f.restype = C.c_float
    m.atan2f.argtypes = [C.c_float, C.c_float]; m.atan2f.restype = C.c_float
    rng = np.random.default_rng(5)
    N = 200000
    x = rng.uniform(-8, 8, N).astype(np.float32)
    x[:1000] = rng.uniform(-1e-4, 1e-4, 1000).astype(np.float32)
    for which, f in ((0, m.sinf), (1, m.cosf)):
        got = engine.debug_libm(which, x)
        want = np.array([f(float(v)) for v in x], np.float32)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), which
    yy = rng.uniform(-6000, 6000, N).astype(np.float32)
    xx = rng.uniform(-6000, 6000, N).astype(np.float32)
    yy[:20000] *= np.float32(1e-3); xx[20000:40000] *= np.float32(1e-3)
    yy[40000:40100] = 0; xx[40100:40200] = 0; xx[40200:40300] = 1
    got = engine.debug_libm(2, yy, xx)
    want = np.array([m.atan2f(float(a), float(b)) for a, b in zip(yy, xx)], np.float32)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.fixture(scope="module")
def pre(engine):
    cache = {}

    def get(log):
        if log not in cache:
            g = load_log(log)
            engine.set_laser(make_laser(g))
            r, bad, seg = engine.preprocess(g["scans"].copy())
            cache[log] = (g, r, bad, seg)
        else:
            engine.set_laser(make_laser(cache[log][0]))
        return cache[log]
    return get


@pytest.mark.parametrize("log", LOGS)
def test_preprocess_bit_exact(pre, log):
    g, r, bad, seg = pre(log)
    for k in range(r.shape[0]):
        got = [_crc(r[k]), _crc(bad[k].astype(np.int32)), _crc(seg[k].astype(np.int32))]
        assert got == [int(v) for v in g["pre_crc"][k]], (log, k)


@pytest.mark.parametrize("log", LOGS)
def test_scan_project(engine, pre, log):
    g, r, bad, seg = pre(log)
    idx = g["proj_idx"]
    pose = g["icp_pose"][idx - 1]
    new_r, new_bad = engine.scan_project(r, bad, seg, idx, pose)
    same_r = (new_r.view(np.uint32) == g["proj_r"].view(np.uint32)).mean()
    same_b = (new_bad.astype(np.int8) == g["proj_bad"]).mean()
    print(f"\n{log}: scan_project bins identical r={same_r:.5f} bad={same_b:.5f}")
    assert same_b == 1.0 and same_r == 1.0, np.argwhere(new_r.view(np.uint32) != g["proj_r"].view(np.uint32))[:10]


@pytest.mark.parametrize("log", LOGS)
@pytest.mark.parametrize("name,algo", [("psm", E.ALGO_PSM), ("icp", E.ALGO_PM_ICP), ("psm_icp", E.ALGO_PSM_ICP)])
def test_matchers_vs_reference_goldens(engine, pre, log, name, algo):
    g, r, bad, seg = pre(log)
    S = r.shape[0]
    res = engine.psm(r, bad, seg, np.arange(S - 1), np.arange(1, S), algo=algo)
    want_pose, want_st = g[name + "_pose"], g[name + "_status"]
    st_eq = (res["status"] != 0).astype(np.int8) == want_st
    ok = want_st == 0
    got = res["x"].astype(np.float32)
    ident = (got.view(np.uint32) == want_pose.view(np.uint32)).all(1)
    d = np.abs(got - want_pose)
    close = (d[:, 0] <= 1e-2) & (d[:, 1] <= 1e-2) & (d[:, 2] <= 1e-5)
    print(f"\n{log} {name}: pairs={S-1} status_eq={st_eq.mean():.5f} identical={ident[ok].mean():.5f} "
          f"within_tol={close[ok].mean():.5f} max|d|={d[ok & st_eq].max(0)}")
    # identical to the reference's own output, bit for bit, throw for throw
    assert st_eq.all(), np.flatnonzero(~st_eq)[:10]
    assert ident[ok].all(), np.flatnonzero(~ident & ok)[:10]


@pytest.mark.parametrize("log", LOGS)
def test_utm_gate_vs_reference_goldens(engine, pre, log):
    """pm_orientation_search's HOKUYO_UTM_30LX branch (ZHPolar_Match.cpp:1182-1191; zhsm_laser.flags =
    ZHSM_LASER_UTM_GATE): pm_psm of every consecutive pair, bit-identical to the compiled reference run with that
    branch on (tests/golden/utm_gate.npz, tools/make_utm_goldens.py).  The gate changes > 90 % of the poses, so a
    missing or misplaced gate cannot pass."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "utm_gate.npz"))
    g, r, bad, seg = pre(log)
    lz = make_laser(g)
    lz.flags = E.LASER_UTM_GATE
    engine.set_laser(lz)
    try:
        S = r.shape[0]
        res = engine.psm(r, bad, seg, np.arange(S - 1), np.arange(1, S), algo=E.ALGO_PSM)
    finally:
        engine.set_laser(make_laser(g))
    want_pose, want_st = z[log + "_pose"], z[log + "_status"]
    got_st = (res["status"] != 0).astype(np.int8)
    assert np.array_equal(got_st, want_st), np.flatnonzero(got_st != want_st)[:10]
    ok = want_st == 0
    got = res["x"].astype(np.float32)
    diff = np.flatnonzero(~(got.view(np.uint32) == want_pose.view(np.uint32)).all(1) & ok)
    assert diff.size == 0, (log, diff[:10])
    assert (np.any(want_pose != g["psm_pose"], 1)).mean() > 0.9   # the fixture really exercises the gate


@pytest.mark.parametrize("log", ["intel-lab"])
def test_fmatch_error_index(engine, pre, log):
    g, r, bad, seg = pre(log)
    S = 400
    res = engine.psm(r, bad, seg, np.arange(S - 1), np.arange(1, S), algo=E.ALGO_FMATCH)
    ok = (g["icp_status"][:S - 1] == 0) & (res["status"] == 0)
    got = np.stack([res["cov"][:, 0], res["cov"][:, 3]], 1).astype(np.float32)
    want = g["err_index"][:S - 1]
    rel = np.abs(got - want) / np.maximum(np.abs(want), 1e-12)
    print(f"\n{log}: pm_error_index2 identical={(got == want)[ok].all(1).mean():.5f} max rel={rel[ok].max():.3g}")
    assert (rel[ok].max(1) < 1e-3).mean() >= 0.999


def test_is_corridor(engine, pre):
    """pm_is_corridor (8a-16) per scan == the reference's own answers (tests/golden/corridor.npz, made by executing
    the reference): every scan of the three logs, plus synthetic corridors / rooms / nearly empty scans that reach the
    corridor branch and the `throw 3` path."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "corridor.npz"))
    for log in LOGS:
        g, r, bad, seg = pre(log)
        got = engine.is_corridor(r, bad, seg)
        assert np.array_equal(got, z[log]), (log, np.nonzero(got != z[log])[0][:10])
    lz = E.laser_from_table("LMS511")
    engine.set_laser(lz)
    r, bad, seg = engine.preprocess(z["synth_scans"].astype(np.float32))
    got = engine.is_corridor(r, bad, seg)
    assert np.array_equal(got, z["synth_flags"]), np.nonzero(got != z["synth_flags"])[0]
    assert set(np.unique(got)) == {-3, 0, 1}
