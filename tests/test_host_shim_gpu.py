"""GPU test of the C++ host shim (laser_slam_b200/host/ZHCanonical_Matcher.h): the reference-shaped class driven the
way the reference's drivers drive it (consecutive-scan FMatch, FMatchKeyFrame, and the batch extension), compared with
the oracle goldens in the units the reference writes back (cm, cm, rad as floats)."""
import os
import subprocess

import numpy as np
import pytest

from tests.support import load_log

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "_build", "shim_runlog")


def _run(tmp_path, g, S, mode):
    raw = np.ascontiguousarray(g["scans"][:S], np.float32)
    fin, fout = tmp_path / "scans.f32", tmp_path / f"out{mode}.txt"
    raw.tofile(fin)
    assert os.path.exists(EXE), "tests/cpp/_build/shim_runlog is not built (run __graft_entry__.build())"
    subprocess.run([EXE, g["laser_name"], repr(float(g["fi_min"])), str(int(g["max_range"])), str(fin), str(S),
                    str(mode), str(fout)], check=True, stdout=subprocess.DEVNULL)
    return np.loadtxt(fout).reshape(-1, 4)


@pytest.mark.parametrize("log,S", [("intel-lab", 300), ("fr079", 120)])
def test_shim_matches_goldens(tmp_path, log, S):
    g = load_log(log)
    want_ret = np.where(g["plicp_valid"][:S - 1] == 1, 0.0, -1.0)
    want = (g["plicp_x"][:S - 1] * [100.0, 100.0, 1.0]).astype(np.float32)
    ok_all = []
    for mode in (0, 1, 2):
        out = _run(tmp_path, g, S, mode)
        if mode == 1:
            assert out[0, 0] == 2 and np.all(out[0, 1:] == 0)  # first key frame (ZHCanonical_Matcher.cpp:318-321)
            out = out[1:]
        assert out.shape[0] == S - 1
        assert np.array_equal(out[:, 0], want_ret)
        d = np.abs(out[:, 1:] - want)
        ok = (d[:, 0] <= 1e-2 + 1e-5 * np.abs(want[:, 0])) & (d[:, 1] <= 1e-2 + 1e-5 * np.abs(want[:, 1])) & (d[:, 2] <= 1e-5)
        ok_all.append(ok.mean())
        assert ok.mean() >= 0.995, (mode, ok.mean())   # 1e-4 m = 1e-2 cm
    print(log, "shim pose agreement per mode", ok_all)


# ---------------------------------------------------------------------------------------------------------------
# CPolarMatch shim (laser_slam_b200/host/ZHPolar_Match.h) against goldens produced by EXECUTING the reference
# ---------------------------------------------------------------------------------------------------------------
def _run_cols(tmp_path, g, S, mode, cols):
    raw = np.ascontiguousarray(g["scans"][:S], np.float32)
    fin, fout = tmp_path / "scans.f32", tmp_path / f"out{mode}.txt"
    raw.tofile(fin)
    subprocess.run([EXE, g["laser_name"], repr(float(g["fi_min"])), str(int(g["max_range"])), str(fin), str(S),
                    str(mode), str(fout)], check=True, stdout=subprocess.DEVNULL)
    return np.loadtxt(fout).reshape(-1, cols)


@pytest.mark.parametrize("log,S", [("fr079", 200), ("intel-lab", 200)])
def test_polar_shim_matches_reference_goldens(tmp_path, log, S):
    g = load_log(log)
    # mode 5: pm_preprocessScan on every scan, then pm_psm and pm_icp from zero poses (testAccuracy, :403-428)
    out = _run_cols(tmp_path, g, S, 5, 10)
    for col, name in ((0, "psm"), (5, "icp")):
        threw = out[:, col]
        pose = out[:, col + 2:col + 5].astype(np.float32)
        st = g[name + "_status"][:S - 1]
        assert np.array_equal(threw != 0, st != 0), name           # `throw 1` exactly where the reference throws
        ok = st == 0
        assert np.array_equal(pose[ok], g[name + "_pose"][:S - 1][ok]), name   # bit-identical floats
        assert np.all(pose[~ok] == 0)                              # a throw leaves lsa's pose untouched
    # mode 3 / 4: FMatch sequential (bFirst on the first pair) and the batch extension
    seq = _run_cols(tmp_path, g, S, 3, 5)
    bat = _run_cols(tmp_path, g, S, 4, 5)
    assert np.array_equal(seq, bat)
    ok = g["icp_status"][:S - 1] == 0
    assert np.array_equal(seq[ok, 2:5].astype(np.float32), g["icp_pose"][:S - 1][ok])
    assert np.all(seq[~ok, :2] == -1)
    want = g["err_index"][:S - 1][ok]
    got = seq[ok, :2].astype(np.float32)
    assert (np.abs(got - want) <= 1e-3 * np.abs(want)).all(1).mean() >= 0.999


# ---------------------------------------------------------------------------------------------------------------
# CICPWarpper shim (laser_slam_b200/host/ZHIcp_Warpper.h) on the reference's wrapper-test clouds
# ---------------------------------------------------------------------------------------------------------------
def test_icp_shim(tmp_path):
    d = np.load(os.path.join(ROOT, "tests", "golden", "zhicp_nodes.npz"))
    exe = os.path.join(ROOT, "tests", "cpp", "_build", "shim_icp")
    assert os.path.exists(exe), "tests/cpp/_build/shim_icp is not built (run __graft_entry__.build())"
    P, G = 12, 8
    rng = np.random.default_rng(5)
    grid = np.concatenate([np.zeros((1, 3)), rng.uniform(-1, 1, (G - 1, 3)) * [0.3, 0.3, 0.15]])
    fin, fout = tmp_path / "pool.bin", tmp_path / "icp.txt"
    with open(fin, "wb") as f:
        np.array([d["off"].size, P, G], np.int32).tofile(f)
        d["off"].astype(np.int32).tofile(f)
        d["cnt"].astype(np.int32).tofile(f)
        d["pairs"][:P].astype(np.int32).tofile(f)
        d["init"][:P].astype(np.float64).tofile(f)
        grid.astype(np.float64).tofile(f)
        d["pts"].astype(np.float32).tofile(f)
    subprocess.run([exe, str(fin), str(fout)], check=True)
    lines = open(fout).read().strip().split("\n")
    single = np.array([[float(v) for v in ln.split()] for ln in lines[0::2]])
    multi = np.array([[float(v) for v in ln.split()] for ln in lines[1::2]])
    o = d["oracle"][:P]
    assert np.allclose(single[:, 1:4], o["pose"], atol=1e-4)
    assert np.array_equal(single[:, 4].astype(int), o["nIterations"])
    assert np.allclose(single[:, 0], o["goodness"], atol=1e-6) and np.array_equal(single[:, 0], single[:, 5])
    cov = o["cov"].reshape(-1, 3, 3)
    tri = np.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], 1)
    assert np.allclose(single[:, 7:13], tri, rtol=1e-6)
    # start 0 of the multi-start run is the single run, so the best of G can only be at least as good
    assert np.all(multi[:, 0] >= single[:, 0] - 1e-7)
