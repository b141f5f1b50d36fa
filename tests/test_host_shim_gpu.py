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
        assert (~ok).sum() <= _golden_mismatch_budget(log), (mode, np.flatnonzero(~ok).tolist())   # 1e-4 m = 1e-2 cm
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


# ---------------------------------------------------------------------------------------------------------------
# the CSM-shaped C entry (include/csm_compat.h): sm_icp / ld_alloc_new / ld_free from plain C, and the REFERENCE'S OWN
# CCanonicalMatcher (compiled unmodified from /root/reference) linked against libzhsm.so -- "link, change nothing"
# ---------------------------------------------------------------------------------------------------------------
def _golden_mismatch_budget(log):
    # pairs on which the CUDA path and the oracle differ by a documented floating-point near-tie:
    # measured 1 of 2671 on intel-lab, none on the other logs
    return {"intel-lab": 1}.get(log, 0)


@pytest.mark.parametrize("log", ["intel-lab", "fr079", "mit-cscail"])
def test_plain_c_sm_icp_matches_goldens(tmp_path, log):
    """tests/cpp/csm_plain.c does what ZHCanonical_Matcher.cpp:193-285 does (two ld_alloc_new, fill, sm_icp, read
    valid / x / iterations) and links against libzhsm.so only; every consecutive pair of the log against the goldens."""
    from oracle import pyoracle as po
    exe = os.path.join(ROOT, "tests", "cpp", "_build", "csm_plain")
    assert os.path.exists(exe), "tests/cpp/_build/csm_plain is not built (run __graft_entry__.build())"
    g = load_log(log)
    scans = g["scans"].copy()
    po.median_filter_batch(scans, 4)
    S = scans.shape[0]
    fin, fout = tmp_path / "scans.f32", tmp_path / "out.txt"
    scans.tofile(fin)
    subprocess.run([exe, str(int(g["n"])), repr(float(g["fi_min"])), repr(float(g["dfi"])), str(int(g["min_range"])),
                    str(int(g["max_range"])), str(fin), str(S), "0", str(fout)], check=True)
    out = np.loadtxt(fout).reshape(S - 1, 9)
    valid = out[:, 0].astype(int)
    assert np.array_equal(valid, g["plicp_valid"])
    both = valid == 1
    bad = ~(pose_close_np(out[:, 1:4], g["plicp_x"]) | ~both) | (out[:, 4].astype(int) != g["plicp_iterations"]) | \
        ((out[:, 7].astype(np.uint32) != g["plicp_hash"]) & both) | ((out[:, 5].astype(int) != g["plicp_nvalid"]) & both)
    print(log, "sm_icp vs goldens: mismatching pairs", np.flatnonzero(bad).tolist())
    assert bad.sum() <= _golden_mismatch_budget(log), np.flatnonzero(bad).tolist()
    # the LDP's output side: laser_sens->corr[] holds the final correspondences (of the best loop, or recomputed without
    # the outlier passes after a restart; sm_result.nvalid is the first loop's count, so the two are not comparable)
    assert (out[both, 8] >= 0.05 * int(g["n"])).all() and (out[~both, 8] == 0).all()


def pose_close_np(a, b):
    d = np.abs(np.asarray(a) - np.asarray(b))
    return (d[:, 0] <= 1e-4) & (d[:, 1] <= 1e-4) & (d[:, 2] <= 1e-5)


@pytest.mark.parametrize("log,S", [("intel-lab", 400), ("fr079", 200)])
def test_reference_wrapper_unmodified_on_libzhsm(tmp_path, log, S, engine):
    """The reference's own src/zhcsm/ZHCanonical_Matcher.cpp, compiled UNMODIFIED against include/csm/*.h and linked
    against libzhsm.so (tests/cpp/ref_zhcsm_driver.cpp; built where /root/reference exists, the binary travels):
    FMatch, FMatchKeyFrame and FMatchKeyFrame2 reproduce the goldens in the units the reference writes back."""
    exe = os.path.join(ROOT, "tests", "cpp", "_build", "ref_zhcsm_on_zhsm")
    if not os.path.exists(exe):
        pytest.skip("built only where /root/reference exists (laser_slam_b200.build.build_host)")
    g = load_log(log)
    raw = np.ascontiguousarray(g["scans"][:S], np.float32)
    fin = tmp_path / "scans.f32"
    raw.tofile(fin)
    want_ret = np.where(g["plicp_valid"][:S - 1] == 1, 0.0, -1.0)
    want = (g["plicp_x"][:S - 1] * [100.0, 100.0, 1.0]).astype(np.float32)
    outs = {}
    for mode, cols in ((0, 4), (1, 4), (2, 10)):
        fout = tmp_path / f"ref{mode}.txt"
        subprocess.run([exe, g["laser_name"], repr(float(g["fi_min"])), str(int(g["max_range"])), str(fin), str(S),
                        str(mode), str(fout)], check=True, stdout=subprocess.DEVNULL)
        out = np.loadtxt(fout).reshape(-1, cols)
        if mode >= 1:
            assert out[0, 0] == 2 and np.all(out[0, 1:4] == 0)  # first key frame (:318-321)
            out = out[1:]
        assert out.shape[0] == S - 1
        assert np.array_equal(out[:, 0], want_ret)
        d = np.abs(out[:, 1:4] - want)
        ok = (d[:, 0] <= 1e-2 + 1e-5 * np.abs(want[:, 0])) & (d[:, 1] <= 1e-2 + 1e-5 * np.abs(want[:, 1])) & (d[:, 2] <= 1e-5)
        assert (~ok).sum() <= _golden_mismatch_budget(log), (mode, np.flatnonzero(~ok).tolist())
        outs[mode] = out
    assert np.array_equal(outs[0], outs[1])
    # FMatchKeyFrame2's covariance == the batched engine's record on the same pairs (same kernel, other entry)
    import laser_slam_b200 as zh
    engine.set_laser(make_laser_local(g))
    scans = engine.median_filter(raw.copy())
    p = engine.default_plicp_params()
    p.do_compute_covariance = 1
    res = engine.plicp(scans, np.arange(S - 1), np.arange(1, S), params=p)
    okv = res["valid"] == 1
    # (the covariance sums are FP64 shared-memory atomicAdds, whose order differs from launch to launch: 1e-6 relative,
    # the tolerance of test_plicp_covariance_matches_oracle)
    assert np.allclose(outs[2][okv, 4:10], res["cov"][okv], rtol=1e-6, atol=1e-12)


def make_laser_local(g):
    from tests.support import make_laser
    return make_laser(g)
