"""CPU tests of the checkers under oracle/ against the committed golden fixtures (tests/golden/*.npz).

- oracle/_ref/libzhpsm_ref.so is the reference's own zhpsm source compiled unmodified: it pins the median filter
  (CCanonicalMatcher::median_filter's twin) and the zhpsm goldens.
- oracle/plicp_oracle.c is a restatement of CSM sm_icp ("parity unpinned": CSM is not vendored); it is pinned
  here against (1) the committed goldens, (2) the logs' ground-truth laser poses through the reference's own
  testAccuracy criterion (src/zhpsm/ZHPolar_Match.cpp:364-447), (3) the compiled zhpsm matcher on the same
  pairs, (4) internal consistency checks CSM itself offers (debug_verify_tricks: tricks == exhaustive search).
"""
import zlib

import numpy as np
import pytest

from oracle import pyoracle as po
from tests.support import LOGS, load_log


@pytest.fixture(scope="module", autouse=True)
def _built():
    po.build()


def _crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def _params(g):
    return float(g["fi_min"]), float(g["dfi"]), int(g["min_range"]), int(g["max_range"])


@pytest.mark.parametrize("log", LOGS)
def test_median_filter_twin_of_reference(log):
    """orc_median_filter == the compiled CPolarMatch::pm_median_filter (bit exact) and both reproduce the
    committed crc32 of the preprocessed ranges."""
    g = load_log(log)
    idx = np.linspace(0, g["scans"].shape[0] - 1, 60).astype(int)
    m = po.RefPolarMatch(g["laser_name"], fi_min=g["fi_min"], max_range_cm=int(g["max_range"])) if po.have_ref() else None
    for k in idx:
        f = po.median_filter(g["scans"][k])
        assert _crc(f) == int(g["pre_crc"][k, 0])
        if m is not None:
            assert np.array_equal(f, m.median(g["scans"][k]))


def test_median_filter_edge_cases():
    # in-place recurrence differs from an out-of-place filter: a single spike next to a step
    r = np.array([1, 9, 1, 1, 9, 9, 1, 9, 9, 9], np.float32)
    f = po.median_filter(r)
    # hand evaluation of ZHCanonical_Matcher.cpp:159-189 (window uses already-written values for i-2, i-1)
    want = r.copy()
    n = len(r)
    for i in range(n):
        w = []
        for j in range(i - 2, i + 3):
            l = min(max(j, 0), n - 1)
            w.append(want[l])
        want[i] = sorted(w)[2]
    assert np.array_equal(f, want)
    # n smaller than the window
    for n in (1, 2, 3, 4):
        r = np.arange(n, 0, -1).astype(np.float32)
        want = r.copy()
        for i in range(n):
            want[i] = sorted(want[min(max(j, 0), n - 1)] for j in range(i - 2, i + 3))[2]
        assert np.array_equal(po.median_filter(r), want)


@pytest.mark.parametrize("log", LOGS)
def test_plicp_oracle_reproduces_goldens(log):
    g = load_log(log)
    S = g["scans"].shape[0]
    filt = np.stack([po.median_filter(s) for s in g["scans"]])
    idx = np.unique(np.concatenate([np.arange(0, 60), np.linspace(60, S - 2, 240).astype(int)]))
    fi_min, dfi, mn, mx = _params(g)
    o = po.zhcsm_fmatch_batch(filt, idx, idx + 1, fi_min, dfi, mn, mx, nthreads=8)
    assert np.array_equal(o["valid"], g["plicp_valid"][idx])
    assert np.array_equal(o["iterations"], g["plicp_iterations"][idx])
    assert np.array_equal(o["corr_hash"], g["plicp_hash"][idx])
    assert np.array_equal(o["nvalid"], g["plicp_nvalid"][idx])
    assert np.array_equal(o["n_loops"], g["plicp_loops"][idx])
    assert np.array_equal(o["x"], g["plicp_x"][idx])  # same binary, same flags: bit identical
    assert np.array_equal(o["error"], g["plicp_error"][idx])


def _rel(p0, p1):
    """pose of p1 in the frame of p0 (OrientedPoint2D::ominus, src/zhpsm/point.cpp:31-69)"""
    d = p1 - p0
    c, s = np.cos(p0[:, 2]), np.sin(p0[:, 2])
    th = np.arctan2(np.sin(d[:, 2]), np.cos(d[:, 2]))
    return np.stack([c * d[:, 0] + s * d[:, 1], -s * d[:, 0] + c * d[:, 1], th], 1)


@pytest.mark.parametrize("log", LOGS)
def test_plicp_goldens_vs_ground_truth(log):
    """testAccuracy criterion (ZHPolar_Match.cpp:364-447): dis_err < 1 m and angle_err < 0.6982 rad counts as a
    success.  The PL-ICP restatement must succeed at least as often as the compiled reference pm_icp on the same
    pairs and its median error vs the log's laser poses must be of the same order."""
    g = load_log(log)
    gt = _rel(g["laser_pose"][:-1], g["laser_pose"][1:])
    v = g["plicp_valid"] == 1
    assert v.mean() > 0.99
    e = np.abs(g["plicp_x"] - gt)
    e[:, 2] = np.abs(np.arctan2(np.sin(e[:, 2]), np.cos(e[:, 2])))
    dis = np.hypot(e[:, 0], e[:, 1])
    ok_pl = v & (dis < 1.0) & (e[:, 2] < 0.6982)
    # compiled reference pm_icp on the same pairs (cm -> m)
    ri = g["icp_pose"].astype(np.float64) * [0.01, 0.01, 1.0]
    er = np.abs(ri - gt)
    er[:, 2] = np.abs(np.arctan2(np.sin(er[:, 2]), np.cos(er[:, 2])))
    ok_ref = (g["icp_status"] == 0) & (np.hypot(er[:, 0], er[:, 1]) < 1.0) & (er[:, 2] < 0.6982)
    # gross failures (the testAccuracy criterion) within 1.5 % of the reference matcher's rate ...
    assert ok_pl.mean() >= ok_ref.mean() - 0.015, (ok_pl.mean(), ok_ref.mean())
    assert ok_pl.mean() > 0.97
    # ... and the bulk of the error distribution no worse than the compiled reference matcher's
    dr = np.hypot(er[:, 0], er[:, 1])
    med_pl = np.median(dis[v])
    med_ref = np.median(dr[g["icp_status"] == 0])
    assert med_pl <= med_ref, (med_pl, med_ref)
    assert (v & (dis < 0.05)).mean() >= ((g["icp_status"] == 0) & (dr < 0.05)).mean()
    # and the two matchers agree with each other to a few cm on the bulk of pairs
    both = v & (g["icp_status"] == 0)
    dd = np.hypot(*(g["plicp_x"][both, :2] - ri[both, :2]).T)
    assert np.median(dd) < 0.03


@pytest.mark.parametrize("log", ["intel-lab", "fr079"])
def test_tricks_vs_exhaustive_search(log):
    """find_correspondences_tricks is NOT an exact nearest-neighbour search: its jump tables may hop over a closer
    point, and where the walk starts (last_best of the previous ray) decides whether they do.  use_corr_tricks=2
    selects an exhaustive search in the restatement: the two agree on the large majority of pairs (measured: 98.9 -
    99.6 % of all pairs of the three logs) but not on all of them -- which is why the CUDA path reproduces the
    sequential tricks walk instead of using a parallel exact search ("correspondence search")."""
    g = load_log(log)
    filt = np.stack([po.median_filter(s) for s in g["scans"][:201]])
    fi_min, dfi, mn, mx = _params(g)
    idx = np.arange(200)
    a = po.zhcsm_fmatch_batch(filt, idx, idx + 1, fi_min, dfi, mn, mx, nthreads=8)
    b = po.zhcsm_fmatch_batch(filt, idx, idx + 1, fi_min, dfi, mn, mx,
                              params=po.default_params(mn, mx, use_corr_tricks=2), nthreads=8)
    assert np.array_equal(a["valid"], b["valid"])
    same = a["corr_hash"] == b["corr_hash"]
    assert same.mean() >= 0.95
    # identical correspondences in every iteration give bit-identical poses; a different final hash usually means
    # a different trajectory, but the median difference stays far below the matcher's accuracy
    d = np.abs(a["x"] - b["x"])
    assert np.median(d.max(1)) < 1e-6


def test_plicp_oracle_first_guess_and_failure_modes():
    g = load_log("intel-lab")
    filt = np.stack([po.median_filter(s) for s in g["scans"][:3]])
    fi_min, dfi, mn, mx = _params(g)
    base = po.zhcsm_fmatch(filt[0], filt[1], fi_min, dfi, mn, mx, want_corr=True)
    assert base["ret"] == 0 and base["valid"] == 1
    # FMatch writes lsa->rx,ry in cm (ZHCanonical_Matcher.cpp:242-244)
    assert np.allclose(base["pose_cm"][:2], base["x"][:2] * 100, atol=1e-3)
    assert ((base["j1"] >= 0) == (base["j2"] >= 0)).all()
    # a first guess near the answer converges to (nearly) the same pose
    o = po.zhcsm_fmatch(filt[0], filt[1], fi_min, dfi, mn, mx, first_guess=base["x"] + [0.02, -0.02, 0.01])
    assert o["valid"] == 1 and np.abs(o["x"] - base["x"]).max() < 5e-3
    # an all-invalid scan fails like the reference: valid == 0, FMatch returns -1 (:275-283)
    dead = np.full_like(filt[0], mx + 1)
    o = po.zhcsm_fmatch(filt[0], dead, fi_min, dfi, mn, mx)
    assert o["valid"] == 0 and o["ret"] == -1
    o = po.zhcsm_fmatch(dead, filt[1], fi_min, dfi, mn, mx)
    assert o["valid"] == 0 and o["ret"] == -1
    # identical scans: zero motion
    o = po.zhcsm_fmatch(filt[0], filt[0], fi_min, dfi, mn, mx)
    assert o["valid"] == 1 and np.abs(o["x"]).max() < 1e-6


def test_quartic_greatest_real_root():
    """poly_greatest_real_root of gpc.c uses gsl_poly_complex_solve; the restatement brackets roots instead.
    Compare with numpy's companion-matrix eigenvalues (the same method GSL uses)."""
    rng = np.random.default_rng(7)
    import ctypes as C
    for _ in range(300):
        roots = rng.uniform(-50, 50, 4)
        if rng.random() < 0.5:  # one complex pair
            re, im = rng.uniform(-20, 20), rng.uniform(0.1, 20)
            poly = np.poly([roots[0], roots[1]])
            poly = np.polymul(poly, [1, -2 * re, re * re + im * im])
            want = max(roots[0], roots[1])
        else:
            poly = np.poly(roots)
            want = roots.max()
        poly = poly * rng.uniform(0.1, 10)
        a = np.ascontiguousarray(poly[::-1], np.float64)  # a[0] + a[1] x + ...
        out = C.c_double()
        ok = po.lib().orc_poly_greatest_real_root(a.ctypes.data, C.cast(C.byref(out), C.c_void_p))
        assert ok == 1
        assert abs(out.value - want) <= 1e-7 * max(1.0, abs(want)), (out.value, want)
    # no real root -> failure like gpc_solve
    a = np.ascontiguousarray(np.poly([1j, -1j, 2j, -2j]).real[::-1], np.float64)
    out = C.c_double()
    assert po.lib().orc_poly_greatest_real_root(a.ctypes.data, C.cast(C.byref(out), C.c_void_p)) == 0


@pytest.mark.skipif(not po.have_ref(), reason="oracle/_ref not built (reference absent)")
@pytest.mark.parametrize("log", LOGS)
def test_compiled_reference_reproduces_zhpsm_goldens(log):
    """The zhpsm goldens are outputs of the reference itself; re-running the compiled reference on a subset
    must reproduce them bit for bit (guards the fixtures and the driver)."""
    g = load_log(log)
    m = po.RefPolarMatch(g["laser_name"], fi_min=g["fi_min"], max_range_cm=int(g["max_range"]))
    S = g["scans"].shape[0]
    idx = np.linspace(0, S - 2, 40).astype(int)
    need = np.unique(np.concatenate([idx, idx + 1]))
    pre = {k: m.preprocess(g["scans"][k]) for k in need}
    for k in need:
        assert [_crc(pre[k][0]), _crc(pre[k][1]), _crc(pre[k][2])] == [int(v) for v in g["pre_crc"][k]]
    for name, algo in (("psm", 0), ("icp", 1), ("psm_icp", 2)):
        for k in idx:
            st, pose, _ = m.match(algo, pre[k], pre[k + 1])
            assert st == int(g[name + "_status"][k])
            if st == 0:
                assert np.array_equal(pose, g[name + "_pose"][k]), (name, k)


@pytest.mark.skipif(not po.have_ref(), reason="oracle/_ref not built (reference absent)")
def test_compiled_reference_reproduces_utm_gate_goldens():
    """tests/golden/utm_gate.npz = the compiled reference with pm_laser_name "HOKUYO_UTM_30LX" (the name switches on
    the delta < PM_MAX_ERROR gate of pm_orientation_search, ZHPolar_Match.cpp:1182-1191); a subset re-run must
    reproduce it bit for bit."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "utm_gate.npz"))
    g = load_log("fr079")
    m = po.RefPolarMatch(g["laser_name"], fi_min=g["fi_min"], max_range_cm=int(g["max_range"]))
    idx = np.linspace(0, g["scans"].shape[0] - 2, 30).astype(int)
    pre = {k: m.preprocess(g["scans"][k]) for k in np.unique(np.concatenate([idx, idx + 1]))}
    m.set_laser_name("HOKUYO_UTM_30LX")
    for k in idx:
        st, pose, _ = m.match(0, pre[k], pre[k + 1])
        assert st == int(z["fr079_status"][k])
        if st == 0:
            assert np.array_equal(pose, z["fr079_pose"][k]), k


# ---------------------------------------------------------------------------------------------------------------
# zhicp oracle (oracle/icp_oracle.c): PARITY UNPINNED (MRPT 0.9.6 absent) -- pinned against regressions by the frozen
# outputs on the reference's own wrapper-test clouds, and checked for the properties the algorithm must have.
# ---------------------------------------------------------------------------------------------------------------
def _icp_nodes():
    import os
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "zhicp_nodes.npz"))


def test_icp_oracle_matches_frozen_outputs():
    d = _icp_nodes()
    r = po.mrpt_icp_batch(d["pts"], d["off"], d["cnt"], d["pairs"][:, 0], d["pairs"][:, 1], d["init"], nthreads=4)
    g = d["oracle"]
    assert np.array_equal(r["nIterations"], g["nIterations"])
    assert np.array_equal(r["nCorrespondences"], g["nCorrespondences"])
    assert np.allclose(r["pose"], g["pose"], rtol=0, atol=1e-12)
    assert np.allclose(r["goodness"], g["goodness"]) and np.allclose(r["quality"], g["quality"], atol=1e-6)


def test_icp_oracle_goodness_consistent_with_reference_logging_rule():
    """src/mapGraph/MapGraph.cpp:1385-1394 logs a pair to data/matched_nodes.log only when the reference's own ICP
    returned 0.4 < goodness < 0.8 (ICP_QUALITY_THRESHOLD, :42); data2's pairs were logged on acceptance (:1586-1590).
    The restatement, run on those very clouds and guesses, must land in the same bands for almost all of them --
    the only behavioural evidence about MRPT's numbers that the reference tree holds."""
    d = _icp_nodes()
    g = d["oracle"]["goodness"]
    names = d["names"][d["pairs"][:, 0]]
    first = np.array([n.startswith("data/") for n in names])
    rejected, accepted = g[first], g[~first]
    assert ((rejected > 0.4) & (rejected < 0.8)).mean() >= 0.9, rejected
    assert (accepted >= 0.8).all(), accepted


def test_icp_oracle_recovers_a_known_transform():
    rng = np.random.default_rng(3)
    # scattered landmarks (point-to-point ICP slides along evenly sampled walls, so walls alone cannot pin a pose)
    x = rng.uniform(0, 6, 400).astype(np.float32)
    y = rng.uniform(0, 6, 400).astype(np.float32)
    for tx, ty, th in ((0.12, -0.08, 0.05), (-0.2, 0.15, -0.08), (0, 0, 0)):
        c, s = np.cos(th), np.sin(th)
        # current cloud = map seen from pose (tx,ty,th): p_cur = R^T (p_map - t)
        cx = (c * (x - tx) + s * (y - ty)).astype(np.float32)
        cy = (-s * (x - tx) + c * (y - ty)).astype(np.float32)
        pts, off, cnt = po.pack_clouds([(x, y), (cx, cy)])
        r = po.mrpt_icp_batch(pts, off, cnt, [0], [1], np.zeros((1, 3)),
                              po.icp_default_options(corresponding_points_decimation=1))[0]
        assert abs(r["pose"][0] - tx) < 5e-3 and abs(r["pose"][1] - ty) < 5e-3 and abs(r["pose"][2] - th) < 2e-3, r
        assert r["goodness"] > 0.95 and 1 < r["nIterations"] <= 100
        assert np.all(np.linalg.eigvalsh(r["cov"].reshape(3, 3)) > 0)


def test_icp_oracle_threshold_schedule():
    """The loop shrinks the gate by ALFA each time the pose stops moving and ends when it falls under
    smallestThresholdDist: identical clouds need exactly ceil(log(0.05/0.75)/log(0.5)) = 4 iterations at
    zero guess (every iteration is a no-move), and maxIterations is a hard cap only while the gate is small."""
    a = np.arange(0, 4, 0.05, dtype=np.float32)
    pts, off, cnt = po.pack_clouds([(a, a * 0.3), (a, a * 0.3)])
    r = po.mrpt_icp_batch(pts, off, cnt, [0], [1], np.zeros((1, 3)),
                          po.icp_default_options(corresponding_points_decimation=1))[0]
    assert r["nIterations"] == 4 and r["goodness"] == 1.0 and np.allclose(r["pose"], 0, atol=1e-7)
    r = po.mrpt_icp_batch(pts, off, cnt, [0], [1], np.zeros((1, 3)),
                          po.icp_default_options(corresponding_points_decimation=1, maxIterations=2))[0]
    # after 2 iterations the gate (0.1875) is still above 0.05: the tail rule keeps halving and iterating
    assert r["nIterations"] > 2


# ---------------------------------------------------------------------------------------------------------------
# loop-closure gating / consensus oracle (oracle/graph_oracle.c)
# ---------------------------------------------------------------------------------------------------------------
def test_graph_oracle_oplus_pinned_by_compiled_reference():
    """The one routine of this path that compiles from the reference (OrientedPoint2D::oplus, point.cpp:63-69) agrees
    bit for bit with the restatement's copy: a node with an identity pose and unit box overlaps itself, and shifting
    the box corners through oplus by hand reproduces the flags."""
    if not po.have_ref():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(0)
    for _ in range(200):
        pose = rng.uniform(-10, 10, 3); pt = rng.uniform(-5, 5, 2)
        c, s = np.cos(pose[2]), np.sin(pose[2])
        want = np.array([pose[0] + pt[0] * c - pt[1] * s, pose[1] + pt[0] * s + pt[1] * c])
        assert np.array_equal(po.ref_oplus_point(pose, pt), want)


def test_graph_oracle_gating_properties():
    from tests.graph_cases import node_table
    pose, ld, ru, cov = node_table()
    N = pose.shape[0]
    f = po.gate_candidates(pose, ld, ru, np.arange(N))
    ov, lp = (f & 1) != 0, (f & 2) != 0
    good_box = ld[:, 0] < ru[:, 0]
    assert ov[np.arange(N), np.arange(N)][good_box].all()           # a node overlaps itself (ratio 1)
    # a node that never got a point keeps (MAX_VALUE, MIN_VALUE) corners; calBounding (:946-955) sorts them, so it turns
    # into a huge box: it contains every proper node, and no proper node covers 40 % of it
    assert ov[good_box][:, ~good_box].all() and not ov[~good_box][:, good_box].any()
    d2 = ((pose[:, None, :2] - pose[None, :, :2]) ** 2).sum(2)
    assert np.array_equal(lp, d2.astype(np.float32) <= 4.0)          # constant covariance: 2 m radius
    assert np.array_equal(lp, lp.T) and not np.array_equal(ov, ov.T)  # the overlap ratio is relative to the current node
    # covariance-scaled radius
    f2 = po.gate_candidates(pose, ld, ru, np.arange(N), cov_xy=cov, beta=225.0)
    rad = (np.abs(cov[:, 0].astype(np.float32)) + np.abs(cov[:, 1].astype(np.float32))) * np.float32(225)
    assert np.array_equal((f2 & 2) != 0, d2.astype(np.float32) <= rad[:, None])
    assert 0.001 < ((f & 3) == 3).mean() < 0.2


def test_graph_oracle_consensus_properties():
    from tests.graph_cases import matched_sets
    poses, good, off = matched_sets()
    keep, fp, fi = po.consensus(poses, good, off)
    for s in range(off.size - 1):
        lo, hi = off[s], off[s + 1]
        n = hi - lo
        if n == 0:
            assert fi[s] == -1
            continue
        if n < 3:
            assert keep[lo:hi].all()                                  # statis() refuses fewer than 3 samples
        k = keep[lo:hi].astype(bool)
        assert k.any() and lo <= fi[s] < hi and keep[fi[s]]
        w = good[lo:hi][k] / good[lo:hi][k].sum()
        assert np.allclose(fp[s], (w[:, None] * poses[lo:hi][k]).sum(0), atol=1e-4)
        d = np.hypot(*(poses[lo:hi, :2] - fp[s, :2]).T)
        assert fi[s] - lo == np.flatnonzero(k)[np.argmin(d[k])]
    assert 0.02 < 1 - keep.mean() < 0.2                               # some mismatches are dropped, most entries stay


def test_division_free_reading_conversion_is_exact():
    """The CUDA kernel rebuilds a sens point from its float reading with q0 = f * 0.01 and two FMAs instead of
    fromPMScan2LDP's (double)f / 100.0 (src/zhcsm/ZHCanonical_Matcher.cpp:198-208).  Exhaustive over every float a
    laser reading can be: 1 mm .. 2^17 cm."""
    assert po.count_div100_mismatches(0.1, 131072.0) == 0
