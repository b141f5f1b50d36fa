"""GPU parity: CUDA PL-ICP (through the C ABI) vs the CPU restatement oracle on the three CARMEN logs.

Tolerances are north_star's: pose 1e-4 m / 1e-5 rad, iteration counts equal, correspondence hash equal."""
import numpy as np
import pytest

from tests.support import LOGS, load_log, make_laser, pose_close

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("log", LOGS)
def test_median_filter_bit_exact(engine, log):
    from oracle import pyoracle as po
    g = load_log(log)
    engine.set_laser(make_laser(g))
    scans = g["scans"][:400].copy()
    want = np.stack([po.median_filter(s) for s in scans])
    got = engine.median_filter(scans.copy())
    assert np.array_equal(got, want)


@pytest.mark.parametrize("log", LOGS)
def test_plicp_matches_oracle_goldens(engine, log):
    g = load_log(log)
    engine.set_laser(make_laser(g))
    scans = engine.median_filter(g["scans"].copy())
    S = scans.shape[0]
    res, j1, j2 = engine.plicp(scans, np.arange(S - 1), np.arange(1, S), want_corr=True)
    valid_eq = res["valid"] == g["plicp_valid"]
    both = (res["valid"] == 1) & (g["plicp_valid"] == 1)
    pose_ok = pose_close(res["x"], g["plicp_x"]) | ~both
    it_eq = res["iterations"] == g["plicp_iterations"]
    hash_eq = (res["corr_hash"] == g["plicp_hash"]) | ~both
    nv_eq = (res["nvalid"] == g["plicp_nvalid"]) | ~both
    print(f"\n{log}: pairs={S-1} valid_eq={valid_eq.mean():.5f} pose_ok={pose_ok.mean():.5f} "
          f"iters_eq={it_eq.mean():.5f} hash_eq={hash_eq.mean():.5f} nvalid_eq={nv_eq.mean():.5f} "
          f"max|dx|={np.abs(res['x'][both]-g['plicp_x'][both]).max(0)}")
    assert valid_eq.all()
    assert pose_ok.mean() >= 0.999
    assert it_eq.mean() >= 0.999
    assert hash_eq.mean() >= 0.999
    # correspondence arrays are self-consistent with the hash stream semantics
    assert ((j1 >= -1) & (j1 < int(g["n"]))).all() and ((j1 < 0) == (j2 < 0)).all()


def test_quartic_greatest_root_solver(engine):
    """gpc_solve's lambda: the warp-cooperative solver vs numpy's companion-matrix roots (what GSL does)."""
    import ctypes as C
    rng = np.random.default_rng(11)
    polys, want = [], []
    for k in range(4000):
        kind = k % 4
        if kind == 0:      # four real roots
            r = rng.uniform(-50, 50, 4)
            p = np.poly(r); w = r.max()
        elif kind == 1:    # two real + a complex pair (possibly to the right of the real roots)
            r = rng.uniform(-30, 30, 2); re, im = rng.uniform(-40, 40), rng.uniform(0.05, 20)
            p = np.polymul(np.poly(r), [1, -2 * re, re * re + im * im]); w = r.max()
        elif kind == 2:    # GPC-like: q = p2(x) - l(x)^2, negative leading coefficient, widely spread scales
            l = np.array([4.0, rng.uniform(1, 1e4), rng.uniform(1, 1e6)])
            p2 = np.array([rng.uniform(0, 1e6), rng.uniform(-1e8, 1e8), rng.uniform(0, 1e12)])
            p = -np.polymul(l, l); p[2:] += p2
            rr = np.roots(p); rr = rr[np.abs(rr.imag) < 1e-9 * (1 + np.abs(rr.real))].real
            if rr.size == 0:
                continue
            w = rr.max()
        else:              # no real root
            re1, im1, re2, im2 = rng.uniform(-5, 5), rng.uniform(0.5, 5), rng.uniform(-5, 5), rng.uniform(0.5, 5)
            p = np.polymul([1, -2 * re1, re1 ** 2 + im1 ** 2], [1, -2 * re2, re2 ** 2 + im2 ** 2]); w = None
        p = p * rng.choice([-1, 1]) * rng.uniform(0.1, 100)
        polys.append(p[::-1].copy()); want.append(w)
    coef = np.ascontiguousarray(np.array(polys), np.float64)
    B = coef.shape[0]
    root = np.zeros(B); ok = np.zeros(B, np.int32)
    engine.L.zhsm_debug_quartic_roots.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    st = engine.L.zhsm_debug_quartic_roots(engine.h, coef.ctypes.data, B, root.ctypes.data, ok.ctypes.data)
    assert st == 0
    bad = 0
    for k in range(B):
        if want[k] is None:
            assert ok[k] == 0, (k, coef[k], root[k])
        else:
            assert ok[k] == 1, (k, coef[k])
            if abs(root[k] - want[k]) > 1e-6 * max(1.0, abs(want[k])):
                bad += 1
    # ill-conditioned (nearly double) greatest roots may differ in the low digits; anything else is a bug
    assert bad <= B // 500, bad


@pytest.mark.parametrize("log", ["intel-lab", "fr079"])
def test_plicp_covariance_matches_oracle(engine, log):
    """compute_covariance_exact (setCov(true), ZHCanonical_Matcher.cpp:79-81,292-296): sigma^2 * cov0, relative 1e-6
    (the kernel evaluates the reference's trigonometry algebraically)."""
    from oracle import pyoracle as po
    g = load_log(log)
    engine.set_laser(make_laser(g))
    S = 200
    scans = engine.median_filter(g["scans"][:S].copy())
    p = engine.default_plicp_params()
    p.do_compute_covariance = 1
    res = engine.plicp(scans, np.arange(S - 1), np.arange(1, S), params=p)
    lz = make_laser(g)
    o = po.zhcsm_fmatch_batch(scans, np.arange(S - 1), np.arange(1, S), lz.fi_min, lz.dfi, lz.min_range_cm,
                              lz.max_range_cm, params=po.default_params(lz.min_range_cm, lz.max_range_cm,
                                                                        do_compute_covariance=1), nthreads=8)
    same = (res["valid"] == 1) & (o["valid"] == 1) & (res["corr_hash"] == o["corr_hash"])
    assert same.mean() > 0.99
    oc = o["cov"].reshape(-1, 3, 3)
    want = np.stack([oc[:, 0, 0], oc[:, 0, 1], oc[:, 0, 2], oc[:, 1, 1], oc[:, 1, 2], oc[:, 2, 2]], 1)
    got = res["cov"]
    scale = np.abs(want).max(1, keepdims=True)
    rel = np.abs(got - want) / scale
    assert np.isfinite(got[same]).all()
    assert rel[same].max() < 1e-6, rel[same].max()
    # symmetric positive semi-definite, and without the flag the record carries zeros
    assert (got[same][:, [0, 3, 5]] >= 0).all()
    res0 = engine.plicp(scans, np.arange(S - 1), np.arange(1, S))
    assert np.all(res0["cov"] == 0) and np.array_equal(res0["x"], res["x"])
