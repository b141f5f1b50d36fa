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


def _oracle(g, scans, ri, si, first_guess=None):
    from oracle import pyoracle as po
    return po.zhcsm_fmatch_batch(scans, ri, si, g["fi_min"], g["dfi"], int(g["min_range"]), int(g["max_range"]),
                                 first_guess=first_guess, nthreads=8)


def _assert_same_as_oracle(res, o, min_frac=0.999, max_bad=None):
    """north_star: poses within 1e-4 m / 1e-5 rad, iteration counts equal, correspondences >= 99.9 % identical.  The
    jobs that differ (documented floating-point near-ties) are listed by index; max_bad bounds their NUMBER
    where the batch is too small for a fraction to mean anything."""
    assert np.array_equal(res["valid"], o["valid"])
    both = (res["valid"] == 1) & (o["valid"] == 1)
    bad = ~(pose_close(res["x"], o["x"]) | ~both) | (res["iterations"] != o["iterations"]) | \
        ((res["corr_hash"] != o["corr_hash"]) & both) | ((res["nvalid"] != o["nvalid"]) & both)
    idx = np.flatnonzero(bad)
    print(f"jobs differing from the oracle: {idx.size} of {bad.size}: {idx.tolist()[:20]}")
    if max_bad is not None:
        assert idx.size <= max_bad, idx.tolist()[:20]
    assert 1.0 - bad.mean() >= min_frac, idx.tolist()[:20]


def test_plicp_degenerate_distances(engine):
    """A scan matched against itself: every point-to-segment distance is 0 or a few ulps, so the trim keys do not
    share the top byte the radix select guesses and it takes its general path from the first byte on; nearly equal
    scans (one reading changed) mix exact zeros with ordinary distances.  kill_outliers_trim sorts in the reference
    (CSM icp_outliers.c behind sm_icp, src/zhcsm/ZHCanonical_Matcher.cpp:274): same result required."""
    g = load_log("fr079")
    engine.set_laser(make_laser(g))
    base = engine.median_filter(g["scans"][:120].copy())
    S = base.shape[0]
    bumped = base.copy()
    bumped[:, 100] += 3.0   # one reading 3 cm off
    half = base.copy()
    half[:, ::2] += 1.0     # every other reading 1 cm off
    scans = np.ascontiguousarray(np.concatenate([base, bumped, half]))
    ri = np.concatenate([np.arange(S), np.arange(S), np.arange(S)]).astype(np.int32)
    si = np.concatenate([np.arange(S), np.arange(S) + S, np.arange(S) + 2 * S]).astype(np.int32)
    guess = np.zeros((ri.size, 3)); guess[1::3] = [0.02, -0.01, 0.003]
    res, _, _ = engine.plicp(scans, ri, si, first_guess=guess, want_corr=True)
    o = _oracle(g, scans, ri, si, first_guess=guess)
    assert np.array_equal(res["valid"], o["valid"]) and (res["valid"] == 1).all()
    assert pose_close(res["x"], o["x"], 1e-9, 1e-9).all()      # measured: 3e-14
    assert np.array_equal(res["iterations"], o["iterations"])
    # Which correspondences the trim keeps is only defined where the distances are: a job that converges from a
    # non-zero guess onto (nearly) identical scans ends with distances that are pure rounding noise of a pose the
    # two solvers compute to 1e-14 of each other, and the kept set follows that noise (measured: 70 % equal, the
    # same with the plain radix select).  Everywhere else the sets must be the reference's.
    defined = (guess == 0).all(1) | (np.arange(ri.size) >= 2 * S)
    assert np.array_equal(res["corr_hash"][defined], o["corr_hash"][defined])
    assert np.array_equal(res["nvalid"][defined], o["nvalid"][defined])
    same = slice(0, S, 3)   # identical scans from a zero first guess: the pose stays at zero
    assert np.abs(res["x"][same]).max() < 1e-9


@pytest.mark.parametrize("threads", [64, 128, 160, 192, 256, 512])
def test_plicp_independent_of_cta_shape(threads):
    """The CTA size (plicp_shape; ZHSM_PLICP_T is its tuning override) changes which thread owns which ray, the order
    of the floating-point block sums, the chunks of the first iterations and of the parallel hash -- never the
    result beyond summation rounding: every shape must agree with the oracle like the default one does."""
    import os
    import laser_slam_b200 as zh
    g = load_log("mit-cscail")
    old = os.environ.get("ZHSM_PLICP_T")
    os.environ["ZHSM_PLICP_T"] = str(threads)
    try:
        eng = zh.Engine(0)
        eng.set_laser(make_laser(g))
        scans = eng.median_filter(g["scans"][:301].copy())
        res, _, _ = eng.plicp(scans, np.arange(300), np.arange(1, 301), want_corr=True)
        eng.close()
    finally:
        if old is None:
            del os.environ["ZHSM_PLICP_T"]
        else:
            os.environ["ZHSM_PLICP_T"] = old
    both = (res["valid"] == 1) & (g["plicp_valid"][:300] == 1)
    assert np.array_equal(res["valid"], g["plicp_valid"][:300])
    # the CTA shape must not change a single record of these 300 pairs (the one known near-tie pair of this log is 1004)
    bad = ~(pose_close(res["x"], g["plicp_x"][:300]) | ~both) | (res["iterations"] != g["plicp_iterations"][:300]) | \
        ((res["corr_hash"] != g["plicp_hash"][:300]) & both)
    assert not bad.any(), (threads, np.flatnonzero(bad).tolist())


def test_plicp_long_scans_poor_guesses_match_oracle():
    """SURVEY 8d C5 in small: 2048-beam 360-degree scans, 16 first guesses per pair up to 0.3 m / 10 degrees off a noisy
    odometry.  With such guesses one walk in twelve runs for hundreds of probes and some cross the whole scan; the
    256-thread instance hands those to the warp (walk_warp_one / walk_warp_both, plicp2_search.cuh), which has to take
    the sequential walk's decisions: pose, iterations and correspondence hash equal to the CPU restatement's."""
    import torch
    import laser_slam_b200 as zh
    from oracle import pyoracle as po
    P, G, n = 24, 16, 2048
    PM_PI = 3.141592654
    lz = zh.Laser()
    lz.n = n; lz.fi_min = np.float32(-PM_PI); lz.dfi = np.float32((360 * PM_PI / 180) / (n - 1.0))
    lz.min_range_cm = 10; lz.max_range_cm = 3000; lz.min_valid_points = n // 5; lz.scan_window = n // 10
    eng = zh.Engine(0)
    eng.set_laser(lz)
    dev = torch.device("cuda:0")
    d_scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev)
    d_truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
    eng.synth_pairs(d_scans, d_truth, 0, P, dmax_xy_m=0.5, dmax_th_rad=np.deg2rad(20.0), seed=20261105)
    eng.median_filter(d_scans); eng.sync()
    scans = d_scans.cpu().numpy(); tr = d_truth.cpu().numpy()
    rng = np.random.default_rng(7)
    odo = tr + rng.normal(0, 1, (P, 3)) * [0.1, 0.1, np.deg2rad(3.0)]
    grid = np.array([[dx, dy, np.deg2rad(dt)] for dx in (-0.3, 0.3) for dy in (-0.3, 0.3) for dt in (-10, -3.3, 3.3, 10)])
    guess = (odo[:, None, :] + grid[None]).reshape(-1, 3)
    ri = np.repeat(np.arange(P, dtype=np.int32) * 2, G); si = ri + 1
    res, _, _ = eng.plicp(scans, ri, si, first_guess=guess, want_corr=True)
    eng.close()
    o = po.zhcsm_fmatch_batch(scans, ri, si, lz.fi_min, lz.dfi, 10, 3000, first_guess=guess, nthreads=8)
    _assert_same_as_oracle(res, o, max_bad=0)   # 384 jobs: measured identical on every one (the newest walk code)
    assert (res["valid"] == 1).mean() > 0.9


def test_overlapping_batches_on_two_streams_of_one_handle():
    """The staging slots and the jump-table scratch belong to the handle: two batches issued back to back on two
    different streams (results left on the device, so neither call blocks) must give the records of the serial run --
    every entry orders its stream behind the handle's previous call (StreamScope, zhsm_internal.h)."""
    import torch
    import laser_slam_b200 as zh
    g = load_log("intel-lab")
    eng = zh.Engine(0)
    eng.set_laser(make_laser(g))
    dev = torch.device("cuda:0")
    host = eng.median_filter(g["scans"][:1201].copy())
    a_h, b_h = host[:601], host[600:1201]
    idx = np.arange(600, dtype=np.int32)
    want_a = eng.plicp(a_h, idx, idx + 1)
    want_b = eng.plicp(b_h, idx, idx + 1)
    s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    for rep in range(3):
        ra = torch.zeros((600, 128), dtype=torch.uint8, device=dev)
        rb = torch.zeros((600, 128), dtype=torch.uint8, device=dev)
        # host scans + host job lists -> staged by the handle; device results -> the calls return while kernels run
        eng.plicp(a_h, idx, idx + 1, results=ra, stream=s1.cuda_stream)
        eng.plicp(b_h, idx, idx + 1, results=rb, stream=s2.cuda_stream)
        torch.cuda.synchronize()
        got_a = ra.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
        got_b = rb.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
        for got, want in ((got_a, want_a), (got_b, want_b)):
            assert np.array_equal(got["corr_hash"], want["corr_hash"]) and np.array_equal(got["iterations"], want["iterations"])
            assert np.array_equal(got["x"], want["x"])
    eng.close()


def test_bad_device_job_index_fails_its_own_job_only():
    """Device-resident job lists are not checked by the host: an out-of-range scan index must fail that job with
    ZHSM_JOB_INVALID_SCAN and leave the rest of the batch intact."""
    import torch
    import laser_slam_b200 as zh
    g = load_log("fr079")
    eng = zh.Engine(0)
    eng.set_laser(make_laser(g))
    dev = torch.device("cuda:0")
    scans = torch.from_numpy(eng.median_filter(g["scans"][:65].copy())).to(dev)
    ri = torch.arange(0, 64, dtype=torch.int32, device=dev)
    si = ri + 1
    good = torch.zeros((64, 128), dtype=torch.uint8, device=dev)
    eng.plicp(scans, ri, si, results=good)
    ri2, si2 = ri.clone(), si.clone()
    ri2[5] = 65; si2[9] = -3; ri2[40] = 1 << 30
    out = torch.zeros((64, 128), dtype=torch.uint8, device=dev)
    eng.plicp(scans, ri2, si2, results=out)
    eng.sync()
    a = good.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
    b = out.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
    badj = np.array([5, 9, 40])
    assert (b["valid"][badj] == 0).all() and (b["status"][badj] == 4).all()
    rest = np.setdiff1d(np.arange(64), badj)
    assert np.array_equal(a[rest].tobytes(), b[rest].tobytes())
    eng.close()


def test_engine_rejects_malformed_device_tensors():
    """Only raw pointers cross the C ABI, so the binding must refuse what the kernels would misread: a float64 or
    sliced scan tensor, job arrays split between host and device, a wrong beam count."""
    import torch
    import laser_slam_b200 as zh
    g = load_log("fr079")
    eng = zh.Engine(0)
    eng.set_laser(make_laser(g))
    dev = torch.device("cuda:0")
    scans = torch.from_numpy(g["scans"][:9].copy()).to(dev)
    ri = torch.arange(0, 8, dtype=torch.int32, device=dev)
    res = torch.zeros((8, 128), dtype=torch.uint8, device=dev)
    bad_calls = [
        lambda: eng.plicp(scans.double(), ri, ri + 1, results=res),               # dtype
        lambda: eng.plicp(scans[:, :360], ri, ri + 1, results=res),               # sliced: not contiguous, wrong beams
        lambda: eng.plicp(scans, ri, (ri + 1).cpu().numpy(), results=res),        # job arrays on two sides
        lambda: eng.plicp(scans, ri, ri + 1, first_guess=np.zeros((8, 3)), results=res),  # first_guess on the host
        lambda: eng.plicp(scans, ri.long(), ri + 1, results=res),                 # int64 indices
        lambda: eng.plicp(scans, ri, ri + 1, results=res[:4]),                    # too few result records
        lambda: eng.median_filter(scans.t().contiguous()),                        # [n][S] instead of [S][n]
    ]
    for k, call in enumerate(bad_calls):
        with pytest.raises(zh.ZhsmError):
            call()
    eng.plicp(scans, ri, ri + 1, results=res)   # and the well-formed call still runs
    eng.sync()
    assert (res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)["valid"] == 1).all()
    eng.close()


def test_walk_pool_same_bytes_as_ray_by_ray():
    """The long-scan instances run a round's walks as one pool behind a shared counter (plicp2_flat.cuh); ZHSM_PLICP_POOL=0
    (read in zhsm_create, tests only) launches the 256-thread shape from an instance built without the pool
    (plicp2_t256r.cu), which calls tricks_one ray by ray as the short-scan instances do.  Same
    decisions on the same operands: the records (diagnostics aside) and every correspondence must be the same BYTES, on
    2048-beam multi-start jobs with poor first guesses (pool, tail budget and the list of unfinished walks all busy) and
    with the covariance on."""
    import os
    import torch
    import laser_slam_b200 as zh
    P, G, n = 32, 16, 2048
    PM_PI = 3.141592654
    lz = zh.Laser()
    lz.n = n; lz.fi_min = np.float32(-PM_PI); lz.dfi = np.float32((360 * PM_PI / 180) / (n - 1.0))
    lz.min_range_cm = 10; lz.max_range_cm = 3000; lz.min_valid_points = n // 5; lz.scan_window = n // 10
    out = {}
    old = os.environ.get("ZHSM_PLICP_POOL")
    try:
        for pool in ("1", "0"):
            os.environ["ZHSM_PLICP_POOL"] = pool
            eng = zh.Engine(0)
            try:
                eng.set_laser(lz)
                dev = torch.device("cuda:0")
                d_scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev)
                d_truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
                eng.synth_pairs(d_scans, d_truth, 0, P, dmax_xy_m=0.5, dmax_th_rad=np.deg2rad(20.0), seed=20261106)
                eng.median_filter(d_scans); eng.sync()
                scans = d_scans.cpu().numpy(); tr = d_truth.cpu().numpy()
                rng = np.random.default_rng(11)
                odo = tr + rng.normal(0, 1, (P, 3)) * [0.1, 0.1, np.deg2rad(3.0)]
                grid = np.array([[dx, dy, np.deg2rad(dt)] for dx in (-0.3, 0.3) for dy in (-0.3, 0.3) for dt in (-10, -3.3, 3.3, 10)])
                guess = (odo[:, None, :] + grid[None]).reshape(-1, 3)
                ri = np.repeat(np.arange(P, dtype=np.int32) * 2, G); si = ri + 1
                p = eng.default_plicp_params()
                p.do_compute_covariance = 1
                out[pool] = eng.plicp(scans, ri, si, first_guess=guess, params=p, want_corr=True)
            finally:
                eng.close()
    finally:
        if old is None:
            os.environ.pop("ZHSM_PLICP_POOL", None)
        else:
            os.environ["ZHSM_PLICP_POOL"] = old
    (a, a1, a2), (b, b1, b2) = out["1"], out["0"]
    assert (a["valid"] == 1).mean() > 0.9
    assert (a["reserved"][:, 3] != b["reserved"][:, 3]).any()     # the two paths really differ (their step counts do)
    assert np.array_equal(a1, b1) and np.array_equal(a2, b2)
    a, b = a.copy(), b.copy()
    assert np.allclose(a["cov"], b["cov"], rtol=1e-6, atol=1e-15)  # (shared-memory atomics)
    a["cov"] = b["cov"]; a["reserved"] = b["reserved"]
    assert np.array_equal(a.tobytes(), b.tobytes())


@pytest.mark.parametrize("log", ["intel-lab", "fr079"])
def test_restarts_as_second_batch_same_bytes(log):
    """sm_icp's restart loop (six perturbed icp_loops, one after the other) runs as a second batch of one-loop jobs for
    small batches (zhsm_api.cu: plicp_fanout; ZHSM_PLICP_FANOUT is the largest such batch, 0 = always in the job).  Each
    loop depends on its start pose only, so the records must be the same BYTES either way -- pose, error, iterations,
    n_loops, nvalid, hash, every correspondence (diagnostics aside; covariance to 1e-6: shared-memory atomics) -- on
    real logs, where 10 % (intel-lab) to 40 % (fr079) of the pairs restart."""
    import os
    import laser_slam_b200 as zh
    g = load_log(log)
    out = {}
    old = os.environ.get("ZHSM_PLICP_FANOUT")
    try:
        for fan in ("100000", "0"):
            os.environ["ZHSM_PLICP_FANOUT"] = fan
            eng = zh.Engine(0)
            try:
                eng.set_laser(make_laser(g))
                scans = eng.median_filter(g["scans"].copy())
                S = scans.shape[0]
                p = eng.default_plicp_params()
                p.do_compute_covariance = 1
                out[fan] = eng.plicp(scans, np.arange(S - 1), np.arange(1, S), params=p, want_corr=True)
            finally:
                eng.close()
    finally:
        if old is None:
            os.environ.pop("ZHSM_PLICP_FANOUT", None)
        else:
            os.environ["ZHSM_PLICP_FANOUT"] = old
    (a, a1, a2), (b, b1, b2) = out["100000"], out["0"]
    restarted = b["n_loops"] > 1
    assert 0.05 < restarted.mean() < 0.6, restarted.mean()
    assert np.array_equal(a1, b1) and np.array_equal(a2, b2)
    a, b = a.copy(), b.copy()
    assert np.allclose(a["cov"], b["cov"], rtol=1e-6, atol=1e-15)
    a["cov"] = b["cov"]; a["reserved"] = b["reserved"]
    bad = np.flatnonzero([x.tobytes() != y.tobytes() for x, y in zip(a, b)])
    assert bad.size == 0, (bad[:10], a[bad[:3]], b[bad[:3]])


def _same_records(a, b, with_cov):
    """Two CTA shapes of the kernel on the same jobs: every integer of the record equal (valid, status, iterations, nvalid,
    correspondence hash, loops), pose and error equal to rounding (the 14 GPC sums are added in a tree whose shape follows
    the thread count: <= 1e-12 m), covariance to 1e-6 (shared-memory atomics)."""
    for f in ("valid", "status", "iterations", "nvalid", "corr_hash", "n_loops"):
        assert np.array_equal(a[f], b[f]), (f, np.flatnonzero(a[f] != b[f])[:10])
    assert np.abs(a["x"] - b["x"]).max() <= 1e-12 and np.allclose(a["error"], b["error"], rtol=1e-10, atol=0)
    if with_cov:
        assert np.allclose(a["cov"], b["cov"], rtol=1e-6, atol=1e-15)


def test_multistart_instance_same_records_and_checks_its_promise():
    """ZHSM_MULTISTART (include/zhsm.h): four jobs that share both scans per CTA over one staged copy of the scans
    (plicp2_ms128.cu).  On 2048-beam multi-start jobs with poor first guesses (the long-walk code runs in it) the records
    must be those of the single-job instance (_same_records), with and without the covariance; shorter scans ignore the
    flag; a run of four that does NOT share its scans fails those jobs (ZHSM_JOB_INVALID_SCAN) and nothing else."""
    import torch
    import laser_slam_b200 as zh
    P, G, n = 24, 16, 2048
    PM_PI = 3.141592654
    lz = zh.Laser()
    lz.n = n; lz.fi_min = np.float32(-PM_PI); lz.dfi = np.float32((360 * PM_PI / 180) / (n - 1.0))
    lz.min_range_cm = 10; lz.max_range_cm = 3000; lz.min_valid_points = n // 5; lz.scan_window = n // 10
    eng = zh.Engine(0)
    try:
        eng.set_laser(lz)
        dev = torch.device("cuda:0")
        d_scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev)
        d_truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
        eng.synth_pairs(d_scans, d_truth, 0, P, dmax_xy_m=0.5, dmax_th_rad=np.deg2rad(20.0), seed=20261105)
        eng.median_filter(d_scans); eng.sync()
        scans = d_scans.cpu().numpy(); tr = d_truth.cpu().numpy()
        rng = np.random.default_rng(7)
        odo = tr + rng.normal(0, 1, (P, 3)) * [0.1, 0.1, np.deg2rad(3.0)]
        grid = np.array([[dx, dy, np.deg2rad(dt)] for dx in (-0.3, 0.3) for dy in (-0.3, 0.3) for dt in (-10, -3.3, 3.3, 10)])
        guess = (odo[:, None, :] + grid[None]).reshape(-1, 3)
        ri = np.repeat(np.arange(P, dtype=np.int32) * 2, G); si = ri + 1
        for cov in (0, 1):
            p = eng.default_plicp_params()
            p.do_compute_covariance = cov
            one, a1, a2 = eng.plicp(scans, ri, si, first_guess=guess, params=p, want_corr=True)
            ms, b1, b2 = eng.plicp(scans, ri, si, first_guess=guess, params=p, want_corr=True, multistart=True)
            _same_records(ms, one, cov)
            assert np.array_equal(a1, b1) and np.array_equal(a2, b2)   # every correspondence of every job
        assert (one["valid"] == 1).mean() > 0.9 and (one["reserved"][:, 3] > 0).all()
        # a broken promise: the second run of four mixes two pairs
        ri2, si2 = ri.copy(), si.copy()
        ri2[5], si2[5] = ri[G], si[G]
        bad = eng.plicp(scans, ri2, si2, first_guess=guess, multistart=True)
        assert bad["status"][5] == 4 and bad["valid"][5] == 0
        keep = np.ones(P * G, bool); keep[5] = False
        _same_records(bad[keep], one[keep], 0)   # (this call ran without the covariance)
        # B not a multiple of 4: the flag is ignored, the single-job instance runs: the same bytes (covariance and diagnostics aside)
        odd = eng.plicp(scans, ri[:-2], si[:-2], first_guess=guess[:-2], params=p, multistart=True)
        odd = odd.copy(); odd["cov"] = one[:-2]["cov"]
        odd["reserved"] = one[:-2]["reserved"]   # diagnostics (walks, steps): they follow the warps' timing on the list of unfinished walks
        assert np.array_equal(odd.tobytes(), one[:-2].tobytes())
    finally:
        eng.close()
    # short scans ignore the flag
    g = load_log("fr079")
    eng = zh.Engine(0)
    try:
        eng.set_laser(make_laser(g))
        s = eng.median_filter(g["scans"][:41].copy())
        r4 = np.repeat(np.arange(10, dtype=np.int32) * 4, 4)
        a = eng.plicp(s, r4, r4 + 1)
        b = eng.plicp(s, r4, r4 + 1, multistart=True)
        assert np.array_equal(a.tobytes(), b.tobytes())
    finally:
        eng.close()
