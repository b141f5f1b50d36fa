"""GPU test of zhsm_plicp_batch_noalloc (include/zhsm.h; no counterpart in the reference):
the entry point that allocates nothing -- its kernels' scratch is a static device array -- must give
the records of the existing zhsm_plicp_batch on the same inputs, on two calls in a row (the array keeps the first
call's contents when the second one runs), and must refuse sizes beyond its stated maximum.

Kept in a file of its own that pytest collects last.  It builds its own engine, so it passes alone in a fresh process
and after the other tests alike."""
import numpy as np
import pytest

from tests.support import load_log, make_laser

pytestmark = pytest.mark.gpu

FIELDS = ("valid", "status", "iterations", "nvalid", "corr_hash", "n_loops")


def _same_records(got, want, what):
    # the tolerance of the PL-ICP parity tests (pose 1e-4 m / 1e-5 rad, iteration counts and correspondence hashes
    # equal); the two entries run the same kernels, so in practice the poses are equal to the last bit
    for f in FIELDS:
        assert np.array_equal(got[f], want[f]), (what, f, np.flatnonzero(got[f] != want[f])[:10])
    d = np.abs(got["x"] - want["x"])
    assert (d[:, :2] <= 1e-4).all() and (d[:, 2] <= 1e-5).all(), (what, d.max(0))
    assert np.allclose(got["error"], want["error"], rtol=1e-9, atol=0), what


def test_noalloc_entry_matches_existing_entry_on_two_calls_in_a_row():
    import torch
    import laser_slam_b200 as zh
    g = load_log("fr079")
    eng = zh.Engine(0)
    try:
        eng.set_laser(make_laser(g))
        n = int(g["n"])
        dev = torch.device("cuda:0")
        host = eng.median_filter(g["scans"][:700].copy())
        scans = torch.from_numpy(host).to(dev)
        S = scans.shape[0]
        # call 1: consecutive-scan odometry; call 2: other scans, another batch size, and jobs that SHARE ref scans
        # (each ref scan matched against its next three scans: the shared jump tables of plicp2_jump.cu) with first
        # guesses -- so whatever call 1 left in the scratch array is wrong for call 2
        ri1 = np.arange(0, 300, dtype=np.int32)
        si1 = ri1 + 1
        ri2 = np.repeat(np.arange(350, 650, 2, dtype=np.int32), 3)
        si2 = ri2 + np.tile(np.arange(1, 4, dtype=np.int32), ri2.size // 3)
        rng = np.random.default_rng(20261104)
        fg2 = rng.uniform(-1, 1, (ri2.size, 3)) * [0.05, 0.05, 0.02]
        calls = [(ri1, si1, None), (ri2, si2, fg2)]
        want = []
        for ri, si, fg in calls:  # the existing entry point, host job lists, host results
            want.append(eng.plicp(scans, torch.from_numpy(ri).to(dev), torch.from_numpy(si).to(dev),
                                  first_guess=None if fg is None else torch.from_numpy(fg).to(dev),
                                  results=torch.zeros((ri.size, 128), dtype=torch.uint8, device=dev), want_corr=True))
        eng.sync()
        want = [(r.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1), j1.cpu().numpy(), j2.cpu().numpy()) for r, j1, j2 in want]
        assert (want[0][0]["valid"] == 1).mean() > 0.9 and (want[1][0]["valid"] == 1).mean() > 0.9

        # every buffer of the new entry point is the caller's, made BEFORE the calls
        bufs = []
        for ri, si, fg in calls:
            bufs.append(dict(ri=torch.from_numpy(ri).to(dev), si=torch.from_numpy(si).to(dev),
                             fg=None if fg is None else torch.from_numpy(fg).to(dev),
                             res=torch.zeros((ri.size, 128), dtype=torch.uint8, device=dev),
                             j1=torch.full((ri.size, n), -7, dtype=torch.int16, device=dev),
                             j2=torch.full((ri.size, n), -7, dtype=torch.int16, device=dev)))
        torch.cuda.synchronize()
        for rep in range(2):  # the pair of calls twice over: call 1 then also runs on call 2's leftovers
            free_before = None
            for k, b in enumerate(bufs):
                b["res"].zero_()
                torch.cuda.synchronize()
                if rep == 1:
                    free_before = torch.cuda.mem_get_info(dev)[0]
                eng.plicp_noalloc(scans, b["ri"], b["si"], b["res"], first_guess=b["fg"], corr_j1=b["j1"], corr_j2=b["j2"])
                eng.sync()
                if rep == 1:  # (the first use on a device may load the module; from then on nothing is allocated)
                    assert torch.cuda.mem_get_info(dev)[0] == free_before, "device memory was allocated by the call"
                got = b["res"].cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
                _same_records(got, want[k][0], f"rep {rep} call {k + 1}")
                assert np.array_equal(b["j1"].cpu().numpy(), want[k][1]) and np.array_equal(b["j2"].cpu().numpy(), want[k][2])

        # sizes beyond the stated maximum are refused with a clear error, nothing is launched
        B_big = (1 << 20) // n + 1
        big = torch.zeros(B_big, dtype=torch.int32, device=dev)
        with pytest.raises(zh.ZhsmError, match="ZHSM_NOALLOC_MAX_RAYS"):
            eng.plicp_noalloc(scans, big, big, torch.zeros((B_big, 128), dtype=torch.uint8, device=dev))
        pool_big = torch.zeros((8193, n), dtype=torch.float32, device=dev)
        with pytest.raises(zh.ZhsmError, match="ZHSM_NOALLOC_MAX_SCANS"):
            eng.plicp_noalloc(pool_big, bufs[0]["ri"], bufs[0]["si"], bufs[0]["res"])
        # ... and the largest batch that fits still runs (970 jobs at 1080 beams; here (1 << 20) // 361 = 2904)
        B_max = (1 << 20) // n
        ri = torch.arange(0, B_max, dtype=torch.int32, device=dev) % (S - 1)
        res = torch.zeros((B_max, 128), dtype=torch.uint8, device=dev)
        eng.plicp_noalloc(scans, ri, ri + 1, res)
        eng.sync()
        got = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
        ref = eng.plicp(host, ri.cpu().numpy(), (ri + 1).cpu().numpy())
        _same_records(got, ref, "largest batch")
    finally:
        eng.close()
