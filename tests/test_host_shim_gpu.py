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
