"""Submap cloud builder (zhsm_submap_cloud) against the reference's OWN occupancy-map code: CPMap and the
Bresenham grid_line of src/mapGraph/{PMap,line}.cpp compiled unmodified into oracle/_ref/libpmap_ref.so (parity
pinned).  Cell indices are integer work: the clouds must be identical, point for point and in the same order."""
import numpy as np
import pytest

from oracle import pyoracle as po
from tests.support import load_log, make_laser

pytestmark = pytest.mark.gpu


def _traj(lp):
    c, s = np.cos(lp[0, 2]), np.sin(lp[0, 2])
    d = lp[:, :2] - lp[0, :2]
    return np.stack([c * d[:, 0] + s * d[:, 1], -s * d[:, 0] + c * d[:, 1], lp[:, 2] - lp[0, 2]], 1)


@pytest.mark.parametrize("log", ["fr079", "intel-lab", "mit-cscail"])
def test_submap_cloud_identical_to_reference(engine, log):
    assert po.have_pmap_ref(), "oracle/_ref/libpmap_ref.so missing (built by __graft_entry__.build() where the reference exists)"
    g = load_log(log)
    engine.set_laser(make_laser(g))
    S = g["scans"].shape[0]
    total = 0
    for start, K in ((0, 1), (10, 5), (100, 40), (S - 130, 120), (300, 60)):
        scans = g["scans"][start:start + K]
        traj = _traj(g["laser_pose"][start:start + K])
        wx, wy, wd = po.ref_submap_cloud(scans, traj, g["fi_min"], g["dfi"], int(g["max_range"]))
        x, y, d = engine.submap_cloud(scans, traj)
        assert np.array_equal(d, wd), (log, start, d, wd)
        assert x.size == wx.size and np.array_equal(x, wx) and np.array_equal(y, wy), (log, start, x.size, wx.size)
        total += x.size
    assert total > 1000


def test_submap_cloud_edge_cases(engine):
    g = load_log("fr079")
    engine.set_laser(make_laser(g))
    n = g["scans"].shape[1]
    far = np.full((3, n), float(g["max_range"]) + 1, np.float32)       # nothing in range: empty cloud
    x, y, d = engine.submap_cloud(far, np.zeros((3, 3)))
    assert x.size == 0
    one = far.copy(); one[1, n // 2] = 250.0                            # a single beam
    traj = np.array([[0, 0, 0], [0.3, -0.2, 0.4], [0, 0, 0]], np.float64)
    wx, wy, wd = po.ref_submap_cloud(one, traj, g["fi_min"], g["dfi"], int(g["max_range"]))
    x, y, d = engine.submap_cloud(one, traj)
    assert np.array_equal(x, wx) and np.array_equal(y, wy) and np.array_equal(d, wd)
    with pytest.raises(Exception):
        engine.submap_cloud(g["scans"][:50], _traj(g["laser_pose"][:50]), cap=4)   # capacity too small is reported
