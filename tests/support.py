"""Test-side helpers: golden fixture loading (tests/golden/*.npz, made by tools/make_goldens.py)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LOGS = ["intel-lab", "fr079", "mit-cscail"]


def load_log(name):
    """Returns dict with 'scans' = float32 [S][n] cm exactly as readCarmon builds PMScan::r
    (src/zhpsm/ZHPolar_Match.cpp:250-279): short readings and the padded last beam become max_range+1."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    g = {k: z[k] for k in z.files}
    n, mx = int(g["n"]), int(g["max_range"])
    r = g["ranges_cm"].astype(np.float32)
    S, num = r.shape
    scans = np.full((S, n), mx + 1, np.float32)
    scans[:, :num] = r
    scans[scans < 10.0] = mx + 1
    g["scans"] = scans
    g["laser_name"] = str(g["laser"])
    return g


def make_laser(g):
    import laser_slam_b200 as zh
    lz = zh.Laser()
    lz.n = int(g["n"]); lz.fi_min = float(g["fi_min"]); lz.dfi = float(g["dfi"])
    lz.min_range_cm = int(g["min_range"]); lz.max_range_cm = int(g["max_range"])
    lz.min_valid_points = int(g["min_valid"]); lz.scan_window = int(g["window"])
    return lz


def pose_close(a, b, tol_xy=1e-4, tol_th=1e-5):
    d = np.abs(np.asarray(a) - np.asarray(b))
    return (d[..., 0] <= tol_xy) & (d[..., 1] <= tol_xy) & (d[..., 2] <= tol_th)
