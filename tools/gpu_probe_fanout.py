"""Feasibility probe: sm_icp's restart loop as a second batch.  Phase A = every job with restart off (one icp_loop);
the jobs whose mean error exceeds the threshold get six more one-loop jobs from x_first (+) the six perturbations
(phase B); sm_icp's bookkeeping on the host (first failing restart ends the loop, strict < keeps the earliest best).
Compared with the kernel's own in-job restart loop: pose, error, iterations, n_loops, nvalid of every job."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
from tests.support import load_log, make_laser
for log in ("intel-lab", "fr079"):
    g = load_log(log)
    eng = zh.Engine(0); eng.set_laser(make_laser(g))
    scans = eng.median_filter(g["scans"].copy()); S = scans.shape[0]; B = S - 1
    ri = np.arange(0, B, dtype=np.int32); si = ri + 1
    p = eng.default_plicp_params()
    full = eng.plicp(scans, ri, si, params=p)
    q = eng.default_plicp_params(); q.restart = 0
    a = eng.plicp(scans, ri, si, params=q)
    need = (a["valid"] == 1) & (a["error"] / np.maximum(a["nvalid"], 1) > p.restart_threshold_mean_error)
    idx = np.flatnonzero(need)
    dt, dth = p.restart_dt, p.restart_dtheta
    pert = np.array([[dt, 0, 0], [-dt, 0, 0], [0, dt, 0], [0, -dt, 0], [0, 0, dth], [0, 0, -dth]])
    guess = (a["x"][idx][:, None, :] + pert[None]).reshape(-1, 3)
    b = eng.plicp(scans, np.repeat(ri[idx], 6), np.repeat(si[idx], 6), first_guess=guess, params=q).reshape(-1, 6)
    out = a.copy()
    for r, j in enumerate(idx):
        best_err, best_x, its, loops = a["error"][j], a["x"][j].copy(), int(a["iterations"][j]), 1
        for k in range(6):
            loops += 1
            if b["valid"][r, k] != 1:
                break
            its += int(b["iterations"][r, k])
            if b["error"][r, k] < best_err:
                best_err, best_x = b["error"][r, k], b["x"][r, k].copy()
        out["x"][j], out["error"][j], out["iterations"][j], out["n_loops"][j] = best_x, best_err, its, loops
    same = (np.array_equal(out["x"].tobytes(), full["x"].tobytes()), np.array_equal(out["error"], full["error"]),
            np.array_equal(out["iterations"], full["iterations"]), np.array_equal(out["n_loops"], full["n_loops"]),
            np.array_equal(out["valid"], full["valid"]), np.array_equal(out["nvalid"], full["nvalid"]))
    bad = np.flatnonzero((out["iterations"] != full["iterations"]) | (out["n_loops"] != full["n_loops"]) | (out["error"] != full["error"]))
    print(f"{log}: {B} pairs, {idx.size} restart; x / error / iterations / n_loops / valid / nvalid identical: {same}; differing jobs {bad[:10].tolist()}")
    eng.close()
