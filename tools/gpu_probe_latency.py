"""Latency of ONE match through the batch entry with host buffers (what the CSM-shaped sm_icp() does per call): the first
400 pairs of intel-lab one at a time, with sm_icp's restarts as a second batch (default) and inside the job
(ZHSM_PLICP_FANOUT=0).  Run: python tools/gpu_probe_latency.py"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
from tests.support import load_log, make_laser
g = load_log("intel-lab")
for fan in ("1000000000", "0"):
    os.environ["ZHSM_PLICP_FANOUT"] = fan
    eng = zh.Engine(0); eng.set_laser(make_laser(g))
    scans = eng.median_filter(g["scans"][:401].copy())
    lat, loops = [], []
    for k in range(400):
        pair = np.ascontiguousarray(scans[k:k + 2])
        t = time.perf_counter()
        r = eng.plicp(pair, np.array([0], np.int32), np.array([1], np.int32))
        lat.append(time.perf_counter() - t); loops.append(int(r["n_loops"][0]))
    lat, loops = np.array(lat[20:]) * 1e6, np.array(loops[20:])
    print(f"fanout={'on' if fan != '0' else 'off'}: one pair per call, us: no restart median {np.median(lat[loops == 1]):.0f} "
          f"(n={int((loops == 1).sum())}), restarted median {np.median(lat[loops > 1]):.0f} (n={int((loops > 1).sum())}), all mean {lat.mean():.0f}")
    eng.close()
