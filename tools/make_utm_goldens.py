#!/usr/bin/env python
"""Freeze tests/golden/utm_gate.npz (run HERE, where /root/reference exists): pm_psm of every consecutive pair of the
three CARMEN logs with pm_orientation_search's "HOKUYO_UTM_30LX" branch on (src/zhpsm/ZHPolar_Match.cpp:1182-1191:
a beam counts only when |new_r - ref_r| < PM_MAX_ERROR).  The reference keys that branch on the sensor's NAME and has
no parameter table for the UTM, so the compiled reference (oracle/_ref/libzhpsm_ref.so) is run with the log's own
geometry and the name replaced -- the only way to execute that branch of the reference."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402
from tests.support import LOGS, load_log  # noqa: E402


def main():
    po.build()
    out = {}
    for log in LOGS:
        g = load_log(log)
        m = po.RefPolarMatch(g["laser_name"], fi_min=np.float32(g["fi_min"]), max_range_cm=int(g["max_range"]))
        pre = [m.preprocess(s) for s in g["scans"]]
        m.set_laser_name("HOKUYO_UTM_30LX")
        S = len(pre)
        pose = np.zeros((S - 1, 3), np.float32)
        st = np.zeros(S - 1, np.int8)
        for k in range(1, S):
            s, p, _ = m.match(0, pre[k - 1], pre[k])
            pose[k - 1], st[k - 1] = p, s
        differs = int((np.any(pose != g["psm_pose"], 1) | (st != g["psm_status"])).sum())
        print(log, "pairs", S - 1, "failures", int(st.sum()), "pairs whose result differs from the ungated search:", differs)
        out[log + "_pose"], out[log + "_status"] = pose, st
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "utm_gate.npz"), **out)


if __name__ == "__main__":
    main()
