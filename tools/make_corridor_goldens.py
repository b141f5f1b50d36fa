#!/usr/bin/env python
"""Freeze tests/golden/corridor.npz: pm_is_corridor (src/zhpsm/ZHPolar_Match.cpp:2058-2186) of every preprocessed scan
of the three CARMEN logs, produced by EXECUTING the reference (oracle/_ref/libzhpsm_ref.so).  Run HERE."""
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
        m = po.RefPolarMatch(g["laser_name"], fi_min=g["fi_min"], max_range_cm=int(g["max_range"]))
        flags = np.zeros(g["scans"].shape[0], np.int8)
        for k, s in enumerate(g["scans"]):
            flags[k] = m.is_corridor(*m.preprocess(s))
        out[log] = flags
        print(log, {int(v): int((flags == v).sum()) for v in np.unique(flags)})
    # synthetic LMS511 scans that exercise the corridor branch, the pi/5-rotated second pass and both throws:
    # a sensor between two parallel walls (width, heading and lateral offset vary), rooms, and nearly empty scans
    rng = np.random.default_rng(20261104)
    m = po.RefPolarMatch("LMS511")
    fi = (np.arange(m.n, dtype=np.float32) * m.dfi + m.fi_min).astype(np.float64)
    syn = np.zeros((96, m.n), np.uint16)
    for k in range(96):
        kind = k % 4
        if kind < 2:      # corridor: walls at y' = +-w/2 in a frame rotated by a, ends open (max range)
            w, a, off = rng.uniform(120, 400), rng.uniform(-1.5, 1.5), rng.uniform(-0.3, 0.3)
            sn = np.sin(fi - a)
            d = np.where(sn > 1e-3, (w / 2 - off * w) / np.maximum(sn, 1e-3), np.where(sn < -1e-3, (-w / 2 - off * w) / np.minimum(sn, -1e-3), 1e9))
            r = np.minimum(d, 5001 if kind == 0 else rng.uniform(800, 3000))
        elif kind == 2:   # rectangular room
            hx, hy = rng.uniform(200, 900, 2)
            c, s_ = np.cos(fi), np.sin(fi)
            r = np.minimum(hx / np.maximum(np.abs(c), 1e-6), hy / np.maximum(np.abs(s_), 1e-6))
        else:             # almost everything out of range: fewer than two line directions -> throw
            r = np.full(m.n, 5001.0)
            j = rng.integers(5, m.n - 5)
            r[j:j + (k // 4) % 3] = 300
        r = r + rng.normal(0, 1.0, m.n) * (r < 5000)
        syn[k] = np.clip(np.round(r), 0, 5001).astype(np.uint16)
    sflags = np.array([m.is_corridor(*m.preprocess(s.astype(np.float32))) for s in syn], np.int8)
    print("synthetic", {int(v): int((sflags == v).sum()) for v in np.unique(sflags)})
    out["synth_scans"], out["synth_flags"] = syn, sflags
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "corridor.npz"), **out)


if __name__ == "__main__":
    main()
