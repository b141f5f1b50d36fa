#!/usr/bin/env python
"""Freeze tests/golden/zhicp_nodes.npz (run HERE, where /root/reference exists).

Inputs : the node point clouds and candidate pairs the reference ships for its ICP wrapper test
         (src/zhicp/data{,2}/<node>.log = "x y" per line in metres, matched_nodes.log = "ref cur num px py pth",
         read exactly like src/zhicp/testWarp.cpp:93-160 does; pairs whose clouds are missing/empty are dropped).
Outputs: clouds (float32, the values fscanf("%f") yields), pairs, initial guesses, and the results of
         oracle/icp_oracle.c with the wrapper's options.  PARITY UNPINNED: MRPT 0.9.6 is absent, so these outputs
         pin the restatement against regressions and are what the CUDA path is compared with; they are not outputs
         of the reference.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402

REF = "/root/reference/src/zhicp"
OUT = os.path.join(ROOT, "tests", "golden", "zhicp_nodes.npz")


def read_cloud(path):
    if not os.path.exists(path):
        return None
    a = np.loadtxt(path, dtype=np.float64, ndmin=2)
    if a.size == 0:
        return None
    return a[:, 0].astype(np.float32), a[:, 1].astype(np.float32)


def main():
    po.build()
    clouds, names, pairs, init = [], {}, [], []
    for d in ("data", "data2"):
        for line in open(os.path.join(REF, d, "matched_nodes.log")):
            f = line.split()
            if len(f) != 6:
                continue
            ids = []
            for node in f[:2]:
                key = (d, int(node))
                if key not in names:
                    c = read_cloud(os.path.join(REF, d, node + ".log"))
                    names[key] = -1 if c is None else len(clouds)
                    if c is not None:
                        clouds.append(c)
                ids.append(names[key])
            if min(ids) < 0:
                continue
            pairs.append(ids)
            init.append([float(v) for v in f[3:6]])
    pairs = np.array(pairs, np.int32)
    init = np.array(init, np.float64)
    pts, off, cnt = po.pack_clouds(clouds)
    res = po.mrpt_icp_batch(pts, off, cnt, pairs[:, 0], pairs[:, 1], init, nthreads=8)
    zero = po.mrpt_icp_batch(pts, off, cnt, pairs[:, 0], pairs[:, 1], np.zeros_like(init), nthreads=8)
    tag = np.array([f"{d}/{n}" for (d, n), k in sorted(names.items(), key=lambda kv: kv[1]) if k >= 0])
    np.savez_compressed(OUT, pts=pts, off=off, cnt=cnt, pairs=pairs, init=init, names=tag,
                        oracle=res, oracle_zero_init=zero)
    print(OUT, os.path.getsize(OUT), "bytes;", len(clouds), "clouds", len(pairs), "pairs")
    print("iters", res["nIterations"], "\ngoodness", np.round(res["goodness"], 3))
    print("pose-init", np.round(res["pose"] - init, 3)[:8])


if __name__ == "__main__":
    main()
