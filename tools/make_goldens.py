#!/usr/bin/env python
"""Freeze golden fixtures under tests/golden/ (run HERE, where /root/reference exists).

Inputs : the three CARMEN logs under /root/reference/data (read once, stored losslessly as uint16 cm).
Outputs: per log one compressed .npz holding
  - raw scans (uint16 cm, exactly what readCarmon produces before the float cast, ZHPolar_Match.cpp:252),
    laser poses, the header fields readCarmon copies into Base_PARAM (:227-235)
  - zhpsm goldens produced by EXECUTING the reference (oracle/_ref/libzhpsm_ref.so):
      crc32 of the preprocessed r/bad/seg of every scan, pm_psm and pm_icp pose+status for every
      consecutive pair (zero initial pose, like testAccuracy :403-428), pm_psm->pm_icp chain,
      full pm_scan_project outputs for a few scans
  - PL-ICP goldens produced by the CPU restatement oracle/plicp_oracle.c ("parity unpinned" vs CSM;
    these pin the restatement against regressions and are what the CUDA path is compared with).
"""
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from carmen import LASERS, PM_D2R, read_carmen  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

LOGS = [("intel-lab", "LMS211"), ("fr079", "LMS511"), ("mit-cscail", "LMS511")]
REF_DATA = "/root/reference/data"
OUT = os.path.join(ROOT, "tests", "golden")


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def main():
    po.build()
    os.makedirs(OUT, exist_ok=True)
    for log, laser in LOGS:
        d = read_carmen(os.path.join(REF_DATA, log + ".log"))
        r = d["r_cm"]
        assert np.all(r == np.round(r)) and r.max() < 65535
        S, num = r.shape
        m = po.RefPolarMatch(laser, fi_min=np.float32(d["start"]), max_range_cm=int(d["max_range_m"] * 100))
        n = m.n
        assert num in (n, n - 1)
        # readCarmon :250-279: short readings and the padded last beam become max_range+1
        scans = np.full((S, n), m.max_range + 1, np.float32)
        scans[:, :num] = r
        scans[scans < 10.0] = m.max_range + 1
        pre = [m.preprocess(s) for s in scans]
        R = np.stack([p[0] for p in pre])
        B = np.stack([p[1] for p in pre])
        G = np.stack([p[2] for p in pre])
        pre_crc = np.array([[crc(R[k]), crc(B[k]), crc(G[k])] for k in range(S)], np.uint32)
        ri, ai = np.arange(S - 1), np.arange(1, S)
        g = {}
        for name, algo in (("psm", 0), ("icp", 1), ("psm_icp", 2)):
            poses, st = m.match_batch(algo, R, B, G, ri, ai, nthreads=8)
            g[name + "_pose"], g[name + "_status"] = poses, st.astype(np.int8)
            print(log, name, "failures", int(st.sum()), "of", S - 1)
        # a few full scan projections (pose = the pm_icp result of that pair)
        proj_idx = np.linspace(1, S - 1, 12).astype(np.int32)
        proj_r, proj_bad = [], []
        for k in proj_idx:
            nr, nb = m.scan_project(R[k], B[k], G[k], g["icp_pose"][k - 1])
            proj_r.append(nr)
            proj_bad.append(nb)
        # error index of the pm_icp result
        ei = np.zeros((S - 1, 2), np.float32)
        for k in range(1, S):
            if g["icp_status"][k - 1] == 0:
                ei[k - 1], _ = m.error_index2((R[k - 1], B[k - 1], G[k - 1]), (R[k], B[k], G[k]), g["icp_pose"][k - 1])
        # PL-ICP restatement (zhcsm FMatch flow: median filter only, then sm_icp)
        filt = np.stack([po.median_filter(s) for s in scans])
        assert np.array_equal(filt, R)  # the two median filters are twins (ZHCanonical_Matcher.cpp:159 / ZHPolar_Match.cpp:1610)
        pl = po.zhcsm_fmatch_batch(filt, ri, ai, m.fi_min, m.dfi, m.min_range, m.max_range, nthreads=8)
        print(log, "plicp valid", int(pl["valid"].sum()), "iters", float(pl["iterations"].mean()))
        np.savez_compressed(
            os.path.join(OUT, log + ".npz"),
            laser=np.array(laser), ranges_cm=r.astype(np.uint16), laser_pose=d["pose"],
            start=np.float64(d["start"]), max_range_m=np.float64(d["max_range_m"]),
            n=np.int32(n), fi_min=np.float32(m.fi_min), dfi=np.float32(m.dfi), min_range=np.int32(m.min_range),
            max_range=np.int32(m.max_range), min_valid=np.int32(m.min_valid), window=np.int32(m.window),
            pre_crc=pre_crc, proj_idx=proj_idx, proj_r=np.stack(proj_r), proj_bad=np.stack(proj_bad).astype(np.int8),
            err_index=ei,
            plicp_valid=pl["valid"].astype(np.int8), plicp_x=pl["x"], plicp_iterations=pl["iterations"],
            plicp_nvalid=pl["nvalid"], plicp_error=pl["error"], plicp_hash=pl["corr_hash"],
            plicp_loops=pl["n_loops"].astype(np.int8), **g)
        print("wrote", log, os.path.getsize(os.path.join(OUT, log + ".npz")) // 1024, "KiB")


if __name__ == "__main__":
    main()
