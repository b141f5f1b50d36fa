"""Tuning probe: PL-ICP consecutive-scan odometry over the CARMEN logs (intel-lab 181 beams, fr079 / mit-cscail 361),
the log x 16 in one batch and the log once; kernel time per call.  Real logs restart (sm_icp's six perturbed loops) on
~10 % of the pairs, which uniform synthetic batches do not.  Run: ZHSM_PLICP_T=<T> python tools/gpu_probe_logs.py"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
from tests.support import load_log, make_laser
dev = torch.device("cuda:0")
for log in ("intel-lab", "fr079", "mit-cscail"):
    g = load_log(log)
    eng = zh.Engine(0); eng.set_laser(make_laser(g))
    raw_h = np.ascontiguousarray(g["scans"]); S, n = raw_h.shape; B = S - 1
    out = []
    for rep in (16, 1):
        raw = torch.from_numpy(raw_h).to(dev).repeat(rep, 1)
        eng.median_filter(raw)
        ri = torch.cat([torch.arange(0, B, dtype=torch.int32, device=dev) + k * S for k in range(rep)]); si = ri + 1
        res = torch.zeros((B * rep, 128), dtype=torch.uint8, device=dev)
        ms = []
        for _ in range(5):
            eng.plicp(raw, ri, si, results=res); eng.sync(); ms.append(eng.last_kernel_ms())
        r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
        out.append(f"x{rep}: {min(ms[1:]):.2f} ms = {B * rep / min(ms[1:]) * 1e3 / 1e3:.0f} k pairs/s")
    print(f"{log} ({n} beams, {B} pairs, {r['iterations'].mean():.1f} iterations, {r['n_loops'].mean():.2f} loops per pair): " + "; ".join(out)
          + f"; hash-sum {int(r['corr_hash'].astype(np.uint64).sum())}")
    eng.close()
