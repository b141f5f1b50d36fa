"""Scratch GPU probe: zhpsm kernels' throughput on a replicated CARMEN log vs the compiled reference on the host."""
import os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
from laser_slam_b200 import engine as E
from oracle import pyoracle as po
from tests.support import load_log, make_laser

log = sys.argv[1] if len(sys.argv) > 1 else "fr079"
rep = int(sys.argv[2]) if len(sys.argv) > 2 else 32
g = load_log(log)
eng = zh.Engine(0)
eng.set_laser(make_laser(g))
dev = torch.device("cuda:0")
scans = torch.from_numpy(g["scans"].copy()).to(dev)
S, n = scans.shape
t = time.time(); _, bad, seg = eng.preprocess(scans); eng.sync(); print(f"preprocess {S} scans: {(time.time()-t)*1e3:.2f} ms")
ri = torch.arange(0, S - 1, dtype=torch.int32, device=dev).repeat(rep)
ai = ri + 1
B = ri.shape[0]
res = torch.zeros((B, 128), dtype=torch.uint8, device=dev)
for name, algo in (("pm_psm", E.ALGO_PSM), ("pm_icp", E.ALGO_PM_ICP), ("fmatch", E.ALGO_FMATCH)):
    for r_ in range(2):
        t = time.time(); eng.psm(scans, bad, seg, ri, ai, algo=algo, results=res); eng.sync(); dt = time.time() - t
    r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
    print(f"{name}: {B} jobs (n={n}) in {dt*1e3:.1f} ms (kernel {eng.last_kernel_ms():.1f} ms) -> {B/dt:.0f} jobs/s; "
          f"valid {r['valid'].mean():.4f} iters {r['iterations'].mean():.2f}")
if po.have_ref():
    m = po.RefPolarMatch(g["laser_name"], fi_min=g["fi_min"], max_range_cm=int(g["max_range"]))
    R, Bd, G = scans.cpu().numpy(), bad.cpu().numpy().astype(np.int32), seg.cpu().numpy().astype(np.int32)
    nt = os.cpu_count()
    for name, algo in (("pm_psm", 0), ("pm_icp", 1)):
        t = time.time(); m.match_batch(algo, R, Bd, G, np.arange(S - 1), np.arange(1, S), nthreads=nt); dt = time.time() - t
        print(f"reference {name}: {S-1} pairs on {nt} threads in {dt:.2f}s -> {(S-1)/dt:.0f} pairs/s")
