"""Scratch probe: how many ray walks the speculate/verify search executes per sequential walk (PL-ICP kernel)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
import bench
P = 20000
eng = zh.Engine(0); eng.set_laser(bench.make_laser(zh))
dev = torch.device("cuda:0")
n = bench.N_BEAMS
scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev); truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
eng.synth_pairs(scans, truth, 0, P, seed=bench.SEED); eng.median_filter(scans)
ri = torch.arange(0, 2 * P, 2, dtype=torch.int32, device=dev); si = ri + 1
res = torch.zeros((P, 128), dtype=torch.uint8, device=dev)
for _ in range(2):
    eng.plicp(scans, ri, si, results=res); eng.sync()
r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
w = r["reserved"].astype(np.float64)
print(f"kernel {eng.last_kernel_ms():.2f} ms for {P} pairs; iterations/pair {r['iterations'].mean():.2f}")
print(f"walks/pair all {w[:,0].mean():.0f} first-pass {w[:,1].mean():.0f} ratio {w[:,0].sum()/w[:,1].sum():.3f}; "
      f"verify rounds with work per iteration {w[:,2].sum()/r['iterations'].sum():.2f}")
