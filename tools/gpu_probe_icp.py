"""Scratch GPU probe: zhicp kernel throughput on BASELINE config 3 (node pairs x 64 initial guesses) vs the CPU oracle."""
import os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
from oracle import pyoracle as po

rep = int(sys.argv[1]) if len(sys.argv) > 1 else 16
d = np.load(os.path.join(ROOT, "tests", "golden", "zhicp_nodes.npz"))
P, G = d["pairs"].shape[0], 64
gx, gy, gt = np.meshgrid(np.linspace(-0.3, 0.3, 4), np.linspace(-0.3, 0.3, 4), np.linspace(-0.15, 0.15, 4), indexing="ij")
grid = np.stack([gx.ravel(), gy.ravel(), gt.ravel()], 1)
init = (d["init"][:, None, :] + grid[None]).reshape(-1, 3)
ri = np.repeat(d["pairs"][:, 0], G); ci = np.repeat(d["pairs"][:, 1], G)
eng = zh.Engine(0)
dev = torch.device("cuda:0")
pts = torch.from_numpy(d["pts"]).to(dev)
dri = torch.from_numpy(np.tile(ri, rep)).to(dev); dci = torch.from_numpy(np.tile(ci, rep)).to(dev)
dinit = torch.from_numpy(np.tile(init, (rep, 1))).to(dev)
B = dri.shape[0]
res = torch.zeros((B, 128), dtype=torch.uint8, device=dev)
for _ in range(3):
    t = time.time(); eng.icp(pts, d["off"], d["cnt"], dri, dci, dinit, results=res); eng.sync(); dt = time.time() - t
r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
print(f"icp: {B} jobs ({P} pairs x {G} starts x {rep}) in {dt*1e3:.1f} ms (kernel {eng.last_kernel_ms():.1f} ms) -> "
      f"{B/dt:.0f} jobs/s = {B/dt/G:.0f} 64-start pairs/s; iters mean {r['iterations'].mean():.1f} max {r['iterations'].max()}; "
      f"goodness>=0.8: {(r['goodness']>=0.8).mean():.3f}")
t = time.time(); best, rec = eng.best_of_starts(res, P * rep, G); eng.sync(); print(f"best_of_starts: {(time.time()-t)*1e3:.2f} ms")
nt = os.cpu_count()
t = time.time(); o = po.mrpt_icp_batch(d["pts"], d["off"], d["cnt"], ri, ci, init, nthreads=nt); dt = time.time() - t
print(f"oracle: {ri.size} jobs on {nt} threads in {dt:.2f}s -> {ri.size/dt:.0f} jobs/s")
k = slice(0, ri.size)
dp = np.abs(r["x"][k] - o["pose"])
print("pose agree (1e-4/1e-5):", ((dp[:, 0] <= 1e-4) & (dp[:, 1] <= 1e-4) & (dp[:, 2] <= 1e-5)).mean(),
      "iters equal:", (r["iterations"][k] == o["nIterations"]).mean(), "max dpose", dp.max(0))
