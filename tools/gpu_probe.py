"""Scratch GPU probe: synthetic 1080-beam pairs -> PL-ICP timing + oracle parity on a sample."""
import os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
from oracle import pyoracle as po

P = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1080
ncheck = int(sys.argv[3]) if len(sys.argv) > 3 else 500
eng = zh.Engine(0)
lz = zh.Laser(); lz.n = n
fov = 270.0 if n == 1080 else 360.0
lz.fi_min = np.float32(-fov / 2 * np.pi / 180); lz.dfi = np.float32(fov * np.pi / 180 / (n - 1))
lz.min_range_cm = 10; lz.max_range_cm = 3000; lz.min_valid_points = n // 5; lz.scan_window = n // 10
eng.set_laser(lz)
print("peaks TFLOP/s: fp32", eng.fp_peak_tflops(32), "fp64", eng.fp_peak_tflops(64))
dev = torch.device("cuda:0")
scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev)
truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
eng.synth_pairs(scans, truth, 0, P); eng.sync()
raw = scans.clone()
eng.median_filter(scans); eng.sync()
ri = torch.arange(0, 2 * P, 2, dtype=torch.int32, device=dev); si = ri + 1
res = torch.zeros((P, 128), dtype=torch.uint8, device=dev)
for rep in range(3):
    t = time.time(); eng.plicp(scans, ri, si, results=res); eng.sync(); dt = time.time() - t
    print(f"rep {rep}: {P} pairs in {dt*1e3:.1f} ms (kernel {eng.last_kernel_ms():.1f} ms) -> {P/dt:.0f} pairs/s")
r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
tr = truth.cpu().numpy()
print("valid", r["valid"].mean(), "iters mean", r["iterations"].mean(), "loops hist", np.bincount(r["n_loops"]))
err = np.abs(r["x"] - tr); v = r["valid"] == 1
print("vs truth: median |dx| %.4f |dy| %.4f |dth| %.5f ; 95%% %.4f %.4f %.5f" % (*np.median(err[v], 0), *np.percentile(err[v], 95, 0)))
# oracle on a sample
h = scans[: 2 * ncheck].cpu().numpy()
t = time.time()
o = po.zhcsm_fmatch_batch(h, np.arange(0, 2 * ncheck, 2), np.arange(1, 2 * ncheck, 2), lz.fi_min, lz.dfi, 10, 3000, nthreads=os.cpu_count())
dt = time.time() - t
print(f"oracle: {ncheck} pairs in {dt:.2f}s on {os.cpu_count()} threads -> {ncheck/dt:.0f} pairs/s")
g = r[:ncheck]
both = (g["valid"] == 1) & (o["valid"] == 1)
d = np.abs(g["x"] - o["x"])
ok = (d[:, 0] < 1e-4) & (d[:, 1] < 1e-4) & (d[:, 2] < 1e-5)
print("parity: valid_eq", (g["valid"] == o["valid"]).mean(), "pose_ok", ok[both].mean(), "iters_eq", (g["iterations"] == o["iterations"]).mean(),
      "hash_eq", (g["corr_hash"] == o["corr_hash"])[both].mean(), "max d", d[both].max(0))
bad = np.where(~ok & both)[0][:10]
for b in bad: print(" mismatch", b, g["x"][b], o["x"][b], g["iterations"][b], o["iterations"][b])
