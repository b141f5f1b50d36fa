"""GPU probe of SURVEY 8d config C5: multi-start PL-ICP on synthetic wide-FOV scans (n=2048, 360 deg, 30 m), G=16
initial guesses per pair around a noisy odometry, best-of-G reduced on the device (zhsm_best_of_starts, lowest mean
error).  Reports matches/s, the success rate against the generator's ground truth, and parity with the CPU
restatement on a sample of jobs."""
import os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
from laser_slam_b200 import engine as E
from oracle import pyoracle as po

P = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
G, n = 16, 2048
PM_PI = 3.141592654
lz = zh.Laser()
lz.n = n; lz.fi_min = np.float32(-PM_PI); lz.dfi = np.float32((360 * PM_PI / 180) / (n - 1.0))
lz.min_range_cm = 10; lz.max_range_cm = 3000; lz.min_valid_points = n // 5; lz.scan_window = n // 10
eng = zh.Engine(0); eng.set_laser(lz)
dev = torch.device("cuda:0")
scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev)
truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
eng.synth_pairs(scans, truth, 0, P, dmax_xy_m=0.5, dmax_th_rad=np.deg2rad(20.0), seed=20261105)
eng.median_filter(scans); eng.sync()
tr = truth.cpu().numpy()
rng = np.random.default_rng(20261105)
odo = tr + rng.normal(0, 1, (P, 3)) * [0.1, 0.1, np.deg2rad(3.0)]
grid = np.array([[dx, dy, np.deg2rad(dt)] for dx in (-0.3, 0.3) for dy in (-0.3, 0.3) for dt in (-10, -3.3, 3.3, 10)])
guess = (odo[:, None, :] + grid[None]).reshape(-1, 3)
ri = np.repeat(np.arange(P, dtype=np.int32) * 2, G); si = ri + 1
d_ri, d_si, d_g = (torch.from_numpy(a).to(dev) for a in (ri, si, guess))
res = torch.zeros((P * G, 128), dtype=torch.uint8, device=dev)
MS = os.environ.get("ZHSM_C5_MULTISTART", "0") != "0"   # four starts of a pair per CTA (ZHSM_MULTISTART); 0: one job per CTA
for _ in range(2):
    t = time.time(); eng.plicp(scans, d_ri, d_si, first_guess=d_g, results=res, multistart=MS); eng.sync(); dt = time.time() - t
kms = eng.last_kernel_ms()
t = time.time(); best, rec = eng.best_of_starts(res, P, G, criterion=E.BEST_MEAN_ERROR); eng.sync(); dtb = time.time() - t
r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(P, G)
b = rec.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
err = np.abs(b["x"] - tr); err[:, 2] = np.abs((err[:, 2] + np.pi) % (2 * np.pi) - np.pi)
ok = (b["valid"] == 1) & (err[:, 0] < 0.02) & (err[:, 1] < 0.02) & (err[:, 2] < np.deg2rad(0.5))
ok_any = ((r["valid"] == 1) & (np.abs(r["x"] - tr[:, None, :]) < [0.02, 0.02, np.deg2rad(0.5)]).all(2))
print("multistart instance" if MS else "single-job instance")
print(f"C5: {P} pairs x {G} starts = {P*G} matches (n={n}) in {dt*1e3:.1f} ms (kernel {kms:.1f} ms) -> {P*G/dt:.0f} matches/s, "
      f"{P/dt:.0f} pairs/s; best_of_starts {dtb*1e3:.2f} ms")
print(f"    valid starts {r['valid'].mean():.3f}; iterations/start {r['iterations'].mean():.1f}; pairs with best start within 2 cm / 0.5 deg of truth: "
      f"{ok.mean():.4f} (any start: {ok_any.any(1).mean():.4f}; single start nearest odometry: {ok_any[:, 0].mean():.4f})")
w = r["reserved"].astype(np.float64)
print(f"    digest: hash-sum {int(r['corr_hash'].astype(np.uint64).sum())} iterations-sum {int(r['iterations'].sum())} x-sum {r['x'][r['valid'] == 1].sum():.15e}")
print(f"    walks per start: {w[..., 0].mean():.0f} executed for {w[..., 1].mean():.0f} sequential ({w[..., 0].sum() / w[..., 1].sum():.2f}x); "
      f"re-search warp rounds {w[..., 2].mean():.0f}; kernel {kms / (P * G) * 148 * 1e3:.0f} us of SM time per start")
if os.environ.get("ZHSM_LIB_VARIANT") == "timing":  # per-phase cycle counters instead of the result (tools/gpu_probe_phases.py)
    q = r.reshape(-1)
    ph = np.concatenate([q["x"], q["cov"], q["error"][:, None], q["goodness"][:, None].astype(np.float64), q["quality"][:, None].astype(np.float64),
                         q["nvalid"][:, None].astype(np.float64), q["reserved"][:, 3:4].astype(np.float64),
                         q["corr_hash"][:, None].astype(np.float64), q["n_loops"][:, None].astype(np.float64),
                         q["reserved"][:, 0:3].astype(np.float64)], axis=1)
    names = ["load", "jump tables", "first-pass walk", "verify + re-search", "kill doubles", "select: bucket search", "kill + GPC sums", "corr hash",
             "solve: pose_diff, cos/sin, barrier", "loop control", "final corr/hash/cov", "solve: cross-warp sums", "solve: quartic coefficients",
             "solve: greatest root", "solve: 4x4 system", "solve: atan2", "trim: keys", "(unused)", "select: histogram passes + barrier"]
    tot = ph.sum(1)
    print(f"    cycles/job {tot.mean():.0f}")
    for k, nm in enumerate(names):
        print(f"      {nm:36s} {ph[:, k].mean():10.0f} cycles/job  {100 * ph[:, k].sum() / tot.sum():5.1f} %")
    sys.exit(0)
if os.environ.get("ZHSM_LIB_VARIANT") == "stats":  # walk-length counters instead of cov[] (tools/gpu_probe_walkstats.py)
    c = r["cov"].reshape(-1, 6); wa = w[..., 0].reshape(-1)
    print(f"    probes/walk {c[:,0].sum()/wa.sum():.2f}; walks > 32 probes: {100*c[:,2].sum()/wa.sum():.2f} % of the walks, {100*c[:,1].sum()/c[:,0].sum():.1f} % of the probes; "
          f"one-direction loop {100*c[:,3].sum()/c[:,0].sum():.1f} % of the probes; longest walk per start: mean {c[:,5].mean():.0f}, max {c[:,5].max():.0f}; "
          f"probes/start {c[:,0].mean():.0f}, in walks > 32: {c[:,1].mean():.0f}, walks > 32 per start {c[:,2].mean():.1f}")
    sys.exit(0)
# parity sample against the CPU restatement
K = 512
fm = scans[: 2 * (K // G)].cpu().numpy()
t = time.time()
o = po.zhcsm_fmatch_batch(fm, ri[:K], si[:K], lz.fi_min, lz.dfi, 10, 3000, first_guess=guess[:K], nthreads=os.cpu_count())
print(f"    oracle: {K} starts on {os.cpu_count()} threads in {time.time() - t:.2f} s -> {K / (time.time() - t):.0f} matches/s")
rr = r.reshape(-1)[:K]
dp = np.abs(rr["x"] - o["x"])
both = (rr["valid"] == 1) & (o["valid"] == 1)
print(f"    parity on {K} jobs: valid_eq {(rr['valid'] == o['valid']).mean():.4f} pose<=1e-4/1e-5 "
      f"{((dp[:, 0] <= 1e-4) & (dp[:, 1] <= 1e-4) & (dp[:, 2] <= 1e-5))[both].mean():.4f} iterations_eq "
      f"{(rr['iterations'] == o['iterations'])[both].mean():.4f} hash_eq {(rr['corr_hash'] == o['corr_hash'])[both].mean():.4f}")
