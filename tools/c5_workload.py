"""The C5 workload of tools/gpu_probe_c5.py (2048 beams, 360 deg, 16 first guesses per pair) for the other tuning
probes: returns (engine, scans, ref_idx, sens_idx, first_guess) with everything on the device."""
import numpy as np
import torch


def make(zh, P, G=16, n=2048, seed=20261105):
    PM_PI = 3.141592654
    lz = zh.Laser()
    lz.n = n; lz.fi_min = np.float32(-PM_PI); lz.dfi = np.float32((360 * PM_PI / 180) / (n - 1.0))
    lz.min_range_cm = 10; lz.max_range_cm = 3000; lz.min_valid_points = n // 5; lz.scan_window = n // 10
    eng = zh.Engine(0); eng.set_laser(lz)
    dev = torch.device("cuda:0")
    scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev)
    truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
    eng.synth_pairs(scans, truth, 0, P, dmax_xy_m=0.5, dmax_th_rad=np.deg2rad(20.0), seed=seed)
    eng.median_filter(scans); eng.sync()
    tr = truth.cpu().numpy()
    rng = np.random.default_rng(seed)
    odo = tr + rng.normal(0, 1, (P, 3)) * [0.1, 0.1, np.deg2rad(3.0)]
    grid = np.array([[dx, dy, np.deg2rad(dt)] for dx in (-0.3, 0.3) for dy in (-0.3, 0.3) for dt in (-10, -3.3, 3.3, 10)])
    guess = (odo[:, None, :] + grid[None]).reshape(-1, 3)
    ri = np.repeat(np.arange(P, dtype=np.int32) * 2, G); si = ri + 1
    d_ri, d_si, d_g = (torch.from_numpy(a).to(dev) for a in (ri, si, guess))
    return eng, scans, d_ri, d_si, d_g
