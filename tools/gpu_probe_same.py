"""Tuning probe: does the working library give the same records as a baseline build (libzhsm_base.so, e.g. HEAD built
in a git worktree)?  Runs the C4 sample with each library in a subprocess and compares the records byte for byte
(diagnostic fields excluded).  Run: python tools/gpu_probe_same.py [pairs] [beams]"""
import os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1080
if len(sys.argv) > 3:  # child
    sys.path.insert(0, ROOT)
    import torch
    import laser_slam_b200 as zh
    eng = zh.Engine(0)
    lz = zh.Laser(); lz.n = n
    fov = 270.0 if n <= 1080 else 360.0
    lz.fi_min = np.float32(-fov / 2 * np.pi / 180); lz.dfi = np.float32(fov * np.pi / 180 / (n - 1))
    lz.min_range_cm = 10; lz.max_range_cm = 3000; lz.min_valid_points = n // 5; lz.scan_window = n // 10
    eng.set_laser(lz)
    dev = torch.device("cuda:0")
    scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev); truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
    eng.synth_pairs(scans, truth, 0, P); eng.median_filter(scans)
    ri = torch.arange(0, 2 * P, 2, dtype=torch.int32, device=dev); si = ri + 1
    res = torch.zeros((P, 128), dtype=torch.uint8, device=dev)
    ms = []
    for _ in range(4):
        eng.plicp(scans, ri, si, results=res); eng.sync(); ms.append(eng.last_kernel_ms())
    np.save(sys.argv[3], res.cpu().numpy())
    print(f"{os.environ.get('ZHSM_LIB_VARIANT', 'work')}: kernel {min(ms[1:]):.2f} ms / {P} pairs -> {P / min(ms[1:]) * 1e3:.0f} pairs/s")
    sys.exit(0)
out = []
for var in ("base", ""):
    env = dict(os.environ)
    if var: env["ZHSM_LIB_VARIANT"] = var
    f = f"/tmp/same_{var or 'work'}.npy"
    subprocess.run([sys.executable, __file__, str(P), str(n), f], env=env, check=True)
    out.append(np.load(f))
sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
a, b = (o.view(zh.RESULT_DTYPE).reshape(-1) for o in out)
for fld in ("valid", "status", "x", "iterations", "n_loops", "nvalid", "error", "corr_hash", "cov"):
    same = np.array_equal(a[fld], b[fld], equal_nan=True) if a[fld].dtype.kind == "f" else np.array_equal(a[fld], b[fld])
    print(f"  {fld:10s} identical: {same}" + ("" if same else f"  ({(a[fld] != b[fld]).sum()} elements differ)"))
