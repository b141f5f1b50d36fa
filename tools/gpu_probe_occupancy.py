"""Tuning probe: PL-ICP kernel time against resident CTAs per SM (ZHSM_PLICP_SMEM_PAD pads the dynamic shared memory so
that fewer CTAs fit).  Time ~ 1/CTAs means the kernel is bound by the latency of each job's dependent chains; a curve
that flattens means a shared resource (issue slots, a pipe) is saturated.  One process per setting (the knob is read in
zhsm_create).  Run: python tools/gpu_probe_occupancy.py [pairs] [beams]"""
import os, subprocess, sys
P = sys.argv[1] if len(sys.argv) > 1 else "20000"
n = sys.argv[2] if len(sys.argv) > 2 else "1080"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code = r"""
import os, sys, numpy as np, torch
sys.path.insert(0, %r)
import laser_slam_b200 as zh
P, n = int(sys.argv[1]), int(sys.argv[2])
eng = zh.Engine(0)
lz = zh.Laser(); lz.n = n
fov = 270.0 if n <= 1440 else 360.0
lz.fi_min = np.float32(-fov / 2 * np.pi / 180); lz.dfi = np.float32(fov * np.pi / 180 / (n - 1))
lz.min_range_cm = 10; lz.max_range_cm = 3000; lz.min_valid_points = n // 5; lz.scan_window = n // 10
eng.set_laser(lz)
dev = torch.device("cuda:0")
scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev); truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
eng.synth_pairs(scans, truth, 0, P); eng.median_filter(scans)
ri = torch.arange(0, 2 * P, 2, dtype=torch.int32, device=dev)
res = torch.zeros((P, 128), dtype=torch.uint8, device=dev)
ms = []
for _ in range(4):
    eng.plicp(scans, ri, ri + 1, results=res); eng.sync(); ms.append(eng.last_kernel_ms())
r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
print("%%.2f ms (min of 3 after warm-up), hash-sum %%d" %% (min(ms[1:]), int(r["corr_hash"].astype(np.uint64).sum())))
""" % ROOT
for pad_kb in [int(v) for v in os.environ.get("PADS_KB", "0,20,58,150").split(",")]:
    env = dict(os.environ, ZHSM_PLICP_SMEM_PAD=str(pad_kb * 1024 + 1))   # odd: the launch reports its CTAs per SM
    out = subprocess.run([sys.executable, "-c", code, P, n], env=env, capture_output=True, text=True)
    occ = [ln for ln in out.stderr.splitlines() if "CTAs per SM" in ln][:1]
    print(f"pad {pad_kb:3d} KB: {out.stdout.strip() or out.stderr.strip()[-300:]} {occ[0] if occ else ''}")
