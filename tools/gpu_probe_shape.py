"""Tuning probe: PL-ICP kernel time per CTA shape (ZHSM_PLICP_T) and beam count; results must not depend on T.
Run: python tools/gpu_probe_shape.py [pairs] [n:T,T,... ...]   e.g.  20000 768:192,256 1080:192,256"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
P = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
specs = sys.argv[2:] or ["768:192,256", "1080:192,256"]
dev = torch.device("cuda:0")
for sp in specs:
    n, ts = sp.split(":"); n = int(n)
    base = None
    for T in [int(x) for x in ts.split(",")]:
        os.environ["ZHSM_PLICP_T"] = str(T)
        eng = zh.Engine(0)
        lz = zh.Laser(); lz.n = n
        lz.fi_min = np.float32(-135 * np.pi / 180); lz.dfi = np.float32(270 * np.pi / 180 / (n - 1))
        lz.min_range_cm = 10; lz.max_range_cm = 3000; lz.min_valid_points = n // 5; lz.scan_window = n // 10
        eng.set_laser(lz)
        scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev); truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
        eng.synth_pairs(scans, truth, 0, P); eng.median_filter(scans)
        ri = torch.arange(0, 2 * P, 2, dtype=torch.int32, device=dev); si = ri + 1
        res = torch.zeros((P, 128), dtype=torch.uint8, device=dev)
        ms = []
        for _ in range(4):
            eng.plicp(scans, ri, si, results=res); eng.sync(); ms.append(eng.last_kernel_ms())
        r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
        key = (r["valid"], r["x"].tobytes(), r["iterations"], r["corr_hash"])
        same = "" if base is None else "  identical to first T: %s" % (all(np.array_equal(a, b) if not isinstance(a, bytes) else a == b for a, b in zip(key, base)))
        base = base or key
        print(f"n={n} T={T}: kernel {min(ms[1:]):.2f} ms / {P} pairs -> {P / min(ms[1:]) * 1e3:.0f} pairs/s; iterations/pair {r['iterations'].mean():.2f}{same}", flush=True)
        del eng
