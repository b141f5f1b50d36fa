"""Tuning probe: distribution of walk lengths (probes of the ref scan per ray walk) in the PL-ICP kernel.  Needs
python -c "from laser_slam_b200 import build as b; b.build(variant='stats', defs=('PLICP_WALK_STATS',))"; the
record's cov[] then carries the counters.  Run: python tools/gpu_probe_walkstats.py [pairs]"""
import os, sys
os.environ.setdefault("ZHSM_LIB_VARIANT", "stats")
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
import bench
P = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
if os.environ.get("ZHSM_PROBE_C5", "0") != "0":   # the 2048-beam multi-start workload instead of the 1080-beam one
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import c5_workload
    eng, scans, ri, si, guess = c5_workload.make(zh, P // 16)
    P = ri.numel()
else:
    eng = zh.Engine(0); eng.set_laser(bench.make_laser(zh))
    dev = torch.device("cuda:0")
    n = bench.N_BEAMS
    scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev); truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
    eng.synth_pairs(scans, truth, 0, P, seed=bench.SEED); eng.median_filter(scans)
    ri = torch.arange(0, 2 * P, 2, dtype=torch.int32, device=dev); si = ri + 1
    guess = None
dev = torch.device("cuda:0")
res = torch.zeros((P, 128), dtype=torch.uint8, device=dev)
eng.plicp(scans, ri, si, first_guess=guess, results=res); eng.sync()
r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
c = r["cov"]; w = r["reserved"][:, 0].astype(np.float64)
print(f"walks/pair {w.mean():.0f}; probes/walk {c[:,0].sum()/w.sum():.2f}")
print(f"walks > 32 probes: {100*c[:,2].sum()/w.sum():.2f} % of the walks, {100*c[:,1].sum()/c[:,0].sum():.1f} % of the probes")
print(f"probes in the one-direction loop: {100*c[:,3].sum()/c[:,0].sum():.1f} % of all, {100*c[:,4].sum()/max(c[:,1].sum(),1):.1f} % of the probes of walks > 32")
print(f"longest walk per pair: mean {c[:,5].mean():.0f}, max {c[:,5].max():.0f}; pairs with a walk > 256: {100*(c[:,5]>256).mean():.1f} %")
