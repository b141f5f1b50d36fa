"""Tuning probe: where one PL-ICP job spends its cycles.  Needs the timing variant of the library
(python -c "from laser_slam_b200 import build as b; b.build(variant='timing', defs=('PLICP_PHASE_TIMING',))"),
in which thread 0 of every CTA accumulates clock64() deltas per phase and the record's x/cov/error/goodness
carry those counters instead of the result.  Run: ZHSM_LIB_VARIANT=timing python tools/gpu_probe_phases.py [pairs]"""
import os, sys
os.environ.setdefault("ZHSM_LIB_VARIANT", "timing")
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
elif os.environ.get("ZHSM_PROBE_LOG"):               # a CARMEN log x 16 (consecutive-scan odometry, first guess 0)
    from tests.support import load_log, make_laser
    g = load_log(os.environ["ZHSM_PROBE_LOG"])
    eng = zh.Engine(0); eng.set_laser(make_laser(g))
    dev = torch.device("cuda:0")
    raw = np.ascontiguousarray(g["scans"]); S = raw.shape[0]
    scans = torch.from_numpy(raw).to(dev).repeat(16, 1); eng.median_filter(scans)
    ri = torch.cat([torch.arange(0, S - 1, dtype=torch.int32, device=dev) + k * S for k in range(16)]); si = ri + 1
    guess = None
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
for _ in range(2):
    eng.plicp(scans, ri, si, first_guess=guess, results=res); eng.sync()
r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
ph = np.concatenate([r["x"], r["cov"], r["error"][:, None], r["goodness"][:, None].astype(np.float64), r["quality"][:, None].astype(np.float64),
                     r["nvalid"][:, None].astype(np.float64), r["reserved"][:, 3:4].astype(np.float64),
                     r["corr_hash"][:, None].astype(np.float64), r["n_loops"][:, None].astype(np.float64),
                     r["reserved"][:, 0:3].astype(np.float64)], axis=1)
names = ["load", "jump tables", "first-pass walk", "verify + re-search", "kill doubles", "select: bucket search (all passes)",
         "kill + GPC sums", "corr hash", "solve: pose_diff, cos/sin, barrier", "loop control", "final corr/hash/cov",
         "solve: cross-warp sums", "solve: quartic coefficients", "solve: greatest root", "solve: 4x4 system", "solve: atan2",
         "trim: keys (dist_to_segment)", "(unused)", "select: histogram passes + barrier"]
tot = ph.sum(1)
print(f"kernel {eng.last_kernel_ms():.2f} ms for {P} pairs; iterations/pair {r['iterations'].mean():.2f}; "
      f"cycles/job {tot.mean():.0f} (thread 0 clock64, includes barrier waits and time-slicing with co-resident CTAs)")
for k, nm in enumerate(names):
    print(f"  {nm:22s} {ph[:, k].mean():10.0f} cycles/job  {100 * ph[:, k].sum() / tot.sum():5.1f} %   per iteration {ph[:, k].sum() / r['iterations'].sum():8.0f}")
w_cyc, w_pack = ph[:, 1], ph[:, 17]
calls = np.mod(w_pack, 256.0); msteps = np.floor(w_pack / 256.0)
print(f"walk routine (thread 0's warp): {calls.mean():.0f} calls/job, {w_cyc.sum() / calls.sum():.0f} cycles per call, "
      f"longest lane {msteps.sum() / calls.sum():.1f} steps per call -> {w_cyc.sum() / msteps.sum():.0f} cycles per step of the longest lane; "
      f"inside the routine: {100 * w_cyc.sum() / (ph[:, 2].sum() + ph[:, 3].sum()):.0f} % of the two search phases")
