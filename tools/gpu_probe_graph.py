"""GPU probe of the kernels either side of the ICP batches: submap clouds of a CARMEN log's nodes
(points / raycast / rowcount / rowscan / emit kernels), candidate gating and consensus -- the launches ncu captures for
profiles/.  Run: python tools/gpu_probe_graph.py [log]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
from tests.loop_cases import build_c3
from tests.support import load_log, make_laser

log = sys.argv[1] if len(sys.argv) > 1 else "mit-cscail"
g = load_log(log)
eng = zh.Engine(0)
eng.set_laser(make_laser(g))
t = time.time()
nodes, clouds, pose, flags, pairs, init = build_c3(g, lambda nd: eng.submap_cloud(nd["scans"], nd["traj"])[:2], eng.gate_candidates)
dt = time.time() - t
print(f"{log}: {len(nodes)} nodes ({np.mean([c[0].size for c in clouds]):.0f} cloud points each), {len(pairs)} pairs: "
      f"clouds + gating + guesses in {dt*1e3:.0f} ms")
t = time.time()
for nd in nodes:
    eng.submap_cloud(nd["scans"], nd["traj"])
print(f"submap_cloud: {(time.time()-t)/len(nodes)*1e3:.2f} ms per node (15 scans x {int(g['n'])} beams, host arrays in and out)")
rng = np.random.default_rng(1)
S = 2000
set_off = np.arange(0, 6 * S + 1, 6, dtype=np.int32)
poses = rng.normal(0, 0.05, (6 * S, 3)) + np.repeat(rng.uniform(-20, 20, (S, 3)), 6, 0)
good = rng.uniform(0.8, 1.0, 6 * S).astype(np.float32)
t = time.time(); eng.consensus(poses, good, set_off); print(f"consensus: {S} sets of 6 in {(time.time()-t)*1e3:.2f} ms")
N = len(nodes)
ld = np.stack([[c[0].min(), c[1].min()] for c in clouds]).astype(np.float64); ru = np.stack([[c[0].max(), c[1].max()] for c in clouds]).astype(np.float64)
t = time.time(); eng.gate_candidates(pose, ld, ru, np.arange(N)); print(f"gate_candidates: {N}x{N} in {(time.time()-t)*1e3:.2f} ms")
