import os, sys
import numpy as np, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
import laser_slam_b200 as zh
import bench, c5_workload
P = 20000
eng = zh.Engine(0); eng.set_laser(bench.make_laser(zh))
dev = torch.device("cuda:0"); n = bench.N_BEAMS
scans = torch.empty((2 * P, n), dtype=torch.float32, device=dev); truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
eng.synth_pairs(scans, truth, 0, P, seed=bench.SEED); eng.median_filter(scans)
ri = torch.arange(0, 2 * P, 2, dtype=torch.int32, device=dev)
res = torch.zeros((P, 128), dtype=torch.uint8, device=dev)
eng.plicp(scans, ri, ri + 1, results=res); eng.sync()
r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
print("C4: restarting jobs %.3f %%, loops per job %.3f" % (100 * (r["n_loops"] > 1).mean(), r["n_loops"].mean()))
eng.close()
eng, scans, ri, si, guess = c5_workload.make(zh, 1024)
res = torch.zeros((ri.numel(), 128), dtype=torch.uint8, device=dev)
eng.plicp(scans, ri, si, first_guess=guess, results=res); eng.sync()
r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
print("C5: restarting jobs %.3f %%, loops per job %.3f" % (100 * (r["n_loops"] > 1).mean(), r["n_loops"].mean()))
